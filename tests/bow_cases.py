"""Synthetic DBoW2 vocabularies and feature sets shared by the golden generator, the CPU oracle tests and the
GPU parity tests (TEST INFRASTRUCTURE).  ORBvoc.txt is not shipped with the reference tree,
so the trees are random: same branching (k = 10), node order of saveToTextFile (a node's k children are
consecutive, created before the recursion descends), idf-like positive weights, a few stop words (weight 0)
and a few early leaves (clusters of one descriptor, HKmeansStep :560-575)."""
import numpy as np

import orbfe

# name: (seed, k, L, weighting, norm, early-leaf %, stop-word %, features, levelsup)
BOW_VOC_CASES = {
    "voc_k10_L4": (3100, 10, 4, 0, 1, 3, 2, 1500, 2),          # TF_IDF + L1 (what ORBvoc uses), FeatureVector at level 2
    "voc_k10_L3_tf_none": (3101, 10, 3, 1, 0, 0, 0, 900, 1),   # TF without normalisation: divided by the word count
    "voc_k6_L4_idf_l2": (3102, 6, 4, 2, 2, 5, 5, 1200, 4),     # IDF (addIfNotExist) + L2; levelsup = L: node 0
    "voc_k10_L3_binary": (3103, 10, 3, 3, 1, 0, 3, 700, 5),    # BINARY; levelsup > L: node 0
}


def make_tree(seed, k, L, early_pct, stop_pct):
    """Returns (parent, is_leaf, desc, weight) in node order; node 0 = root."""
    rng = np.random.default_rng(seed)
    parent, leaf, weight, level = [0], [0], [0.0], [0]
    desc = [np.zeros(32, np.uint8)]

    def grow(pid, lev):
        first = len(parent)
        for _ in range(k):
            parent.append(pid); leaf.append(0); weight.append(0.0); level.append(lev)
            desc.append(rng.integers(0, 256, 32, dtype=np.uint8))
        for c in range(first, first + k):
            if lev < L and rng.integers(0, 100) >= early_pct:
                grow(c, lev + 1)
            else:
                leaf[c] = 1
                weight[c] = 0.0 if rng.integers(0, 100) < stop_pct else float(rng.uniform(0.5, 9.0))

    grow(0, 1)
    return (np.array(parent, np.int32), np.array(leaf, np.uint8), np.stack(desc), np.array(weight, np.float64))


def make_full_tree(seed, k, L, stop_pct=1):
    return orbfe.synth.vocabulary(seed, k, L, stop_pct)


def features(seed, tree_desc, n):
    """Descriptors near random tree nodes (so descents are decisive) mixed with pure noise."""
    rng = np.random.default_rng(seed + 77)
    pick = rng.integers(1, len(tree_desc), n)
    d = tree_desc[pick].copy()
    for i in range(n):
        nb = int(rng.integers(0, 90))
        for b in rng.choice(256, nb, replace=False):
            d[i, b >> 3] ^= 1 << (b & 7)
    noise = rng.random(n) < 0.15
    d[noise] = rng.integers(0, 256, (int(noise.sum()), 32), dtype=np.uint8)
    d[n // 2] = d[3]                       # repeated descriptor: same word twice
    return d


def voc_inputs(name):
    seed, k, L, weighting, norm, early, stop, n, levelsup = BOW_VOC_CASES[name]
    parent, leaf, desc, weight = make_tree(seed, k, L, early, stop)
    return dict(k=k, L=L, weighting=weighting, norm=norm, parent=parent, leaf=leaf, desc=desc, weight=weight,
                feat=features(seed, desc, n), levelsup=levelsup)
