"""Python restatement of the DBoW2 vocabulary tree used by ORB-SLAM3 (TEST INFRASTRUCTURE).

Follows Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h (loadFromTextFile :1350-1428, saveToTextFile :1433-1454,
transform :1137-1206 and :1230-1271), BowVector.cpp:30-98, FeatureVector.cpp:31-45, FORB.cpp:81-101.
Objects are kept as in the reference: a node list with `children` lists, std::map -> dict + sorted().
Written independently of oracle/cpu_ref/bow_ref.cc; the golden fixtures come from this file.
"""
import math

import numpy as np

TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3          # DBoW2::WeightingType
NORM_NONE, NORM_L1, NORM_L2 = 0, 1, 2         # LNorm that mustNormalize() hands back (L1 for ORBvoc's L1_NORM scoring)
_POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


class Node:
    __slots__ = ("id", "parent", "children", "descriptor", "weight", "word_id")

    def __init__(self, nid):
        self.id, self.parent, self.children, self.descriptor, self.weight, self.word_id = nid, 0, [], None, 0.0, 0

    def is_leaf(self):
        return not self.children


class Vocabulary:
    def __init__(self, k, L, weighting=TF_IDF, norm=NORM_L1):
        self.k, self.L, self.weighting, self.norm = k, L, weighting, norm
        self.nodes = [Node(0)]
        self.words = []

    def add_node(self, pid, is_leaf, descriptor, weight):
        """One line of the text format (:1393-1424)."""
        nid = len(self.nodes)
        nd = Node(nid)
        nd.parent, nd.descriptor, nd.weight = pid, np.asarray(descriptor, np.uint8).copy(), float(weight)
        self.nodes.append(nd)
        self.nodes[pid].children.append(nid)
        if is_leaf:
            nd.word_id = len(self.words)
            self.words.append(nd)
        return nid

    @classmethod
    def load_text(cls, path, norm=NORM_L1):
        with open(path) as f:
            k, L, n1, n2 = (int(x) for x in f.readline().split())
            v = cls(k, L, n2, norm)
            for line in f:
                t = line.split()
                if not t:
                    continue
                v.add_node(int(t[0]), int(t[1]) > 0, [int(x) for x in t[2:34]], float(t[34]))
        return v

    def save_text(self, path, scoring=0):
        with open(path, "w") as f:
            f.write(f"{self.k} {self.L}  {scoring} {self.weighting}\n")
            for nd in self.nodes[1:]:
                f.write(f"{nd.parent} {1 if nd.is_leaf() else 0} " + " ".join(str(int(b)) for b in nd.descriptor) +
                        f" {nd.weight!r}\n")

    def transform_one(self, feature, levelsup):
        """:1230-1271.  Returns (word_id, weight, nid)."""
        nid_level = self.L - levelsup
        nid = 0
        final_id, level = 0, 0
        while True:
            level += 1
            nodes = self.nodes[final_id].children
            final_id = nodes[0]
            best_d = float(_POP[feature ^ self.nodes[final_id].descriptor].sum())
            for i in nodes[1:]:
                d = float(_POP[feature ^ self.nodes[i].descriptor].sum())
                if d < best_d:
                    best_d, final_id = d, i
            if level == nid_level:
                nid = final_id
            if self.nodes[final_id].is_leaf():
                break
        if nid_level > 0 and level < nid_level:
            nid = final_id        # the reference leaves nid uninitialised here; defined as the leaf itself
        return self.nodes[final_id].word_id, self.nodes[final_id].weight, nid

    def transform(self, features, levelsup):
        """:1137-1206.  Returns (BowVector as sorted [(word, value)], FeatureVector as sorted [(node, [idx...])])."""
        v, fv = {}, {}
        tf = self.weighting in (TF, TF_IDF)
        for i, f in enumerate(features):
            wid, w, nid = self.transform_one(np.asarray(f, np.uint8), levelsup)
            if w > 0:
                if wid in v:
                    if tf:
                        v[wid] += w
                else:
                    v[wid] = w
                fv.setdefault(nid, []).append(i)
        keys = sorted(v)
        if tf and v and self.norm == NORM_NONE:
            nd = float(len(v))
            for kx in keys:
                v[kx] /= nd
        if self.norm != NORM_NONE:
            nrm = 0.0
            if self.norm == NORM_L1:
                for kx in keys:
                    nrm += math.fabs(v[kx])
            else:
                for kx in keys:
                    nrm += v[kx] * v[kx]
                nrm = math.sqrt(nrm)
            if nrm > 0.0:
                for kx in keys:
                    v[kx] /= nrm
        return [(kx, v[kx]) for kx in keys], [(n, fv[n]) for n in sorted(fv)]

    def arrays(self):
        """Flattened form taken by the C ABI / the C++ oracle: parent, is_leaf, descriptors, weights (node order)."""
        n = len(self.nodes)
        parent = np.zeros(n, np.int32)
        leaf = np.zeros(n, np.uint8)
        desc = np.zeros((n, 32), np.uint8)
        w = np.zeros(n, np.float64)
        for nd in self.nodes[1:]:
            parent[nd.id], leaf[nd.id], desc[nd.id], w[nd.id] = nd.parent, nd.is_leaf(), nd.descriptor, nd.weight
        return parent, leaf, desc, w
