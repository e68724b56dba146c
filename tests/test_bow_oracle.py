"""CPU: the C++ DBoW2-transform oracle (oracle/cpu_ref/bow_ref.cc) against the golden vectors of the independent
Python restatement (oracle/py/bow_py.py), plus small known-answer cases."""
import os
import zlib

import numpy as np
import pytest

import bow_cases as bc
from common import GOLDEN
from oracle_binding import VocRef


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c & 0xFFFFFFFF


def voc_of(c):
    return VocRef(c["k"], c["L"], c["weighting"], c["norm"], c["parent"], c["leaf"], c["desc"], c["weight"])


@pytest.mark.parametrize("name", list(bc.BOW_VOC_CASES))
def test_transform_matches_golden(name):
    c = bc.voc_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"bow_{name}.npz"))
    assert crc(c["parent"], c["desc"], c["weight"], c["feat"]) == int(g["in_crc"])
    words, values, nodes, off, idx = voc_of(c).transform(c["feat"], c["levelsup"])
    assert np.array_equal(words, g["words"]) and np.array_equal(nodes, g["nodes"])
    assert np.array_equal(values, g["values"])              # float64, bit for bit (sequential sums in map order)
    assert np.array_equal(off, g["offsets"]) and np.array_equal(idx, g["indices"])
    if c["norm"] == 1:
        assert abs(values.sum() - 1.0) < 1e-12


def test_two_level_tree_known_answers():
    """k = 2, L = 2: root -> {A, B}, A -> {a0, a1}, B -> {b0 (stop word), b1}."""
    z, f = np.zeros(32, np.uint8), np.full(32, 255, np.uint8)
    half = np.concatenate([np.zeros(16, np.uint8), np.full(16, 255, np.uint8)])
    q = np.concatenate([np.full(8, 255, np.uint8), np.zeros(24, np.uint8)])
    parent = np.array([0, 0, 0, 1, 1, 2, 2], np.int32)
    leaf = np.array([0, 0, 0, 1, 1, 1, 1], np.uint8)
    desc = np.stack([z, z, f, z, q, f, half])
    w = np.array([0, 0, 0, 2.0, 3.0, 0.0, 5.0])
    v = VocRef(2, 2, 0, 1, parent, leaf, desc, w)
    assert v.nwords == 4
    h1 = half.copy()
    h1[0] = 255                                 # 136 from A, 120 from B -> B; 120 from b0, 8 from b1 -> b1
    feats = np.stack([z, q, q, f, h1])          # -> a0, a1, a1, b0 (stop word), b1
    word, wt, node = v.descend(feats, 1)
    assert word.tolist() == [0, 1, 1, 2, 3] and wt.tolist() == [2.0, 3.0, 3.0, 0.0, 5.0]
    assert node.tolist() == [1, 1, 1, 2, 2]                 # level L - 1 = 1
    words, values, nodes, off, idx = v.transform(feats, 1)
    assert words.tolist() == [0, 1, 3]
    assert np.array_equal(values, np.array([2.0, 6.0, 5.0]) / 13.0)
    assert nodes.tolist() == [1, 2] and off.tolist() == [0, 3, 4] and idx.tolist() == [0, 1, 2, 4]
    # ties go to the first child (strict '<', TemplatedVocabulary.h:1256)
    assert v.descend(np.stack([half]), 1)[2].tolist() == [1]     # 128 from A's z and from B's f


def test_text_format_roundtrip(tmp_path):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "py"))
    import bow_py
    c = bc.voc_inputs("voc_k10_L3_binary")
    v = bow_py.Vocabulary(c["k"], c["L"], c["weighting"], c["norm"])
    for i in range(1, len(c["parent"])):
        v.add_node(int(c["parent"][i]), bool(c["leaf"][i]), c["desc"][i], c["weight"][i])
    p = str(tmp_path / "voc.txt")
    v.save_text(p)
    v2 = bow_py.Vocabulary.load_text(p, c["norm"])
    a, b = v.arrays(), v2.arrays()
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert np.array_equal(a[0], c["parent"]) and np.array_equal(a[1], c["leaf"]) and np.array_equal(a[3], c["weight"])
