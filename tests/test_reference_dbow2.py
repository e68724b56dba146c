"""CPU: the reference's OWN vendored DBoW2 (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, FORB.cpp, BowVector.cpp,
FeatureVector.cpp, ScoringObject.cpp), compiled unmodified into oracle/_ref/libdbow2_orbvoc.so against the OpenCV / boost
stand-ins of oracle/refshim, against the C++ oracle (oracle/cpu_ref/bow_ref.cc) and the golden vectors of the Python
restatement.  The vocabulary goes through DBoW2's own loadFromTextFile (the ORBvoc.txt format)."""
import os
import sys

import numpy as np
import pytest

import bow_cases as bc
import oracle_binding as ob
from common import GOLDEN, ROOT
from test_bow_oracle import voc_of

pytestmark = pytest.mark.skipif(not ob.ref_voc_available(), reason="oracle/_ref not built (needs /root/reference; make -C oracle ref)")
SCORING_OF_NORM = {1: 0, 2: 1, 0: 5}        # DBoW2::ScoringType: L1_NORM, L2_NORM, DOT_PRODUCT (= no normalisation)


def write_text(c, path):
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import bow_py
    v = bow_py.Vocabulary(c["k"], c["L"], c["weighting"], c["norm"])
    for i in range(1, len(c["parent"])):
        v.add_node(int(c["parent"][i]), bool(c["leaf"][i]), c["desc"][i], c["weight"][i])
    v.save_text(str(path), SCORING_OF_NORM[c["norm"]])
    # DBoW2's loader loops `while(!f.eof()) getline(...)` (TemplatedVocabulary.h:1391-1394): a file that ends in a newline
    # makes it parse one more, empty, line into a phantom node whose parent / leaf flag are uninitialised stack values
    # (undefined behaviour; in practice a weight-0 sibling of the last node with an unset descriptor).  The format itself
    # has no such node and neither have the oracle, the Python restatement or orbfe's loader, so the pin feeds DBoW2 the
    # file without its final newline.  test_phantom_node_of_trailing_newline documents the effect.
    txt = open(path).read().rstrip("\n")
    open(path, "w").write(txt)


@pytest.mark.parametrize("name", list(bc.BOW_VOC_CASES))
def test_dbow2_source_equals_oracle_and_golden(name, tmp_path):
    c = bc.voc_inputs(name)
    write_text(c, tmp_path / "voc.txt")
    ref = ob.DBoW2Vocabulary(tmp_path / "voc.txt")
    orc = voc_of(c)
    assert ref.nwords == orc.nwords
    got = ref.transform(c["feat"], c["levelsup"])
    want = orc.transform(c["feat"], c["levelsup"])
    for a, b, what in zip(got, want, ("words", "values", "node ids", "offsets", "indices")):
        assert np.array_equal(a, b), what                                   # values: float64 bit for bit
    g = np.load(os.path.join(GOLDEN, f"bow_{name}.npz"))
    assert np.array_equal(got[0], g["words"]) and np.array_equal(got[1], g["values"]) and np.array_equal(got[2], g["nodes"])
    assert np.array_equal(got[3], g["offsets"]) and np.array_equal(got[4], g["indices"])


def test_dbow2_source_full_tree_from_synth_generator(tmp_path):
    """The k = 10 full tree of orbfe.synth.vocabulary (what bench.py and the GPU tests use at 1.1 M nodes), here L = 4."""
    import orbfe
    parent, leaf, desc, weight = orbfe.synth.vocabulary(4242, 10, 4, 1)
    c = dict(k=10, L=4, weighting=0, norm=1, parent=parent, leaf=leaf, desc=desc, weight=weight)
    write_text(c, tmp_path / "voc.txt")
    ref, orc = ob.DBoW2Vocabulary(tmp_path / "voc.txt"), voc_of(c)
    feat = bc.features(99, desc, 2000)
    for levelsup in (4, 2, 0):
        for a, b in zip(ref.transform(feat, levelsup), orc.transform(feat, levelsup)):
            assert np.array_equal(a, b)
    assert len(ref.transform(feat[:0], 4)[0]) == 0                           # empty feature set


def test_phantom_node_of_trailing_newline(tmp_path):
    c = bc.voc_inputs("voc_k10_L3_binary")
    write_text(c, tmp_path / "voc.txt")
    clean = ob.DBoW2Vocabulary(tmp_path / "voc.txt")
    open(tmp_path / "voc_nl.txt", "w").write(open(tmp_path / "voc.txt").read() + "\n")
    phantom = ob.DBoW2Vocabulary(tmp_path / "voc_nl.txt")
    assert clean.nwords == voc_of(c).nwords
    assert phantom.nwords in (clean.nwords, clean.nwords + 1)       # +1 whenever the stale leaf flag was set
