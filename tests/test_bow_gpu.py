"""GPU parity of the DBoW2 transform (through the C ABI) against the C++ oracle and the golden vectors: word ids,
node ids, CSR and the float64 BowVector values bit for bit."""
import os

import numpy as np
import pytest

import bow_cases as bc
import orbfe
from common import GOLDEN
from oracle_binding import VocRef

pytestmark = pytest.mark.gpu


def gpu_voc(c):
    return orbfe.ORBVocabulary(c["k"], c["L"], c["weighting"], c["norm"], c["parent"], c["leaf"], c["desc"], c["weight"])


def ref_voc(c):
    return VocRef(c["k"], c["L"], c["weighting"], c["norm"], c["parent"], c["leaf"], c["desc"], c["weight"])


@pytest.mark.parametrize("name", list(bc.BOW_VOC_CASES))
def test_transform_golden_and_oracle(name):
    c = bc.voc_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"bow_{name}.npz"))
    v = gpu_voc(c)
    assert v.size() == int(c["leaf"].sum()) and v.nodes() == len(c["parent"])
    (words, values), fv = v.transform(c["feat"], c["levelsup"])
    assert np.array_equal(words, g["words"]) and np.array_equal(values, g["values"])
    assert np.array_equal(fv.node_ids, g["nodes"]) and np.array_equal(fv.offsets, g["offsets"])
    assert np.array_equal(fv.indices, g["indices"])
    rw, rv, rn, ro, ri = ref_voc(c).transform(c["feat"], c["levelsup"])
    assert np.array_equal(words, rw) and np.array_equal(values, rv) and np.array_equal(fv.indices, ri)


@pytest.mark.parametrize("n", [0, 1, 2, 33, 1000, 5000])
def test_transform_sizes(n):
    c = bc.voc_inputs("voc_k10_L4")
    feat = bc.features(900 + n, c["desc"], max(n, 8))[:n]
    (words, values), fv = gpu_voc(c).transform(feat, 2)
    rw, rv, rn, ro, ri = ref_voc(c).transform(feat, 2) if n else (np.zeros(0, np.uint32), np.zeros(0), np.zeros(0, np.uint32),
                                                                 np.zeros(1, np.int32), np.zeros(0, np.int32))
    assert np.array_equal(words, rw) and np.array_equal(values, rv)
    assert np.array_equal(fv.node_ids, rn) and np.array_equal(fv.offsets, ro) and np.array_equal(fv.indices, ri)


def test_orbvoc_sized_tree_and_batch_from_extractor():
    """A full k = 10, L = 6 tree (1.1 M nodes, 35 MB: the size of ORBvoc) fed from the extractor's device results,
    levelsup = 4 as Frame::ComputeBoW; then the FeatureVectors drive SearchByBoW like Tracking does."""
    import torch
    parent, leaf, desc, weight = bc.make_full_tree(3200, 10, 6)
    assert len(parent) == 1111111
    v = orbfe.ORBVocabulary(10, 6, 0, 1, parent, leaf, desc, weight)
    ref = VocRef(10, 6, 0, 1, parent, leaf, desc, weight)
    w, h, nf, B = 752, 480, 1000, 3
    imgs = orbfe.synth.frames(3300, B, w, h)
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    res = ex.extract_batch(imgs, (0, 0))
    d_kps, d_desc, d_n, d_mono, cap = ex.device_results()
    dev = lambda shape, dt: torch.zeros(shape, dtype=dt, device="cuda")
    words, values = dev((B, cap), torch.int32), dev((B, cap), torch.float64)
    nodes, off, idx = dev((B, cap), torch.int32), dev((B, cap + 1), torch.int32), dev((B, cap), torch.int32)
    nw, nn = dev((B,), torch.int32), dev((B,), torch.int32)
    ex.synchronize()
    v.transform_batch_device(d_desc, cap * 32, d_n, cap, B, 4, words.data_ptr(), values.data_ptr(), nw.data_ptr(), nodes.data_ptr(),
                             off.data_ptr(), idx.data_ptr(), nn.data_ptr())
    v.synchronize()
    fvs = []
    for b in range(B):
        rw, rv, rn, ro, ri = ref.transform(res[b][2], 4)
        k, m = int(nw[b]), int(nn[b])
        assert k == len(rw) and m == len(rn)
        assert np.array_equal(words[b, :k].cpu().numpy().view(np.uint32), rw)
        assert np.array_equal(values[b, :k].cpu().numpy(), rv)
        assert np.array_equal(nodes[b, :m].cpu().numpy().view(np.uint32), rn)
        assert np.array_equal(off[b, :m + 1].cpu().numpy(), ro)
        assert np.array_equal(idx[b, :int(ro[-1])].cpu().numpy(), ri)
        assert m <= 100
        fvs.append(orbfe.FeatureVectorCSR(rn, ro, ri))
    # the host-buffer entry point gives the same
    (hw, hv), hfv = v.transform(res[0][2], 4)
    assert np.array_equal(hfv.indices, fvs[0].indices) and np.array_equal(hw, ref.transform(res[0][2], 4)[0])


def test_load_text_file(tmp_path):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "py"))
    import bow_py
    c = bc.voc_inputs("voc_k10_L3_tf_none")
    pv = bow_py.Vocabulary(c["k"], c["L"], c["weighting"], c["norm"])
    for i in range(1, len(c["parent"])):
        pv.add_node(int(c["parent"][i]), bool(c["leaf"][i]), c["desc"][i], c["weight"][i])
    p = str(tmp_path / "voc.txt")
    pv.save_text(p)
    v = orbfe.ORBVocabulary.loadFromTextFile(p, norm=c["norm"])
    assert v.size() == len(pv.words) and v.nodes() == len(pv.nodes)
    g = np.load(os.path.join(GOLDEN, "bow_voc_k10_L3_tf_none.npz"))
    (words, values), fv = v.transform(c["feat"], c["levelsup"])
    assert np.array_equal(words, g["words"]) and np.array_equal(values, g["values"]) and np.array_equal(fv.indices, g["indices"])
    with pytest.raises(orbfe.OrbfeError):
        orbfe.ORBVocabulary.loadFromTextFile(str(tmp_path / "missing.txt"))


@pytest.mark.parametrize("name", ["voc_k10_L4", "voc_k6_L4_idf_l2"])
def test_cuda_equals_dbow2_source(name, tmp_path):
    """The CUDA transform (vocabulary loaded by orbfe's ORBvoc text loader) against the reference's own vendored DBoW2
    (oracle/_ref/libdbow2_orbvoc.so, prebuilt; see tests/test_reference_dbow2.py) reading the same text file."""
    import oracle_binding as ob
    if not ob.ref_voc_available():
        pytest.skip("oracle/_ref not built")
    from test_reference_dbow2 import write_text
    c = bc.voc_inputs(name)
    write_text(c, tmp_path / "voc.txt")
    ref = ob.DBoW2Vocabulary(tmp_path / "voc.txt")
    v = orbfe.ORBVocabulary.loadFromTextFile(str(tmp_path / "voc.txt"), c["norm"])
    assert v.size() == ref.nwords
    for levelsup in (c["levelsup"], 0, 4):
        (words, values), fv = v.transform(c["feat"], levelsup)
        rw, rv, rn, ro, ri = ref.transform(c["feat"], levelsup)
        assert np.array_equal(words, rw) and np.array_equal(values, rv)
        if levelsup == 0:
            # an early leaf (above level L - levelsup) leaves DBoW2's `NodeId nid` uninitialised (TemplatedVocabulary.h:1162,
            # :1259-1260): the FeatureVector entry of such a feature is undefined in the reference; include/orbfe.h defines it
            # as the leaf's own id.  Only the BowVector is comparable here.
            continue
        assert np.array_equal(fv.node_ids, rn) and np.array_equal(fv.offsets, ro) and np.array_equal(fv.indices, ri)
