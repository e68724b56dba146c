"""GPU parity of the matchers (through the C ABI) against the CPU oracle on identical inputs and
against the committed golden vectors.  Bar (north_star): match indices / distances bit-exact given
identical descriptors; mvuRight / mvDepth float32-exact."""
import os
import zlib

import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe
from common import GOLDEN
from oracle_binding import OrbRef

pytestmark = pytest.mark.gpu


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c & 0xFFFFFFFF


def gold(name):
    return np.load(os.path.join(GOLDEN, f"match_{name}.npz"))


def extract_ref(img, nf, lap):
    ref = OrbRef(nf, *mc.ORB)
    mono, k, d = ref(img, lap)
    return mono, k, d, ref


@pytest.fixture(scope="module")
def matcher():
    return {r: orbfe.ORBmatcher(r, True) for r in (0.7, 0.8)}


def test_knn2_ties_and_golden():
    m = orbfe.ORBmatcher()
    q, t = mc.knn_inputs()
    g = gold("knn_ties")
    idx, dist = m.knn2(q, t)
    assert np.array_equal(idx, g["idx"]) and np.array_equal(dist, g["dist"])
    match, n = m.match_lapping(q, t, 0.7)
    assert np.array_equal(match, g["match"]) and n == (g["match"] >= 0).sum()


@pytest.mark.parametrize("nq,nt", [(1, 1), (3, 1), (5, 2), (127, 129), (128, 128), (1000, 4000), (4000, 4000), (300, 70000),
                                   (0, 10), (10, 0)])
def test_knn2_sizes(nq, nt):
    """Ragged / empty / split-train shapes against the oracle (ties resolved to the lowest train index)."""
    m = orbfe.ORBmatcher()
    q = orbfe.synth.descriptors(nq * 31 + nt, max(nq, 1))[:nq]
    t = orbfe.synth.descriptors(nq + nt * 17, max(nt, 1))[:nt]
    if nt > 100 and nq > 2:
        t[nt // 2] = t[3]        # ties across tiles / splits
        q[2] = t[3]
    idx, dist = m.knn2(q, t)
    ridx, rdist = ob.knn2(q, t) if nq else (idx, dist)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)


def test_knn2_batch_device_from_extractor_results():
    """Sequence matching frame t <-> t+1 straight from the extractor's device results."""
    import torch
    w, h, nf, B = 752, 480, 1000, 4
    imgs = orbfe.synth.frames(3000, B, w, h)
    ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h, max_batch=B)
    res = ex.extract_batch(imgs, (0, 0))
    d_kps, d_desc, d_n, d_mono, cap = ex.device_results()
    m = orbfe.ORBmatcher()
    idx = torch.full((B - 1, cap, 2), -7, dtype=torch.int32, device="cuda")
    dist = torch.full((B - 1, cap, 2), -7, dtype=torch.int32, device="cuda")
    m.knn2_batch_device(d_desc, cap * 32, d_n, cap, d_desc + cap * 32, cap * 32, d_n + 4, cap, B - 1, idx.data_ptr(),
                        dist.data_ptr())
    m.synchronize()
    idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
    for b in range(B - 1):
        q, t = res[b][2], res[b + 1][2]
        ridx, rdist = ob.knn2(q, t)
        assert np.array_equal(idx[b, :len(q)], ridx) and np.array_equal(dist[b, :len(q)], rdist)


@pytest.mark.parametrize("name", list(mc.LAP_CASES))
def test_lapping_match(name):
    c = mc.lap_inputs(name, extract_ref)
    g = gold(name)
    m = orbfe.ORBmatcher()
    match, n = m.match_lapping(c["q"], c["t"], 0.7)
    assert np.array_equal(match, ob.knn2_ratio(c["q"], c["t"], 0.7))
    assert crc(c["q"], c["t"]) == int(g["in_crc"]), "inputs drifted from the committed fixture"
    assert np.array_equal(match, g["match"])
    assert n == (match >= 0).sum() > 100


@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
@pytest.mark.parametrize("mode", ["two_handles", "batch_pair"])
def test_stereo_match(name, mode):
    cam, nf, seed = mc.STEREO_CASES[name]
    w, h = cam["w"], cam["h"]
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    m = orbfe.ORBmatcher()
    if mode == "two_handles":      # Frame.cc:119-122: one extractor per eye
        exL = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h)
        exR = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h)
        _, kL, dL = exL(imL, None, (0, 0))
        _, kR, dR = exR(imR, None, (0, 0))
        u, d, n = m.stereo_match(exL, exR, bf, b)
    else:                          # the pair as one batch of two on one handle
        ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h, max_batch=2)
        (_, kL, dL), (_, kR, dR) = ex.extract_batch(np.stack([imL, imR]), (0, 0))
        u, d, n = m.stereo_match(ex, ex, bf, b, 0, 1)
    u, d = u[:len(kL)], d[:len(kL)]
    # oracle on the very same keypoints / descriptors (its pyramids are bit-identical to the device's)
    rL, rR = OrbRef(nf, *mc.ORB), OrbRef(nf, *mc.ORB)
    rL(imL, (0, 0))
    rR(imR, (0, 0))
    ru, rd, rsad = ob.stereo_match(rL, rR, kL, dL, kR, dR, bf, b)
    assert np.array_equal(d > 0, rd > 0)
    assert np.array_equal(u, ru) and np.array_equal(d, rd)
    assert n == (rd > 0).sum() > 0.4 * len(kL)
    g = gold(name)
    assert crc(kL, dL, kR, dR) == int(g["in_crc"]), "inputs drifted from the committed fixture"
    assert np.array_equal(u, g["u_right"]) and np.array_equal(d, g["depth"])


def test_stereo_batch_device():
    import torch
    cam, nf, seed = mc.STEREO_CASES["stereo_euroc"]
    w, h, P = cam["w"], cam["h"], 3
    ims = []
    for p in range(P):
        ims += list(orbfe.synth.stereo_pair(seed + p, w, h))
    ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h, max_batch=2 * P)
    res = ex.extract_batch(np.stack(ims), (0, 0))
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    m = orbfe.ORBmatcher()
    du = torch.zeros((P, ex.cap), dtype=torch.float32, device="cuda")
    dd = torch.zeros((P, ex.cap), dtype=torch.float32, device="cuda")
    m.stereo_match_batch_device(ex, P, bf, b, du.data_ptr(), dd.data_ptr())
    m.synchronize()
    du, dd = du.cpu().numpy(), dd.cpu().numpy()
    for p in range(P):
        (_, kL, dL), (_, kR, dR) = res[2 * p], res[2 * p + 1]
        rL, rR = OrbRef(nf, *mc.ORB), OrbRef(nf, *mc.ORB)
        rL(ims[2 * p], (0, 0))
        rR(ims[2 * p + 1], (0, 0))
        ru, rd, _ = ob.stereo_match(rL, rR, kL, dL, kR, dR, bf, b)
        assert np.array_equal(du[p, :len(kL)], ru) and np.array_equal(dd[p, :len(kL)], rd)


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_search_by_bow(name, matcher):
    c = mc.bow_inputs(name, extract_ref)
    g = gold(name)
    same_inputs = crc(c["d1"], c["d2"], c["mp1"], c["mp2"]) == int(g["in_crc"])
    assert same_inputs, "inputs drifted from the committed fixture"
    m7 = orbfe.ORBmatcher(0.7, c["check_ori"])
    m8 = orbfe.ORBmatcher(0.8, c["check_ori"])
    b2a, n = m7.SearchByBoW_KF_F(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["k2"], c["nleft"])
    _, rb2a, rn = ob.search_bow(c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], None, c["k2"]["angle"], 0.7,
                                c["check_ori"], strict=False, nleft=c["nleft"])
    assert n == rn and np.array_equal(b2a, rb2a)
    if same_inputs:
        assert n == int(g["n_kf_f"]) and np.array_equal(b2a, g["kf_f"])
    a2b, n = m8.SearchByBoW_KF_KF(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["mp2"], c["k2"])
    ra2b, _, rn = ob.search_bow(c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], c["mp2"], c["k2"]["angle"], 0.8,
                                c["check_ori"], strict=True)
    assert n == rn and np.array_equal(a2b, ra2b)
    if same_inputs:
        assert n == int(g["n_kf_kf"]) and np.array_equal(a2b, g["kf_kf"])


def test_search_by_bow_many_keyframes_one_frame():
    """Relocalisation shape: several keyframes against one frame in one call; each equals the single call."""
    cam, nf = mc.EUROC, 1000
    w, h = cam["w"], cam["h"]

    def ext(seed, shift):
        return extract_ref(orbfe.synth.frame(seed, w, h, 0, shift), nf, (0, 0))

    _, kF, dF, _ = ext(4000, 4)
    kfs = []
    for i, shift in enumerate((0, 2, 7, 11)):
        _, k, d, _ = ext(4000, shift)
        kfs.append((d, mc.fv_of(d), orbfe.synth.mask(100 + i, len(d), 70), k))
    _, k, d, _ = ext(4001, 0)                       # an unrelated keyframe: few or no matches
    kfs.append((d, mc.fv_of(d), orbfe.synth.mask(200, len(d), 70), k))
    m = orbfe.ORBmatcher(0.75, True)
    many, counts = m.SearchByBoW_KFs_F(kfs, dF, mc.fv_of(dF), kF)
    for (d, fv, mp, k), got, cnt in zip(kfs, many, counts):
        _, rb2a, rn = ob.search_bow(d, fv, mp, k["angle"], dF, mc.fv_of(dF), None, kF["angle"], 0.75, True, strict=False)
        assert cnt == rn and np.array_equal(got, rb2a)
    assert counts[0] > 100 and counts[-1] < counts[0] // 4


@pytest.mark.parametrize("name", list(mc.TRI_CASES))
def test_search_for_triangulation(name):
    sc, isc, s2, is2, per, umax = OrbRef(1000, *mc.ORB).tables()
    c = mc.tri_inputs(name, extract_ref, sc, s2)
    g = gold(name)
    m = orbfe.ORBmatcher(0.6, c["check_ori"])
    pairs, n, m12 = m.SearchForTriangulation(c["d1"], c["fv1"], c["mp1"], c["k1"], c["ur1"], c["d2"], c["fv2"], c["mp2"], c["k2"],
                                             c["ur2"], c["F12"], c["ep"], s2, sc, c["only_stereo"], c["coarse"])
    rm12, rn = ob.search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"],
                                       c["ur2"], s2, sc, c["F12"], c["ep"], c["check_ori"], c["only_stereo"], c["coarse"])
    assert n == rn and np.array_equal(m12, rm12)
    assert len(pairs) == n and np.all(np.diff(pairs[:, 0]) > 0)
    assert crc(c["d1"], c["d2"], c["mp1"], c["mp2"], c["F12"]) == int(g["in_crc"]), "inputs drifted from the committed fixture"
    assert n == int(g["n"]) and np.array_equal(m12, g["m12"])


def test_empty_and_disjoint_feature_vectors():
    m = orbfe.ORBmatcher(0.7, True)
    d1, d2 = orbfe.synth.descriptors(1, 50), orbfe.synth.descriptors(2, 60)
    k1, k2 = np.zeros(50, orbfe.KEYPOINT_DTYPE), np.zeros(60, orbfe.KEYPOINT_DTYPE)
    fv1 = orbfe.FeatureVectorCSR.from_node_of_feature(np.full(50, 3))
    fv2 = orbfe.FeatureVectorCSR.from_node_of_feature(np.full(60, 4))      # no shared node
    b2a, n = m.SearchByBoW_KF_F(d1, fv1, np.ones(50, np.uint8), k1, d2, fv2, k2)
    assert n == 0 and (b2a == -1).all() and len(b2a) == 60
    a2b, n = m.SearchByBoW_KF_KF(d1, fv1, np.zeros(50, np.uint8), k1, d2, fv1.__class__.from_node_of_feature(np.full(60, 3)),
                                 np.ones(60, np.uint8), k2)             # shared node, but no MapPoints in KF1
    assert n == 0 and (a2b == -1).all()


@pytest.mark.parametrize("name", list(mc.FISHEYE_CASES))
def test_fisheye_stereo(name):
    """Frame::ComputeStereoFishEyeMatches on the GPU vs the oracle and the golden: the 2-NN / ratio stage is
    bit-exact (same candidate set: margins < 1e29 at the same rows); the KB8 triangulation ends in a 4x4 SVD,
    so accept/reject is compared where neither side is within 1e-3 (relative) of a threshold and depths /
    points to rtol 2e-3."""
    c = mc.fisheye_inputs(name)
    g = gold(name)
    s2 = OrbRef(1000, *mc.ORB).tables()[2]
    m = orbfe.ORBmatcher(0.7, True)
    l2r, r2l, depth, p3d, n = m.ComputeStereoFishEyeMatches(c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], c["mono_r"], c["cam1"],
                                                            c["cam2"], c["R12"], c["t12"], s2)
    rl2r, rr2l, rdepth, rp3d, mg, rn = ob.fisheye_stereo(c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"],
                                                         c["mono_r"], c["R12"], c["t12"], s2)
    tried = mg < 1e29
    assert (l2r[~tried] == -1).all() and (l2r[:c["mono_l"]] == -1).all()
    for ref_l2r, ref_depth, ref_p3d, margins in ((rl2r, rdepth, rp3d, mg), (g["l2r"], g["depth"], g["p3d"], np.minimum(mg, g["margins"]))):
        safe = margins > 1e-3
        assert np.array_equal(l2r[safe], ref_l2r[safe])
        ok = safe & (ref_l2r >= 0)
        assert ok.sum() > 200
        assert np.allclose(depth[ok], ref_depth[ok], rtol=2e-3, atol=1e-5)
        assert np.allclose(p3d[ok], ref_p3d[ok], rtol=2e-3, atol=2e-4)
    assert abs(n - rn) <= (tried & (mg <= 1e-3)).sum()
    assert n == (l2r >= 0).sum()
    assert (depth[l2r < 0] == -1).all()
    # r2l: a right feature claimed by several left ones keeps one of them (the reference keeps the last in
    # query order: Frame.cc:1148 overwrites), and it points back
    want = np.full(len(c["kR"]), -1, np.int32)
    for i in np.nonzero(l2r >= 0)[0]:
        want[l2r[i]] = i
    assert np.array_equal(r2l, want)


def test_fisheye_stereo_empty_sides():
    c = mc.fisheye_inputs("fisheye_tumvi")
    s2 = OrbRef(1000, *mc.ORB).tables()[2]
    m = orbfe.ORBmatcher(0.7, True)
    nL = len(c["kL"])
    # no lapping features on the left / on the right
    out = m.ComputeStereoFishEyeMatches(c["kL"], c["dL"], nL, c["kR"], c["dR"], c["mono_r"], c["cam1"], c["cam2"], c["R12"], c["t12"], s2)
    assert out[4] == 0 and (out[0] == -1).all() and (out[2] == -1).all()
    out = m.ComputeStereoFishEyeMatches(c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], len(c["kR"]), c["cam1"], c["cam2"], c["R12"],
                                        c["t12"], s2)
    assert out[4] == 0 and (out[0] == -1).all() and (out[1] == -1).all()
    # a single right lapping feature: knn has < 2 neighbours, nothing passes (Frame.cc:1141)
    out = m.ComputeStereoFishEyeMatches(c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], len(c["kR"]) - 1, c["cam1"], c["cam2"],
                                        c["R12"], c["t12"], s2)
    assert out[4] == 0


@pytest.mark.parametrize("name", list(mc.TRI_KB8_CASES))
def test_search_for_triangulation_kb8(name):
    """SearchForTriangulation_ with Kannala-Brandt cameras vs the oracle and the golden: rows whose candidate
    pairs all sit > 1e-3 (relative) away from a threshold of the triangulation test must agree exactly."""
    sc, isc, s2 = OrbRef(1000, *mc.ORB).tables()[:3]
    c = mc.tri_kb8_inputs(name)
    g = gold(name)
    m = orbfe.ORBmatcher(0.6, c["check_ori"])
    pairs, n, m12, pts = m.SearchForTriangulationKB8(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["mp2"], c["k2"],
                                                     c["cam_a"], c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"],
                                                     s2, s2, sc, False, c["coarse"])
    rm12, rn, rpts, mg = ob.search_triangulation_kb8(c["k1"], c["d1"], c["fv1"], c["mp1"], s2, c["k2"], c["d2"], c["fv2"], c["mp2"],
                                                     s2, sc, c["cam_a"], c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"],
                                                     c["ep"], c["check_ori"], False, c["coarse"])
    assert n == (m12 >= 0).sum() == len(pairs) and n > 60
    for ref, refn, refp, margins in ((rm12, rn, rpts, mg), (g["m12"], int(g["n"]), g["points"], np.minimum(mg, g["margins"]))):
        # match indices are in the bit-exact class: a row may differ only where some accept / reject threshold of its
        # candidate pairs lies within 1e-3 (relative) of the computed value (4x4 SVD: Jacobi in double here, cv::SVD in the
        # golden).  With the rotation check such a flip can move a histogram bin across the 10 % rule and with it rows that
        # are themselves safe: that cascade is accepted only when an unsafe row really flipped.
        safe = margins > 1e-3
        differs = m12 != ref
        if not (c["check_ori"] and (differs & ~safe).any()):
            assert not (differs & safe).any(), np.flatnonzero(differs & safe)
            assert n == refn or (differs & ~safe).any()
        else:
            assert differs.mean() < 0.01
        ok = (m12 >= 0) & (m12 == ref)
        if not c["coarse"]:
            assert np.allclose(pts[ok], refp[ok], rtol=2e-3, atol=2e-4)
    assert m.SearchForTriangulationKB8(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["mp2"], c["k2"], c["cam_a"],
                                       c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"], s2, s2, sc, True,
                                       c["coarse"])[1] == 0


def test_distinctive_descriptors():
    """MapPoint::ComputeDistinctiveDescriptors batched over 400 map points (0..64 observations, tied medians)."""
    d, off = mc.distinctive_inputs()
    best = orbfe.ORBmatcher().ComputeDistinctiveDescriptors(d, off)
    assert np.array_equal(best, gold("distinctive")["best"]) and np.array_equal(best, ob.distinctive_descriptors(d, off))
    d2, off2 = mc.distinctive_inputs(5201, 3000)
    assert np.array_equal(orbfe.ORBmatcher().ComputeDistinctiveDescriptors(d2, off2), ob.distinctive_descriptors(d2, off2))
    assert len(orbfe.ORBmatcher().ComputeDistinctiveDescriptors(d[:0], off[:1])) == 0


def test_undistort_keypoints():
    """Frame::UndistortKeyPoints on the GPU == the oracle (itself bit-identical to cv2.undistortPoints)."""
    from test_oracle_vs_cv2 import UND_CASES, UND_K, und_points
    k = und_points()
    m = orbfe.ORBmatcher()
    for D in UND_CASES:
        assert m.UndistortKeyPoints(k, UND_K, D).tobytes() == ob.undistort_keypoints(k, UND_K, D).tobytes()
    assert len(m.UndistortKeyPoints(k[:0], UND_K, UND_CASES[0])) == 0
