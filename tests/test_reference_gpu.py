"""GPU: the CUDA matchers (through the C ABI) compared DIRECTLY with the reference's own ORBmatcher.cc (oracle/_ref/
liborbslam3_orbmatcher.so, prebuilt; tests/test_reference_orbmatcher.py says what that library is): SearchByBoW x2,
SearchForTriangulation_, SearchForInitialization, SearchByProjection (local map points, last frame, Sim3, Fuse)."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe
import proj_cases as pc
from test_match_oracle import extract, tables
from test_proj_gpu import gpu_search
from test_proj_oracle import run_kf, run_lf, run_mp

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ob.ref_matcher_available(), reason="oracle/_ref not built")]


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_search_by_bow_vs_reference_source(name):
    c = mc.bow_inputs(name, extract)
    b2a, n = orbfe.ORBmatcher(0.7, c["check_ori"]).SearchByBoW_KF_F(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["k2"], c["nleft"])
    _, rb2a, rn = ob.ref_search_bow(0, c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], None, c["k2"]["angle"], 0.7,
                                    c["check_ori"], c["nleft"])
    assert n == rn and np.array_equal(b2a, rb2a) and n > 100
    a2b, n = orbfe.ORBmatcher(0.8, c["check_ori"]).SearchByBoW_KF_KF(c["d1"], c["fv1"], c["mp1"], c["k1"], c["d2"], c["fv2"], c["mp2"], c["k2"])
    ra2b, _, rn = ob.ref_search_bow(1, c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], c["mp2"], c["k2"]["angle"], 0.8,
                                    c["check_ori"])
    assert n == rn and np.array_equal(a2b, ra2b)


@pytest.mark.parametrize("name", list(mc.TRI_CASES))
def test_search_for_triangulation_vs_reference_source(name):
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    m = orbfe.ORBmatcher(0.6, c["check_ori"])
    pairs, n = m.SearchForTriangulation(c["d1"], c["fv1"], c["mp1"], c["k1"], c["ur1"], c["d2"], c["fv2"], c["mp2"], c["k2"], c["ur2"],
                                        c["F12"], c["ep"], s2, sc, c["only_stereo"], c["coarse"])[:2]
    rm12, rn = ob.ref_search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                                           s2, sc, c["F12"], c["ep"], 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    want = np.stack([np.flatnonzero(rm12 >= 0), rm12[rm12 >= 0]], 1)
    assert n == rn and np.array_equal(np.asarray(pairs).reshape(-1, 2), want)


@pytest.mark.parametrize("name", list(pc.INIT_CASES))
def test_search_for_initialization_vs_reference_source(name):
    c = pc.init_inputs(name)
    n, m12, prev = orbfe.ORBmatcher(c["nnratio"], c["ori"]).SearchForInitialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"])
    rn, rm12, rprev = ob.ref_search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    assert n == rn and np.array_equal(m12, rm12) and np.array_equal(prev, rprev)


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_local_points_vs_reference_source(name):
    c = pc.mp_inputs(name)
    res, n, _, _ = run_mp(c, gpu_search(c["nnratio"]))
    rn, rres = ob.ref_search_by_projection_mappoints(c["F"], c["mp"], c["th"], c["nnratio"], c["far"], c["th_far"])
    assert n == rn and np.array_equal(res, rres)


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_track_last_frame_vs_reference_source(name):
    c = pc.lf_inputs(name)
    Trl = np.eye(4, dtype=np.float32)
    Trl[:3, 3] = (-0.101, 0.0, 0.0)
    res, n, _, _ = run_lf(c, gpu_search(0.9, c["ori"]))
    rn, rres = ob.ref_search_by_projection_lastframe(c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], pc.CAM, Trl, c["ori"])
    assert n == rn and np.array_equal(np.where(res >= 0, res, -1), rres)


@pytest.mark.parametrize("name", list(pc.KF_CASES))
def test_keyframe_side_vs_reference_source(name):
    c = pc.kf_inputs(name)
    res, n, _, _ = run_kf(c, gpu_search(0.9, False))
    if c["kind"] == "sim3":
        rn, rres = ob.ref_search_by_projection_sim3(c["KF"], c["pts"], c["Scw"], c["th"], c["ratio"], pc.CAM, pc.LOG_SF)
    else:
        sim3 = c["kind"] == "fuse_sim3"
        rn, rres = ob.ref_fuse(c["KF"], c["pts"], None if sim3 else c["Tcw"], c["Scw"] if sim3 else None, c["th"], c["kind"] == "fuse_right",
                               pc.CAM, pc.LOG_SF)
    assert n == rn and np.array_equal(res, rres)


@pytest.mark.skipif(not ob.ref_mappoint_available(), reason="oracle/_ref/liborbslam3_mappoint.so not built")
@pytest.mark.parametrize("seed,npoints", [(5200, 400), (5203, 4000)])
def test_distinctive_descriptors_vs_reference_source(seed, npoints):
    """k_distinctive == MapPoint::ComputeDistinctiveDescriptors of the reference's own MapPoint.cc (mono keyframes and
    fisheye keyframes that contribute a left and a right row)."""
    d, off = mc.distinctive_inputs(seed, npoints)
    best = orbfe.ORBmatcher().ComputeDistinctiveDescriptors(d, off)
    assert np.array_equal(best, ob.ref_distinctive_descriptors(d, off, False))
    assert np.array_equal(best, ob.ref_distinctive_descriptors(d, off, True))


frame_lib = pytest.mark.skipif(not ob.ref_frame_available(), reason="oracle/_ref/liborbslam3_frame.so not built")


@frame_lib
@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_match_vs_reference_source(name):
    """k_stereo_search / k_stereo_filter == Frame::ComputeStereoMatches of the reference's own Frame.cc, float32-exact, on
    the features and pyramids the device extracted."""
    from oracle_binding import OrbRef
    cam, nf, seed = mc.STEREO_CASES[name]
    w, h = cam["w"], cam["h"]
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h, max_batch=2)
    (_, kL, dL), (_, kR, dR) = ex.extract_batch(np.stack([imL, imR]), (0, 0))
    u, d, n = orbfe.ORBmatcher().stereo_match(ex, ex, bf, b, 0, 1)
    levL = [ex.image_pyramid(l, 0)[0] for l in range(8)]
    levR = [ex.image_pyramid(l, 1)[0] for l in range(8)]
    ru, rd = ob.ref_stereo_matches((nf,) + mc.ORB, levL, levR, kL, dL, kR, dR, bf, b)
    assert np.array_equal(u[:len(kL)], ru) and np.array_equal(d[:len(kL)], rd)
    assert n == (rd > 0).sum() > 0.4 * len(kL)


@frame_lib
@pytest.mark.parametrize("name", list(mc.FISHEYE_CASES))
def test_fisheye_stereo_vs_reference_source(name):
    """ComputeStereoFishEyeMatches on the GPU vs the reference's own Frame.cc + KannalaBrandt8.cpp (decisions away from the
    thresholds: both sides end in their own 4x4 null-vector solve)."""
    c = mc.fisheye_inputs(name)
    s2 = tables()[2]
    l2r, r2l, depth, p3d, n = orbfe.ORBmatcher(0.7, True).ComputeStereoFishEyeMatches(
        c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], c["mono_r"], c["cam1"], c["cam2"], c["R12"], c["t12"], s2)
    args = (c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], c["mono_r"], c["R12"], c["t12"], s2)
    rl2r, rr2l, rdepth, rp3d, rn = ob.ref_fisheye_stereo(*args)
    mg = ob.fisheye_stereo(*args)[4]
    safe = mg > 1e-3
    assert np.array_equal(l2r[safe], rl2r[safe])
    ok = safe & (rl2r >= 0)
    assert ok.sum() > 200
    assert np.allclose(depth[ok], rdepth[ok], rtol=2e-3, atol=1e-5) and np.allclose(p3d[ok], rp3d[ok], rtol=2e-3, atol=2e-4)
    assert abs(n - rn) <= ((mg < 1e29) & ~safe).sum()
    both = (l2r >= 0) & (rl2r >= 0)
    assert np.array_equal(r2l[l2r[both]] >= 0, rr2l[rl2r[both]] >= 0)


@frame_lib
def test_undistort_keypoints_vs_reference_source():
    from test_reference_frame import UND_CASES, UND_K, und_points
    k = und_points()
    m = orbfe.ORBmatcher()
    for D in UND_CASES:
        assert m.UndistortKeyPoints(k, UND_K, D).tobytes() == ob.ref_undistort_keypoints(k, UND_K, D).tobytes()


@frame_lib
@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_frame_end_to_end_vs_reference_constructor(name):
    """Two images -> the product (batched extraction of both eyes, stereo matcher, undistortion) vs the reference's own stereo
    Frame constructor (Frame.cc:98-200: its ORBextractor on two threads, UndistortKeyPoints, ComputeStereoMatches) run from the
    same two images: mvKeys, mvKeysRight, descriptors, mvuRight, mvDepth, mvKeysUn byte for byte."""
    cam, nf, seed = mc.STEREO_CASES[name]
    w, h = cam["w"], cam["h"]
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    bf = float(np.float32(cam["bf"]))
    K4 = np.array([cam["fx"], cam.get("fy", cam["fx"]), cam.get("cx", w / 2), cam.get("cy", h / 2)], np.float32)
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)
    b = float(np.float32(np.float32(bf) / K4[0]))
    ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h, max_batch=2)
    (_, kL, dL), (_, kR, dR) = ex.extract_batch(np.stack([imL, imR]), (0, 0))
    m = orbfe.ORBmatcher()
    u, d, n = m.stereo_match(ex, ex, bf, b, 0, 1)
    kun = m.UndistortKeyPoints(kL, K4, D)
    F = ob.ref_stereo_frame(imL, imR, (nf,) + mc.ORB, K4, D, bf)
    assert F["kps"].tobytes() == kL.tobytes() and F["kps_r"].tobytes() == kR.tobytes()
    assert np.array_equal(F["desc"], dL) and np.array_equal(F["desc_r"], dR)
    assert np.array_equal(F["u_right"], u[:len(kL)]) and np.array_equal(F["depth"], d[:len(kL)])
    assert F["kps_un"].tobytes() == kun.tobytes()
    assert n == (F["depth"] > 0).sum()


@frame_lib
def test_fisheye_frame_end_to_end_vs_reference_constructor():
    """Two images -> the product (extraction of each eye with its lapping area, fisheye stereo matcher) vs the reference's own
    stereo-fisheye Frame constructor (Frame.cc:1017-1114) run from the same images."""
    from test_reference_frame import fisheye_frame_case
    imL, imR, nf, lapL, lapR = fisheye_frame_case()
    h, w = imL.shape
    F = ob.ref_fisheye_frame(imL, imR, (nf,) + mc.ORB, mc.KB8_LEFT, mc.KB8_RIGHT, lapL, lapR, mc.TLR)
    ex = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h)
    monoL, kL, dL = ex(imL, None, lapL)
    monoR, kR, dR = ex(imR, None, lapR)
    assert (F["nleft"], F["nright"], F["mono_left"], F["mono_right"]) == (len(kL), len(kR), monoL, monoR)
    assert F["kps"][:len(kL)].tobytes() == kL.tobytes() and F["kps"][len(kL):].tobytes() == kR.tobytes()
    assert np.array_equal(F["desc"], np.concatenate([dL, dR]))
    s2 = tables()[2]
    l2r, r2l, depth, p3d, n = orbfe.ORBmatcher(0.7, True).ComputeStereoFishEyeMatches(kL, dL, monoL, kR, dR, monoR, mc.KB8_LEFT, mc.KB8_RIGHT,
                                                                                      mc.TLR[:, :3], mc.TLR[:, 3], s2)
    mg = ob.fisheye_stereo(mc.KB8_LEFT, mc.KB8_RIGHT, kL, dL, monoL, kR, dR, monoR, mc.TLR[:, :3], mc.TLR[:, 3], s2)[4]
    safe = mg > 1e-3
    assert (mg < 1e29).sum() > 100
    assert np.array_equal(l2r[safe], F["l2r"][safe])
    ok = safe & (F["l2r"] >= 0)
    if ok.any():
        assert np.allclose(depth[ok], F["depth"][ok], rtol=2e-3, atol=1e-5) and np.allclose(p3d[ok], F["p3d"][ok], rtol=2e-3, atol=2e-4)


@frame_lib
def test_mono_frame_and_bow_end_to_end_vs_reference_constructor(tmp_path):
    """BASELINE config 0 end to end: one 752x480 image -> the product (extraction with the lapping range (0, 1000), undistortion,
    BoW transform with the vocabulary read by orbfe's ORBvoc loader) vs the reference's own monocular Frame constructor +
    Frame::ComputeBoW (Frame.cc:284-378, 721-728; DBoW2 reading the same text file): mvKeys, descriptors, mvKeysUn, mBowVec (float64
    bit patterns) and mFeatVec byte for byte."""
    from test_reference_dbow2 import write_text
    img = orbfe.synth.frame(1000, 752, 480)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)
    parent, leaf, desc, weight = orbfe.synth.vocabulary(4242, 10, 5, 1)
    c = dict(k=10, L=5, weighting=0, norm=1, parent=parent, leaf=leaf, desc=desc, weight=weight)
    write_text(c, tmp_path / "voc.txt")
    F = ob.ref_mono_frame(img, (1000,) + mc.ORB, K4, D, tmp_path / "voc.txt")
    ex = orbfe.ORBextractor(1000, *mc.ORB, max_width=752, max_height=480)
    mono, k, d = ex(img, None, (0, 1000))
    assert F["kps"].tobytes() == k.tobytes() and np.array_equal(F["desc"], d) and len(k) >= 1000
    assert F["kps_un"].tobytes() == orbfe.ORBmatcher().UndistortKeyPoints(k, K4, D).tobytes()
    v = orbfe.ORBVocabulary.loadFromTextFile(str(tmp_path / "voc.txt"), 1)
    (words, values), fv = v.transform(d, 4)
    rw, rv, rn, ro, ri = F["bow"]
    assert np.array_equal(words, rw) and np.array_equal(values, rv)
    assert np.array_equal(fv.node_ids, rn) and np.array_equal(fv.offsets, ro) and np.array_equal(fv.indices, ri)


def _gpu_stereo_configs():
    from test_reference_frame import _stereo_random_configs
    return _stereo_random_configs(20261108, 5)


@frame_lib
@pytest.mark.parametrize("cfg", _gpu_stereo_configs(), ids=lambda c: "%dx%d_n%d_s%.2f_l%d_t%d-%d" % c)
def test_stereo_frame_random_configs_vs_reference_constructor(cfg):
    """Random shapes / budgets / scale factors / level counts / thresholds: product (batched extraction + stereo matcher +
    undistortion) vs the reference's own stereo Frame constructor run from the same two images, byte for byte."""
    w, h, nf, sf, nl, ini, mn = cfg
    imL, imR = orbfe.synth.stereo_pair(9100 + w + h, w, h)
    K4 = np.array([0.6 * w, 0.6 * w, w / 2 - 3.25, h / 2 + 1.5], np.float32)
    D = np.array([-0.21, 0.05, 0.0004, -0.0003], np.float32)
    bf = float(np.float32(0.11 * K4[0]))
    b = float(np.float32(np.float32(bf) / K4[0]))
    F = ob.ref_stereo_frame(imL, imR, (nf, sf, nl, ini, mn), K4, D, bf)
    ex = orbfe.ORBextractor(nf, sf, nl, ini, mn, max_width=w, max_height=h, max_batch=2)
    (_, kL, dL), (_, kR, dR) = ex.extract_batch(np.stack([imL, imR]), (0, 0))
    m = orbfe.ORBmatcher()
    u, d, n = m.stereo_match(ex, ex, bf, b, 0, 1)
    assert F["kps"].tobytes() == kL.tobytes() and F["kps_r"].tobytes() == kR.tobytes()
    same = (F["desc"] == dL).all(1).sum() + (F["desc_r"] == dR).all(1).sum()
    assert same >= 0.9999 * (len(dL) + len(dR))
    if same == len(dL) + len(dR):
        assert np.array_equal(F["u_right"], u[:len(kL)]) and np.array_equal(F["depth"], d[:len(kL)])
    assert F["kps_un"].tobytes() == m.UndistortKeyPoints(kL, K4, D).tobytes()
