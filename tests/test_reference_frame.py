"""The reference's own src/Frame.cc, src/CameraModels/KannalaBrandt8.cpp and Pinhole.cpp (compiled unmodified into
oracle/_ref/liborbslam3_frame.so, recipe in oracle/Makefile) against the C++ oracle and the golden fixtures:
Frame::ComputeStereoMatches (Frame.cc:794-964), Frame::ComputeStereoFishEyeMatches (:1116-1156) with
KannalaBrandt8::TriangulateMatches, Frame::UndistortKeyPoints (:730-763)."""
import os

import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
from common import GOLDEN
from test_match_oracle import extract, tables

pytestmark = pytest.mark.skipif(not ob.ref_frame_available(), reason="oracle/_ref/liborbslam3_frame.so not built")


def gold(name):
    return np.load(os.path.join(GOLDEN, f"match_{name}.npz"))


def levels(ref, n=8):
    return [ref.level(l) for l in range(n)]


@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_matches_oracle_equals_reference_source(name):
    c = mc.stereo_inputs(name, extract)
    (_, kL, dL, rL), (_, kR, dR, rR) = c["L"], c["R"]
    u, d = ob.ref_stereo_matches((c["nf"],) + mc.ORB, levels(rL), levels(rR), kL, dL, kR, dR, c["bf"], c["b"])
    uo, do, _ = ob.stereo_match(rL, rR, kL, dL, kR, dR, c["bf"], c["b"])
    g = gold(name)
    assert np.array_equal(u, uo) and np.array_equal(d, do)                      # float32-exact
    assert np.array_equal(u, g["u_right"]) and np.array_equal(d, g["depth"])
    assert (d > 0).sum() > 0.4 * len(kL)


@pytest.mark.parametrize("name", list(mc.FISHEYE_CASES))
def test_fisheye_stereo_oracle_vs_reference_source(name):
    """The reference's own knnMatch / Lowe / TriangulateMatches control flow.  Its cv::SVD is a stand-in (double Jacobi), the
    oracle's is its own double Jacobi, the golden's is the real cv2.SVDecomp: decisions compared away from the thresholds."""
    c = mc.fisheye_inputs(name)
    s2 = tables()[2]
    args = (c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], c["mono_r"], c["R12"], c["t12"], s2)
    l2r, r2l, depth, p3d, n = ob.ref_fisheye_stereo(*args)
    ol2r, or2l, odepth, op3d, mg, on = ob.fisheye_stereo(*args)
    g = gold(name)
    safe = np.minimum(mg, g["margins"]) > 1e-3
    assert safe.sum() > 0.95 * (g["margins"] < 1e29).sum()
    assert np.array_equal(l2r[safe], ol2r[safe]) and np.array_equal(l2r[safe], g["l2r"][safe])
    ok = safe & (l2r >= 0)
    assert ok.sum() > 200
    assert np.allclose(depth[ok], odepth[ok], rtol=2e-3, atol=1e-5) and np.allclose(depth[ok], g["depth"][ok], rtol=2e-3, atol=1e-5)
    assert np.allclose(p3d[ok], op3d[ok], rtol=2e-3, atol=2e-4)
    assert abs(n - on) <= (~safe & (g["margins"] < 1e29)).sum()
    # unmatched rows keep the reference's defaults
    assert (depth[l2r < 0] == -1).all()
    m = l2r[l2r >= 0]
    assert (r2l[m] >= 0).all()


def und_points(n=3000, seed=77):
    import orbfe
    rng = np.random.default_rng(seed)
    k = np.zeros(n, orbfe.KEYPOINT_DTYPE)
    k["x"], k["y"] = rng.uniform(0, 752, n).astype(np.float32), rng.uniform(0, 480, n).astype(np.float32)
    k["octave"], k["angle"], k["size"], k["class_id"] = rng.integers(0, 8, n), rng.uniform(0, 360, n), 31, -1
    return k


UND_K = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
UND_CASES = [np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32),
             np.array([-0.3, 0.1, 0.001, -0.0005, -0.02], np.float32),
             np.array([0.0, 0.07, 0.0, 0.0], np.float32)]


@pytest.mark.parametrize("ci", range(len(UND_CASES)))
def test_undistort_keypoints_oracle_equals_reference_source(ci):
    """Frame::UndistortKeyPoints' own lines (the k1 == 0 early exit, the N x 2 matrix, reshape, the copy back that keeps
    octave / angle / size) on top of the undistortPoints restatement that is held to cv2."""
    k = und_points()
    out = ob.ref_undistort_keypoints(k, UND_K, UND_CASES[ci])
    assert out.tobytes() == ob.undistort_keypoints(k, UND_K, UND_CASES[ci]).tobytes()
    if UND_CASES[ci][0] == 0:
        assert out.tobytes() == k.tobytes()             # Frame.cc:732-736
    else:
        assert np.abs(out["x"] - k["x"]).max() > 1.0


def test_kb8_project_unproject_oracle_equals_reference_source():
    """KannalaBrandt8::unproject (KannalaBrandt8.cpp:96-123, Newton iterations with the precision stop) of the reference's own
    object code == oracle/cpu_ref/kb8_ref.cc bit for bit.  project (:24-41) ends in `fx * r * cos(psi) + cx` with a float
    psi: whether that is cosf or the double cos depends on which <math.h> / `using namespace std` the translation unit happens
    to see through the OpenCV headers (the stand-in build sees std::cos(float), the oracle and the GPU take the double one);
    the two differ by at most one float ulp of the pixel coordinate, which is the tolerance here."""
    rng = np.random.default_rng(5)
    P = np.stack([rng.uniform(-3, 3, 4000), rng.uniform(-2, 2, 4000), rng.uniform(0.2, 10, 4000)], 1).astype(np.float32)
    L = ob.lib()
    for cam in (mc.KB8_LEFT, mc.KB8_RIGHT):
        cam = np.ascontiguousarray(cam, np.float32)
        uv = ob.ref_kb8_project(cam, P)
        ray = ob.ref_kb8_unproject(cam, uv)
        ouv, oray = np.zeros_like(uv), np.zeros_like(ray)
        for i in range(len(P)):
            L.orbref_kb8_project(ob._p(cam), ob._p(P[i]), ob._p(ouv[i]))
            L.orbref_kb8_unproject(ob._p(cam), float(ouv[i, 0]), float(ouv[i, 1]), ob._p(oray[i]))
        assert np.abs(uv - ouv).max() <= 6.2e-5 and (uv == ouv).mean() > 0.8       # ulp(512) = 6.1e-5
        ray_o = ob.ref_kb8_unproject(cam, ouv)
        assert np.array_equal(ray_o, oray)
        assert np.allclose(ray[:, :2] * P[:, 2:3], P[:, :2], atol=2e-3)


@pytest.mark.parametrize("nleft_frac", [None, 0.55], ids=["mono", "fisheye_rig"])
def test_features_in_area_oracle_equals_reference_source(nleft_frac):
    """Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea of the reference's own Frame.cc: same indices in the same
    (cell-major, push) order as the oracle's grid, for random windows incl. level ranges, undistorted points outside the image
    bounds (dropped by PosInGrid) and windows hanging over the border."""
    import orbfe
    import proj_cases as pc
    from orbfe import projection as pj
    rng = np.random.default_rng(11)
    n = 3000
    k = np.zeros(n, orbfe.KEYPOINT_DTYPE)
    k["x"] = rng.uniform(-30, 780, n).astype(np.float32)           # some outside (-12.5, 760) x (-8, 490)
    k["y"] = rng.uniform(-20, 500, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n)
    nleft = -1 if nleft_frac is None else int(n * nleft_frac)
    F = pj.GridFrame(k, np.zeros((n, 32), np.uint8), (-12.5, -8.0, 760.25, 490.5), pc.SCALE, nleft=nleft)
    nq = 1500
    win = np.zeros((nq, 6), np.float32)
    win[:, 0], win[:, 1] = rng.uniform(-60, 820, nq), rng.uniform(-50, 540, nq)
    win[:, 2] = rng.choice([2.5, 4.0, 7.2, 15.0, 40.0, 100.0], nq)
    lo = rng.integers(-1, 6, nq)
    hi = np.where(rng.random(nq) < 0.5, -1, lo + rng.integers(0, 4, nq))
    win[:, 3], win[:, 4] = lo, hi
    if nleft != -1:
        win[:, 5] = rng.random(nq) < 0.5
    got = ob.ref_features_in_area(F, win)
    total = 0
    for q in range(nq):
        want = ob.features_in_area(F, win[q, 0], win[q, 1], win[q, 2], int(win[q, 3]), int(win[q, 4]), bool(win[q, 5]))
        assert np.array_equal(got[q], want), q
        total += len(want)
    assert total > 8 * nq


def test_pinhole_project_equals_reference_source():
    """Pinhole::project of the reference's own Pinhole.cpp == the float32 restatement the window construction uses
    (orbfe/projection.py pinhole_project)."""
    from orbfe import projection as pj
    rng = np.random.default_rng(3)
    P = np.stack([rng.uniform(-4, 4, 5000), rng.uniform(-3, 3, 5000), rng.uniform(0.1, 30, 5000)], 1).astype(np.float32)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    uv = ob.ref_pinhole_project(K4, P)
    got = np.array([pj.pinhole_project(K4, p) for p in P], np.float32)
    assert np.array_equal(uv, got)


@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_frame_constructor_end_to_end(name):
    """Two images -> the reference's own stereo Frame constructor (its ORBextractor on two threads, UndistortKeyPoints,
    ComputeStereoMatches) == the oracle's extraction + stereo matching + undistortion: key points, descriptors, mvuRight,
    mvDepth, mvKeysUn and the image bounds byte for byte."""
    c = mc.stereo_inputs(name, extract)
    (_, kL, dL, rL), (_, kR, dR, rR) = c["L"], c["R"]
    cam = c["cam"]
    K4 = np.array([cam["fx"], cam.get("fy", cam["fx"]), cam.get("cx", cam["w"] / 2), cam.get("cy", cam["h"] / 2)], np.float32)
    for D in (np.zeros(4, np.float32), np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)):
        F = ob.ref_stereo_frame(c["imL"], c["imR"], (c["nf"],) + mc.ORB, K4, D, c["bf"])
        assert F["kps"].tobytes() == kL.tobytes() and np.array_equal(F["desc"], dL)
        assert F["kps_r"].tobytes() == kR.tobytes() and np.array_equal(F["desc_r"], dR)
        b = float(np.float32(np.float32(c["bf"]) / K4[0]))
        u, d, _ = ob.stereo_match(rL, rR, kL, dL, kR, dR, c["bf"], b)
        assert np.array_equal(F["u_right"], u) and np.array_equal(F["depth"], d)
        assert F["kps_un"].tobytes() == ob.undistort_keypoints(kL, K4, D).tobytes()
        if D[0] == 0:
            assert F["bounds"].tolist() == [0.0, 0.0, float(cam["w"]), float(cam["h"])]      # Frame.cc:765-792
        else:
            assert F["bounds"][0] < 0 and F["bounds"][2] > cam["w"]


def fisheye_frame_case():
    w, h, nf, seed, lapL, lapR = mc.LAP_CASES["lap_tumvi"]
    import orbfe
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    return imL, imR, nf, lapL, lapR


def test_fisheye_frame_constructor_end_to_end():
    """Two images -> the reference's own stereo-fisheye Frame constructor (extraction with the lapping areas on two threads,
    ComputeStereoFishEyeMatches, stacked descriptors) vs the oracle pipeline: key points (mono part first, lapping part last),
    monoLeft / monoRight and the stacked descriptors byte for byte; matches where no threshold is within 1e-3."""
    imL, imR, nf, lapL, lapR = fisheye_frame_case()
    F = ob.ref_fisheye_frame(imL, imR, (nf,) + mc.ORB, mc.KB8_LEFT, mc.KB8_RIGHT, lapL, lapR, mc.TLR)
    monoL, kL, dL, _ = extract(imL, nf, lapL)
    monoR, kR, dR, _ = extract(imR, nf, lapR)
    assert (F["nleft"], F["nright"], F["mono_left"], F["mono_right"]) == (len(kL), len(kR), monoL, monoR)
    assert F["kps"][:len(kL)].tobytes() == kL.tobytes() and F["kps"][len(kL):].tobytes() == kR.tobytes()
    assert np.array_equal(F["desc"], np.concatenate([dL, dR]))
    assert 0 < monoL < len(kL) and 0 < monoR < len(kR)
    s2 = tables()[2]
    l2r, r2l, depth, p3d, mg, n = ob.fisheye_stereo(mc.KB8_LEFT, mc.KB8_RIGHT, kL, dL, monoL, kR, dR, monoR, mc.TLR[:, :3], mc.TLR[:, 3], s2)
    tried = mg < 1e29
    assert tried.sum() > 100                       # Lowe-ratio survivors that reached the triangulation
    safe = mg > 1e-3
    assert np.array_equal(F["l2r"][safe], l2r[safe])
    assert (F["l2r"][~tried] == -1).all()
    ok = safe & (l2r >= 0)
    if ok.any():
        assert np.allclose(F["depth"][ok], depth[ok], rtol=2e-3, atol=1e-5) and np.allclose(F["p3d"][ok], p3d[ok], rtol=2e-3, atol=2e-4)


def test_mono_frame_constructor_and_bow_end_to_end(tmp_path):
    """BASELINE config 0 as the reference runs it: one 752x480 image -> the reference's own monocular Frame constructor
    (ORBextractor(1000, 1.2, 8, 20, 7) with the lapping range (0, 1000), UndistortKeyPoints) -> Frame::ComputeBoW with a k = 10
    vocabulary loaded by DBoW2's own loader == the oracle's extraction, undistortion and BoW transform, byte for byte
    (BowVector values: float64 bit patterns)."""
    import orbfe
    from test_bow_oracle import voc_of
    from test_reference_dbow2 import write_text
    img = orbfe.synth.frame(1000, 752, 480)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)
    parent, leaf, desc, weight = orbfe.synth.vocabulary(4242, 10, 5, 1)      # L = 5: levelsup 4 groups features by the 10 level-1 nodes
    c = dict(k=10, L=5, weighting=0, norm=1, parent=parent, leaf=leaf, desc=desc, weight=weight)
    write_text(c, tmp_path / "voc.txt")
    F = ob.ref_mono_frame(img, (1000,) + mc.ORB, K4, D, tmp_path / "voc.txt")
    mono, k, d, _ = extract(img, 1000, (0, 1000))
    assert F["mono"] == -1                      # the constructor resets monoLeft after the extraction (Frame.cc:356)
    assert F["kps"].tobytes() == k.tobytes() and np.array_equal(F["desc"], d)
    assert F["kps_un"].tobytes() == ob.undistort_keypoints(k, K4, D).tobytes()
    want = voc_of(c).transform(d, 4)
    for a, b, what in zip(F["bow"], want, ("words", "values", "node ids", "offsets", "indices")):
        assert np.array_equal(a, b), what
    assert len(F["bow"][0]) > 300 and len(F["bow"][2]) == 10


def test_kb8_epipolar_constrain_oracle_vs_reference_source():
    """KannalaBrandt8::epipolarConstrain_ (the Matx code path SearchForTriangulation_ reaches with fisheye cameras:
    TriangulateMatches_ / Triangulate_, KannalaBrandt8.cpp:409-480, 514-537) vs the oracle's per-pair predicate, on the matched
    pairs of a fisheye fixture plus mismatched pairs: same decision wherever no threshold is within 1e-3, same point."""
    c = mc.fisheye_inputs("fisheye_tumvi")
    s2 = tables()[2]
    l2r = ob.fisheye_stereo(c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"], c["mono_r"], c["R12"], c["t12"], s2)[0]
    rng = np.random.default_rng(8)
    iL = np.concatenate([np.flatnonzero(l2r >= 0), rng.integers(c["mono_l"], len(c["kL"]), 1500)])
    iR = np.concatenate([l2r[l2r >= 0], rng.integers(c["mono_r"], len(c["kR"]), 1500)])
    k1, k2 = c["kL"][iL], c["kR"][iR]
    sig, unc = s2[k1["octave"]], s2[k2["octave"]]
    ok, depth, p3d = ob.ref_kb8_epipolar_constrain(c["cam1"], c["cam2"], k1, k2, c["R12"], c["t12"], sig, unc)
    odepth, op3d, mg = ob.kb8_triangulate_pairs(c["cam1"], c["cam2"], k1, k2, c["R12"], c["t12"], sig, unc)
    safe = mg > 1e-3
    assert safe.mean() > 0.95
    assert np.array_equal(ok[safe], odepth[safe] > 0.0001)
    assert np.array_equal(ok, depth > 0.0001)
    both = safe & ok
    assert both.sum() > 300 and (~ok & safe).sum() > 300
    assert np.allclose(depth[both], odepth[both], rtol=2e-3, atol=1e-5) and np.allclose(p3d[both], op3d[both], rtol=2e-3, atol=2e-4)


def _stereo_random_configs(seed, count):
    from test_reference_orbextractor import random_configs
    return [c for c in random_configs(seed, 3 * count) if c[4] >= 3][:count]


def _cfg_id(c):
    return "%dx%d_n%d_s%.2f_l%d_t%d-%d" % c


@pytest.mark.parametrize("cfg", _stereo_random_configs(20261107, 6), ids=_cfg_id)
def test_stereo_frame_constructor_random_configs(cfg):
    """Random image shapes, feature budgets, scale factors, level counts and FAST thresholds: the reference's stereo Frame
    constructor from two images == the oracle's extraction + stereo matcher + undistortion, byte for byte."""
    import orbfe
    w, h, nf, sf, nl, ini, mn = cfg
    imL, imR = orbfe.synth.stereo_pair(9000 + w + h, w, h)
    K4 = np.array([0.6 * w, 0.6 * w, w / 2 - 3.25, h / 2 + 1.5], np.float32)
    D = np.array([-0.21, 0.05, 0.0004, -0.0003], np.float32)
    bf = float(np.float32(0.11 * K4[0]))
    F = ob.ref_stereo_frame(imL, imR, (nf, sf, nl, ini, mn), K4, D, bf)
    rL, rR = ob.OrbRef(nf, sf, nl, ini, mn), ob.OrbRef(nf, sf, nl, ini, mn)
    _, kL, dL = rL(imL, (0, 0))
    _, kR, dR = rR(imR, (0, 0))
    assert F["kps"].tobytes() == kL.tobytes() and np.array_equal(F["desc"], dL)
    assert F["kps_r"].tobytes() == kR.tobytes() and np.array_equal(F["desc_r"], dR)
    b = float(np.float32(np.float32(bf) / K4[0]))
    u, d, _ = ob.stereo_match(rL, rR, kL, dL, kR, dR, bf, b)
    assert np.array_equal(F["u_right"], u) and np.array_equal(F["depth"], d)
    assert F["kps_un"].tobytes() == ob.undistort_keypoints(kL, K4, D).tobytes()
    assert (d > 0).sum() > 0.2 * len(kL)
