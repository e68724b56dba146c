"""CPU: the C++ matcher oracle (oracle/cpu_ref/match_ref.cc) against the committed golden vectors
produced by the independent Python restatement (oracle/py/match_py.py + cv2), on inputs rebuilt with
the C++ extractor oracle.  Plus small known-answer tests of the primitives."""
import os
import zlib

import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe
from common import GOLDEN
from oracle_binding import OrbRef


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c & 0xFFFFFFFF


def extract(img, nf, lap):
    ref = OrbRef(nf, *mc.ORB)
    mono, k, d = ref(img, lap)
    return mono, k, d, ref


def gold(name):
    return np.load(os.path.join(GOLDEN, f"match_{name}.npz"))


def tables():
    return OrbRef(1000, *mc.ORB).tables()


def test_descriptor_distance_swar_equals_popcount():
    d = orbfe.synth.descriptors(3, 64)
    for i in range(0, 64, 2):
        want = int(np.unpackbits(d[i] ^ d[i + 1]).sum())
        assert ob.descriptor_distance(d[i], d[i + 1]) == want
        assert orbfe.ORBmatcher.DescriptorDistance(d[i], d[i + 1]) == want
    assert ob.descriptor_distance(d[0], d[0]) == 0
    assert ob.descriptor_distance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


def test_three_maxima_rules():
    """ORBmatcher.cc:2515-2556: strict '>' keeps the first of equal bins, the 0.1*max1 cut drops weak bins."""
    h = [0] * 30
    h[3], h[7], h[9] = 50, 50, 4
    assert ob.three_maxima(h) == (3, 7, -1)          # 4 < 0.1 * 50
    h[9] = 5
    assert ob.three_maxima(h) == (3, 7, 9)
    assert ob.three_maxima([0] * 30) == (-1, -1, -1)
    h = [0] * 30
    h[12] = 9
    assert ob.three_maxima(h) == (12, -1, -1)


@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_matches_golden(name):
    c = mc.stereo_inputs(name, extract)
    (_, kL, dL, rL), (_, kR, dR, rR) = c["L"], c["R"]
    g = gold(name)
    assert crc(kL, dL, kR, dR) == int(g["in_crc"])
    u, d, sad = ob.stereo_match(rL, rR, kL, dL, kR, dR, c["bf"], c["b"])
    assert np.array_equal(sad, g["sad"])
    assert np.array_equal(u, g["u_right"]) and np.array_equal(d, g["depth"])   # float32-exact
    assert (d > 0).sum() > 0.4 * len(kL)
    # ground truth of the synthetic pair: disparity is one of three integer planes
    ok = d > 0
    disp = kL["x"][ok] - u[ok]
    true = np.array([orbfe.synth.disparity(int(y), c["cam"]["h"]) for y in kL["y"][ok]])
    assert np.mean(np.abs(disp - true) < 1.5) > 0.95


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_bow_matches_golden(name):
    c = mc.bow_inputs(name, extract)
    g = gold(name)
    assert crc(c["d1"], c["d2"], c["mp1"], c["mp2"]) == int(g["in_crc"])
    a2b, b2a, n = ob.search_bow(c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], None, c["k2"]["angle"], 0.7,
                                c["check_ori"], strict=False, nleft=c["nleft"])
    assert n == int(g["n_kf_f"]) and np.array_equal(b2a, g["kf_f"])
    assert n > 100
    a2b, b2a, n = ob.search_bow(c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], c["mp2"], c["k2"]["angle"],
                                0.8, c["check_ori"], strict=True)
    assert n == int(g["n_kf_kf"]) and np.array_equal(a2b, g["kf_kf"])
    assert n == (a2b >= 0).sum() == (b2a >= 0).sum()


@pytest.mark.parametrize("name", list(mc.TRI_CASES))
def test_triangulation_matches_golden(name):
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    g = gold(name)
    assert crc(c["d1"], c["d2"], c["mp1"], c["mp2"], c["F12"]) == int(g["in_crc"])
    m12, n = ob.search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"],
                                     c["ur2"], s2, sc, c["F12"], c["ep"], c["check_ori"], c["only_stereo"], c["coarse"])
    assert n == int(g["n"]) and np.array_equal(m12, g["m12"])
    assert n > 30


@pytest.mark.parametrize("name", list(mc.LAP_CASES))
def test_lapping_matches_golden(name):
    c = mc.lap_inputs(name, extract)
    g = gold(name)
    assert crc(c["q"], c["t"]) == int(g["in_crc"])
    idx, dist = ob.knn2(c["q"], c["t"])
    assert np.array_equal(idx, g["idx"]) and np.array_equal(dist, g["dist"])
    assert np.array_equal(ob.knn2_ratio(c["q"], c["t"], 0.7), g["match"])


def test_knn2_tie_rule_golden():
    q, t = mc.knn_inputs()
    g = gold("knn_ties")
    idx, dist = ob.knn2(q, t)
    assert np.array_equal(idx, g["idx"]) and np.array_equal(dist, g["dist"])
    assert idx[0].tolist() == [7, 100] and idx[5].tolist() == [1500, 1501]
    # fewer than two train rows: missing neighbours are -1 (Frame.cc:1141 checks size() >= 2)
    idx, dist = ob.knn2(q[:3], t[:1])
    assert idx[:, 1].tolist() == [-1, -1, -1] and (idx[:, 0] == 0).all()
    assert (ob.knn2_ratio(q[:3], t[:1]) == -1).all()


@pytest.mark.parametrize("name", list(mc.FISHEYE_CASES))
def test_fisheye_stereo_matches_golden(name):
    """KB8 triangulation ends in a 4x4 SVD (cv::SVD in the reference, cv2.SVDecomp in the golden generator, a
    double-precision Jacobi solve here): float tolerance, decisions compared away from the thresholds."""
    c = mc.fisheye_inputs(name)
    g = gold(name)
    assert crc(c["kL"], c["dL"], c["kR"], c["dR"]) == int(g["in_crc"])
    s2 = tables()[2]
    l2r, r2l, depth, p3d, mg, n = ob.fisheye_stereo(c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"], c["kR"], c["dR"],
                                                    c["mono_r"], c["R12"], c["t12"], s2)
    safe = np.minimum(mg, g["margins"]) > 1e-3          # both sides agree the decision is not knife-edge
    assert safe.sum() > 0.95 * (g["margins"] < 1e29).sum()
    assert np.array_equal(l2r[safe], g["l2r"][safe])
    ok = safe & (g["l2r"] >= 0)
    assert ok.sum() > 200
    assert np.allclose(depth[ok], g["depth"][ok], rtol=2e-3, atol=1e-5)
    assert np.allclose(p3d[ok], g["p3d"][ok], rtol=2e-3, atol=2e-4)
    assert abs(n - int(g["n"])) <= (~safe & (g["margins"] < 1e29)).sum()
    # every right index claimed at most once per surviving left match
    m = l2r[l2r >= 0]
    assert np.array_equal(r2l[m] >= 0, np.ones(len(m), bool))


def tri_kb8_call(fn, c, s2, sc, **kw):
    a = dict(check_ori=c["check_ori"], only_stereo=c["only_stereo"], coarse=c["coarse"])
    a.update(kw)
    return fn(c["k1"], c["d1"], c["fv1"], c["mp1"], s2, c["k2"], c["d2"], c["fv2"], c["mp2"], s2, sc, c["cam_a"], c["cam_b"],
              c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"], a["check_ori"], a["only_stereo"], a["coarse"])


@pytest.mark.parametrize("name", list(mc.TRI_KB8_CASES))
def test_triangulation_kb8_matches_golden(name):
    """SearchForTriangulation_ with KB8 cameras: the accept test is a triangulation (4x4 SVD), so rows whose
    candidates all sit away from the thresholds must agree exactly, points to float tolerance."""
    sc, isc, s2 = tables()[:3]
    c = mc.tri_kb8_inputs(name)
    g = gold(name)
    assert crc(c["k1"], c["d1"], c["k2"], c["d2"], c["R12"]) == int(g["in_crc"])
    m12, n, pts, mg = tri_kb8_call(ob.search_triangulation_kb8, c, s2, sc)
    if not c["check_ori"]:
        safe = np.minimum(mg, g["margins"]) > 1e-3
        assert np.array_equal(m12[safe], g["m12"][safe])
        assert abs(n - int(g["n"])) <= (~safe).sum()
    elif (np.minimum(mg, g["margins"]) > 1e-3).all():
        assert np.array_equal(m12, g["m12"]) and n == int(g["n"])     # the rotation histogram couples all rows
    else:
        assert np.mean(m12 == g["m12"]) > 0.99
    assert n > 60
    ok = (m12 >= 0) & (m12 == g["m12"]) & (not c["coarse"])
    assert np.allclose(pts[ok], g["points"][ok], rtol=2e-3, atol=2e-4)
    if c["nleft_a"] >= 0:                                             # all four camera pairings occur
        r1, r2 = np.nonzero(m12 >= 0)[0] >= c["nleft_a"], m12[m12 >= 0] >= c["nleft_b"]
        assert len(set(zip(r1.tolist(), r2.tolist()))) == 4
    # bOnlyStereo: a fisheye keyframe has no mvuRight >= 0 (ORBmatcher.cc:1280-1283), nothing matches
    assert tri_kb8_call(ob.search_triangulation_kb8, c, s2, sc, only_stereo=True)[1] == 0


def test_kb8_project_unproject_roundtrip():
    cam = mc.KB8_LEFT
    L = ob.lib()
    rng = np.random.default_rng(1)
    for _ in range(200):
        u, v = rng.uniform(80, 430, 2)      # inside the 90-degree cone (unproject clamps theta_d at pi/2)
        ray = np.zeros(3, np.float32)
        L.orbref_kb8_unproject(ob._p(cam), float(u), float(v), ob._p(ray))
        uv = np.zeros(2, np.float32)
        L.orbref_kb8_project(ob._p(cam), ob._p(ray), ob._p(uv))
        assert abs(uv[0] - u) < 2e-3 and abs(uv[1] - v) < 2e-3


def test_distinctive_descriptors_golden():
    d, off = mc.distinctive_inputs()
    g = gold("distinctive")
    assert crc(d, off) == int(g["in_crc"])
    best = ob.distinctive_descriptors(d, off)
    assert np.array_equal(best, g["best"])
    assert best[0] == -1 and best[1] == 0 and (np.diff(off)[best < 0] == 0).all()
