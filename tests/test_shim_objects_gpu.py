"""The object layer of the ORBmatcher shim (include/orbfe_shim/ORBmatcher.h + ORBmatcherObjects.h): every Search* / Fuse
member with the reference's signature, driven through the SAME C wrappers and the same stand-in Frame / KeyFrame /
MapPoint objects as the reference's own ORBmatcher.cc (oracle/refshim/ref_matcher_capi.cc compiled twice: against the
reference's ORBmatcher.h + ORBmatcher.cc -> oracle/_ref/liborbslam3_orbmatcher.so, and against the shim header ->
tests/cpp/_build/liborbfe_shim_orbmatcher.so, which calls liborbfe.so on the GPU).  Results must be identical."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import proj_cases as pc
from test_match_oracle import extract, tables

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (ob.ref_matcher_available() and ob.shim_matcher_available()),
                                                  reason="oracle/_ref or the shim matcher library not built")]


def both(fn, *a, **k):
    want = fn(*a, **k)
    with ob.shim_matcher_library():
        got = fn(*a, **k)
    return want, got


def same(want, got):
    assert len(want) == len(got)
    for w, g in zip(want, got):
        assert np.array_equal(np.asarray(w), np.asarray(g))


def test_three_maxima_and_descriptor_distance():
    rng = np.random.default_rng(7)
    for _ in range(200):
        counts = rng.integers(0, 6, 30) * rng.integers(0, 2, 30)
        w, g = both(ob.ref_three_maxima, counts)
        assert w == g
    a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
    w, g = both(ob.ref_descriptor_distance, a, b)
    assert w == g == int(np.unpackbits(a ^ b).sum())


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_search_by_bow_objects(name):
    c = mc.bow_inputs(name, extract)
    w, g = both(ob.ref_search_bow, 0, c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], None, c["k2"]["angle"], 0.7,
                c["check_ori"], c["nleft"])
    same(w, g)
    assert w[2] > 100
    w, g = both(ob.ref_search_bow, 1, c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], c["mp2"], c["k2"]["angle"], 0.8,
                c["check_ori"])
    same(w, g)


@pytest.mark.parametrize("name", list(mc.TRI_CASES))
def test_search_for_triangulation_objects(name):
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    w, g = both(ob.ref_search_triangulation, c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                s2, sc, c["F12"], c["ep"], 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    same(w, g)


@pytest.mark.parametrize("name", list(pc.INIT_CASES))
def test_search_for_initialization_objects(name):
    c = pc.init_inputs(name)
    w, g = both(ob.ref_search_for_initialization, c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    same(w, g)
    assert w[0] > 100


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_local_points_objects(name):
    c = pc.mp_inputs(name)
    w, g = both(ob.ref_search_by_projection_mappoints, c["F"], c["mp"], c["th"], c["nnratio"], c["far"], c["th_far"])
    same(w, g)
    assert w[0] > 300


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_track_last_frame_objects(name):
    c = pc.lf_inputs(name)
    Trl = np.eye(4, dtype=np.float32)
    Trl[:3, 3] = (-0.101, 0.0, 0.0)
    w, g = both(ob.ref_search_by_projection_lastframe, c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], pc.CAM, Trl, c["ori"])
    same(w, g)
    assert w[0] > 200


@pytest.mark.parametrize("name", list(pc.KF_CASES))
def test_keyframe_side_objects(name):
    c = pc.kf_inputs(name)
    if c["kind"] == "sim3":
        w, g = both(ob.ref_search_by_projection_sim3, c["KF"], c["pts"], c["Scw"], c["th"], c["ratio"], pc.CAM, pc.LOG_SF)
    else:
        sim3 = c["kind"] == "fuse_sim3"
        w, g = both(ob.ref_fuse, c["KF"], c["pts"], None if sim3 else c["Tcw"], c["Scw"] if sim3 else None, c["th"], c["kind"] == "fuse_right",
                    pc.CAM, pc.LOG_SF)
    same(w, g)
    assert w[0] > 200


@pytest.mark.parametrize("name", list(pc.SIM3_CASES))
def test_search_by_sim3_objects(name):
    c = pc.sim3_inputs(name)
    w, g = both(ob.ref_search_by_sim3, c, pc.CAM, pc.LOG_SF)
    same(w, g)
    assert w[0] > 100


@pytest.mark.parametrize("orbdist", [100, 64])
def test_relocalisation_objects(orbdist):
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist): the scenario of
    tests/test_reference_orbmatcher.py::test_search_by_projection_relocalisation."""
    c = pc.lf_inputs("proj_last_mono")
    rng = np.random.default_rng(77)
    last, m = c["last"], len(c["last"]["has_mp"])
    dist = np.sqrt((last["world_pos"].astype(np.float64) ** 2).sum(1)).astype(np.float32)
    kf = dict(has_mp=last["has_mp"], bad=rng.random(m) < 0.03, already_found=rng.random(m) < 0.2, world_pos=last["world_pos"],
              min_dist=(dist * rng.uniform(0.3, 1.05, m)).astype(np.float32), max_dist=(dist * rng.uniform(0.95, 4.0, m)).astype(np.float32),
              angle=last["angle"], desc=last["desc"])
    w, g = both(ob.ref_search_by_projection_reloc, c["Cur"], kf, c["Tcw"], 10.0, orbdist, pc.CAM, np.float32(np.log(1.2)))
    same(w, g)
    assert w[0] > 60


# ---- the same object layer on the reference's OWN camera classes (liborbslam3_orbmatcher_cam.so vs liborbfe_shim_orbmatcher_cam.so) ----
cam_libs = pytest.mark.skipif(not (ob.ref_matcher_cam_available() and ob.shim_matcher_cam_available()),
                              reason="the *_cam libraries are not built (make -C oracle ref shim)")


@cam_libs
@pytest.mark.parametrize("name", list(mc.TRI_KB8_CASES) + list(mc.TRI_KB8_EXTRA))
def test_search_for_triangulation_objects_kannala_brandt(name):
    """SearchForTriangulation_ on keyframes that hold real KannalaBrandt8 objects: the shim reads GetType() / getParameter(), forms the
    relative poses of the four camera pairings from the keyframes' poses and runs orbfe_search_triangulation_kb8; the reference
    runs KannalaBrandt8::epipolarConstrain_ per pair.  Rows may differ only where a threshold of the triangulation test lies within
    1e-3 (relative) of the computed value (the oracle reports that margin per row)."""
    from oracle_binding import OrbRef
    from test_reference_matcher_cameras import _sides
    sc, isc, s2 = OrbRef(1000, *mc.ORB).tables()[:3]
    c = mc.tri_kb8_inputs(name)
    A, B = _sides(c, s2, sc, c["cam_a"], c["cam_b"])
    want, wn = ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], False, c["coarse"])
    got, gn = ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], False, c["coarse"], shim=True)
    mg = ob.search_triangulation_kb8(c["k1"], c["d1"], c["fv1"], c["mp1"], s2, c["k2"], c["d2"], c["fv2"], c["mp2"], s2, sc, c["cam_a"],
                                     c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"], c["check_ori"], False,
                                     c["coarse"])[3]
    assert gn == (got >= 0).sum() and gn > 60
    safe, differs = mg > 1e-3, got != want
    if not (c["check_ori"] and (differs & ~safe).any()):
        assert not (differs & safe).any(), np.flatnonzero(differs & safe)
        assert gn == wn or (differs & ~safe).any()
    else:
        assert differs.mean() < 0.01
    assert ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], True, c["coarse"], shim=True)[1] == 0


@cam_libs
@pytest.mark.parametrize("name", [k for k, v in mc.TRI_CASES.items() if v[7] is None])
def test_search_for_triangulation_objects_real_pinhole(name):
    """Keyframes that hold real Pinhole objects, F12 as LocalMapping::ComputeF12_ forms it from the poses: identical pairs."""
    from test_reference_matcher_cameras import _sides
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    cam = mc.TRI_CASES[name][0]
    K = np.zeros((2, 8), np.float32)
    K[:, :4] = (cam["fx"], cam["fy"], cam["cx"], cam["cy"])
    c["Tcw1"], c["Tcw2"], c["Trl"] = np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32)
    c["Tcw2"][0, 3] = -0.1
    A, B = _sides(c, s2, sc, K, K, c["ur1"], c["ur2"])
    want, wn = ob.ref_search_for_triangulation_cam(0, A, B, 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    got, gn = ob.ref_search_for_triangulation_cam(0, A, B, 0.6, c["check_ori"], c["only_stereo"], c["coarse"], shim=True)
    assert wn == gn and np.array_equal(want, got) and gn > 20
