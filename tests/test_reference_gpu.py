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
