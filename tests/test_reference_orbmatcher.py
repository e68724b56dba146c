"""CPU: the reference's OWN src/ORBmatcher.cc + include/ORBmatcher.h, compiled unmodified into
oracle/_ref/liborbslam3_orbmatcher.so (the include guards of Frame.h / KeyFrame.h / MapPoint.h are pre-defined and the
data-holder stand-ins of oracle/refshim/orbslam3_standin.h force-included), against the C++ oracle and the golden
vectors of the Python restatements: DescriptorDistance, ComputeThreeMaxima, both SearchByBoW overloads,
SearchForInitialization and SearchByProjection(Frame&, vpMapPoints, ...)."""
import os

import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe
import proj_cases as pc
from common import GOLDEN
from test_match_oracle import extract

pytestmark = pytest.mark.skipif(not ob.ref_matcher_available(), reason="oracle/_ref not built (needs /root/reference; make -C oracle ref)")


def test_descriptor_distance_and_three_maxima():
    d = orbfe.synth.descriptors(11, 200)
    for i in range(0, 200, 2):
        assert ob.ref_descriptor_distance(d[i], d[i + 1]) == ob.descriptor_distance(d[i], d[i + 1]) == int(np.unpackbits(d[i] ^ d[i + 1]).sum())
    rng = np.random.default_rng(5)
    for _ in range(300):
        h = rng.integers(0, 12, 30) * (rng.random(30) < rng.random())
        if rng.random() < 0.3:
            h[rng.integers(0, 30, 3)] = h.max()                     # ties between the maxima
        assert ob.ref_three_maxima(h) == ob.three_maxima(h)


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_search_by_bow(name):
    c = mc.bow_inputs(name, extract)
    g = np.load(os.path.join(GOLDEN, f"match_{name}.npz"))
    ang1, ang2 = c["k1"]["angle"], c["k2"]["angle"]
    # SearchByBoW(KeyFrame*, Frame&, ...)
    a2b, b2a, n = ob.ref_search_bow(0, c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], None, ang2, 0.7, c["check_ori"], c["nleft"])
    oa2b, ob2a, on = ob.search_bow(c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], None, ang2, 0.7, c["check_ori"], strict=False,
                                   nleft=c["nleft"])
    assert n == on == int(g["n_kf_f"]) and np.array_equal(b2a, ob2a) and np.array_equal(b2a, g["kf_f"]) and n > 100
    # SearchByBoW(KeyFrame*, KeyFrame*, ...)
    a2b, b2a, n = ob.ref_search_bow(1, c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], c["mp2"], ang2, 0.8, c["check_ori"])
    oa2b, ob2a, on = ob.search_bow(c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], c["mp2"], ang2, 0.8, c["check_ori"], strict=True)
    assert n == on == int(g["n_kf_kf"]) and np.array_equal(a2b, oa2b) and np.array_equal(a2b, g["kf_kf"])


@pytest.mark.parametrize("name", list(pc.INIT_CASES))
def test_search_for_initialization(name):
    c = pc.init_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    n, m12, prev = ob.ref_search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    on, om12, oprev = ob.search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    assert n == on == int(g["n"]) and np.array_equal(m12, om12) and np.array_equal(m12, g["vnMatches12"])
    assert np.array_equal(prev, oprev) and np.array_equal(prev, g["prev_matched"])


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_by_projection_local_points(name):
    from test_proj_oracle import run_mp
    c = pc.mp_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    n, res = ob.ref_search_by_projection_mappoints(c["F"], c["mp"], c["th"], c["nnratio"], c["far"], c["th_far"])
    ores, on, _, _ = run_mp(c, ob.search_by_projection)
    assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["mappoint_of_feature"])


@pytest.mark.parametrize("name", list(mc.TRI_CASES))
def test_search_for_triangulation(name):
    from test_match_oracle import tables
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    g = np.load(os.path.join(GOLDEN, f"match_{name}.npz"))
    m12, n = ob.ref_search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                                         s2, sc, c["F12"], c["ep"], 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    om12, on = ob.search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                                       s2, sc, c["F12"], c["ep"], c["check_ori"], c["only_stereo"], c["coarse"])
    assert n == on and np.array_equal(m12, om12)
    assert n == int(g["n"]) and np.array_equal(m12, g["m12"])


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_search_by_projection_last_frame(name):
    """SearchByProjection(CurrentFrame, LastFrame, th, bMono) :2173: the reference projects with its own cv::Mat lines."""
    from test_proj_oracle import run_lf
    c = pc.lf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    Trl = np.eye(4, dtype=np.float32)
    Trl[:3, 3] = (-0.101, 0.0, 0.0)
    n, res = ob.ref_search_by_projection_lastframe(c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], pc.CAM, Trl, c["ori"])
    ores, on, _, _ = run_lf(c, ob.search_by_projection)
    assert n == on == int(g["n"])
    assert np.array_equal(res, np.where(ores >= 0, ores, -1)) and np.array_equal(res, np.where(g["lastframe_feature_of_feature"] >= 0, g["lastframe_feature_of_feature"], -1))


@pytest.mark.parametrize("name", list(pc.KF_CASES))
def test_keyframe_side_searches(name):
    """SearchByProjection(KeyFrame*, Scw, ...) :473, Fuse :1629 (incl. bRight) and Fuse(Scw) :1823 of the reference itself: the
    Sim3 decomposition, projection, distance / viewing-angle gates, PredictScale and the chi-square gate are its own lines."""
    from test_proj_oracle import run_kf
    c = pc.kf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    ores, on, _, _ = run_kf(c, ob.search_by_projection)
    if c["kind"] == "sim3":
        n, res = ob.ref_search_by_projection_sim3(c["KF"], c["pts"], c["Scw"], c["th"], c["ratio"], pc.CAM, pc.LOG_SF)
        assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["point_of_feature"])
    else:
        sim3 = c["kind"] == "fuse_sim3"
        n, res = ob.ref_fuse(c["KF"], c["pts"], None if sim3 else c["Tcw"], c["Scw"] if sim3 else None, c["th"], c["kind"] == "fuse_right",
                             pc.CAM, pc.LOG_SF)
        assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["feature_of_point"])


@pytest.mark.parametrize("name", list(pc.SIM3_CASES))
def test_search_by_sim3(name):
    from test_proj_oracle import run_sim3
    c = pc.sim3_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    n, m12 = ob.ref_search_by_sim3(c, pc.CAM, pc.LOG_SF)
    vn1, vn2, pairs, on = run_sim3(c, ob.search_by_projection)
    want = np.full(c["KF1"].n, -1, np.int32)
    for i1, i2 in pairs:
        want[i1] = i2
    assert n == on == int(g["n"]) and np.array_equal(m12, want) and n > 100


def test_search_by_projection_relocalisation():
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) :2391 of the reference itself vs the engine on the
    windows of projection.queries_from_keyframe (same scenario as tests/test_proj_gpu.py::test_relocalisation_overload)."""
    from orbfe import projection as pj
    from test_proj_oracle import owners
    c = pc.lf_inputs("proj_last_mono")
    rng = np.random.default_rng(77)
    last, m = c["last"], len(c["last"]["has_mp"])
    dist = np.sqrt((last["world_pos"].astype(np.float64) ** 2).sum(1)).astype(np.float32)
    kf = dict(has_mp=last["has_mp"], bad=rng.random(m) < 0.03, already_found=rng.random(m) < 0.2, world_pos=last["world_pos"],
              min_dist=(dist * rng.uniform(0.3, 1.05, m)).astype(np.float32), max_dist=(dist * rng.uniform(0.95, 4.0, m)).astype(np.float32),
              angle=last["angle"], desc=last["desc"])
    q, owner = pj.queries_from_keyframe(c["Cur"], kf, c["Tcw"], 10.0, c["project"], np.float32(np.log(1.2)))
    for orbdist in (100, 64):
        n, res = ob.ref_search_by_projection_reloc(c["Cur"], kf, c["Tcw"], 10.0, orbdist, pc.CAM, np.float32(np.log(1.2)))
        assign, choice, on = ob.search_by_projection(c["Cur"], q, 0, orbdist, 0.9, False, True)
        assert n == on and n > 60 and np.array_equal(res, np.where(assign >= 0, owners(assign, owner), -1))
