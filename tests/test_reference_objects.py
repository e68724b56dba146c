"""CPU: the reference's matcher on the reference's OWN object model.  oracle/_ref/liborbslam3_objects.so compiles src/ORBmatcher.cc,
KeyFrame.cc, Frame.cc, MapPoint.cc, ORBextractor.cc and both camera models together, each against its own header; the wrappers fill
reference Frames, make KeyFrames with the reference's constructor (feature copies, SetPose, mTlr) and real MapPoints, and call the
matcher.  No data-holder stand-in for Frame / KeyFrame / MapPoint is involved (compare tests/test_reference_orbmatcher.py, where
oracle/refshim/orbslam3_standin.h supplies those classes)."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
from oracle_binding import OrbRef
from test_reference_matcher_cameras import _sides

pytestmark = pytest.mark.skipif(not ob.ref_objects_available(), reason="oracle/_ref not built (needs /root/reference; make -C oracle ref)")


@pytest.mark.parametrize("name", list(mc.TRI_KB8_CASES) + list(mc.TRI_KB8_EXTRA))
def test_triangulation_search_on_real_keyframes_kannala_brandt(name):
    """The keyframe's own pose algebra (KeyFrame::SetPose, GetRightRotation_ / GetRightTranslation_ out of mTlr) feeds the
    reference's loop and KannalaBrandt8::epipolarConstrain_; rows outside the 1e-3 threshold margin must equal the oracle's."""
    sc, isc, s2 = OrbRef(1000, *mc.ORB).tables()[:3]
    c = mc.tri_kb8_inputs(name)
    A, B = _sides(c, s2, sc, c["cam_a"], c["cam_b"])
    m12, n = ob.ref_objects_search_for_triangulation(1, A, B, 0.6, c["check_ori"], False, c["coarse"])
    om12, on, _, mg = ob.search_triangulation_kb8(c["k1"], c["d1"], c["fv1"], c["mp1"], s2, c["k2"], c["d2"], c["fv2"], c["mp2"], s2, sc,
                                                  c["cam_a"], c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"],
                                                  c["check_ori"], False, c["coarse"])
    assert n == (m12 >= 0).sum() and n > 60
    safe, differs = mg > 1e-3, m12 != om12
    if not (c["check_ori"] and (differs & ~safe).any()):
        assert not (differs & safe).any(), np.flatnonzero(differs & safe)
        assert n == on or (differs & ~safe).any()
    else:
        assert differs.mean() < 0.01
    # and the stand-in keyframe of liborbslam3_orbmatcher_cam.so gives what the real one gives
    if ob.ref_matcher_cam_available():
        sm12, sn = ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], False, c["coarse"])
        assert sn == n or (differs & ~safe).any() or (mg <= 1e-3).any()
        assert (sm12 == m12)[safe].all()


@pytest.mark.parametrize("name", [k for k, v in mc.TRI_CASES.items() if v[7] is None])
def test_triangulation_search_on_real_keyframes_pinhole(name):
    from test_match_oracle import extract, tables
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    cam = mc.TRI_CASES[name][0]
    K = np.zeros((2, 8), np.float32)
    K[:, :4] = (cam["fx"], cam["fy"], cam["cx"], cam["cy"])
    c["Tcw1"], c["Tcw2"], c["Trl"] = np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32)
    c["Tcw2"][0, 3] = -0.1
    A, B = _sides(c, s2, sc, K, K, c["ur1"], c["ur2"])
    m12, n = ob.ref_objects_search_for_triangulation(0, A, B, 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    om12, on = ob.search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                                       s2, sc, c["F12"], c["ep"], c["check_ori"], c["only_stereo"], c["coarse"])
    assert n == on and np.array_equal(m12, om12) and n > 20


@pytest.mark.parametrize("name", [k for k, v in mc.BOW_CASES.items() if v[5] == -1])
def test_search_by_bow_between_real_keyframes(name):
    """SearchByBoW(KeyFrame*, KeyFrame*) with real MapPoints behind GetMapPointMatches(): the matched MapPoint of every pKF1 feature,
    mapped back to its pKF2 feature index, equals the oracle's."""
    from test_match_oracle import extract, tables
    sc, isc, s2, is2, per, umax = tables()
    c = mc.bow_inputs(name, extract)
    cam = np.zeros((2, 8), np.float32)
    cam[:, :4] = (500, 500, 320, 240)
    eye = np.eye(4, dtype=np.float32)
    A = dict(kps=c["k1"], desc=c["d1"], fv=c["fv1"], mp=c["mp1"], nleft=-1, u_right=None, sigma2=s2, scale=sc, cam=cam, Tcw=eye, Trl=eye)
    B = dict(kps=c["k2"], desc=c["d2"], fv=c["fv2"], mp=c["mp2"], nleft=-1, u_right=None, sigma2=s2, scale=sc, cam=cam, Tcw=eye, Trl=eye)
    a2b, n = ob.ref_objects_search_by_bow_kf_kf(A, B, 0.8, c["check_ori"])
    oa2b, _, on = ob.search_bow(c["d1"], c["fv1"], c["mp1"], c["k1"]["angle"], c["d2"], c["fv2"], c["mp2"], c["k2"]["angle"], 0.8,
                                c["check_ori"], strict=True)
    assert n == on and np.array_equal(a2b, oa2b) and n > 50


@pytest.mark.parametrize("name", list(__import__("proj_cases").MP_CASES))
def test_search_by_projection_local_points_on_real_frames(name):
    """SearchByProjection(Frame&, vpMapPoints, ...) :44 (Tracking::SearchLocalPoints): a reference Frame whose 64 x 48 grid its own
    AssignFeaturesToGrid filled, reference MapPoints carrying the tracking fields Frame::isInFrustum would have written."""
    import os
    import proj_cases as pc
    from common import GOLDEN
    from test_proj_oracle import run_mp
    c = pc.mp_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    with ob.objects_matcher_library():
        n, res = ob.ref_search_by_projection_mappoints(c["F"], c["mp"], c["th"], c["nnratio"], c["far"], c["th_far"])
    ores, on, _, _ = run_mp(c, ob.search_by_projection)
    assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["mappoint_of_feature"])


@pytest.mark.parametrize("name", list(__import__("proj_cases").LF_CASES))
def test_search_by_projection_last_frame_on_real_frames(name):
    """SearchByProjection(CurrentFrame, LastFrame, th, bMono) :2173 (Tracking::TrackWithMotionModel) on two reference Frames (their
    own SetPose / UpdatePoseMatrices, grid and GetFeaturesInArea) and reference MapPoints (SetWorldPos / GetWorldPos)."""
    import os
    import proj_cases as pc
    from common import GOLDEN
    from test_proj_oracle import run_lf
    c = pc.lf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    Trl = np.eye(4, dtype=np.float32)
    Trl[:3, 3] = (-0.101, 0.0, 0.0)
    with ob.objects_matcher_library():
        n, res = ob.ref_search_by_projection_lastframe(c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], pc.CAM, Trl, c["ori"])
    ores, on, _, _ = run_lf(c, ob.search_by_projection)
    assert n == on == int(g["n"])
    assert np.array_equal(res, np.where(ores >= 0, ores, -1))


@pytest.mark.parametrize("name", [k for k in __import__("proj_cases").KF_CASES if __import__("proj_cases").kf_inputs(k)["kind"] == "sim3"])
def test_search_by_projection_sim3_on_real_keyframes(name):
    """SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming) :473 (loop closing): a reference KeyFrame (grid copied
    by its constructor, its own IsInImage / GetFeaturesInArea) and reference MapPoints (their own PredictScale and distance bounds)."""
    import os
    import proj_cases as pc
    from common import GOLDEN
    from test_proj_oracle import run_kf
    c = pc.kf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    ores, on, _, _ = run_kf(c, ob.search_by_projection)
    with ob.objects_matcher_library():
        n, res = ob.ref_search_by_projection_sim3(c["KF"], c["pts"], c["Scw"], c["th"], c["ratio"], pc.CAM, pc.LOG_SF)
    assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["point_of_feature"])


@pytest.mark.parametrize("name", list(__import__("proj_cases").SIM3_CASES))
def test_search_by_sim3_on_real_keyframes(name):
    import os
    import proj_cases as pc
    from common import GOLDEN
    from test_proj_oracle import run_sim3
    c = pc.sim3_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    with ob.objects_matcher_library():
        n, m12 = ob.ref_search_by_sim3(c, pc.CAM, pc.LOG_SF)
    vn1, vn2, pairs, on = run_sim3(c, ob.search_by_projection)
    want = np.full(c["KF1"].n, -1, np.int32)
    for i1, i2 in pairs:
        want[i1] = i2
    assert n == on == int(g["n"]) and np.array_equal(m12, want) and n > 100


def test_search_by_projection_relocalisation_on_real_objects():
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) :2391 with a reference Frame, KeyFrame and MapPoints."""
    import proj_cases as pc
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
        with ob.objects_matcher_library():
            n, res = ob.ref_search_by_projection_reloc(c["Cur"], kf, c["Tcw"], 10.0, orbdist, pc.CAM, np.float32(np.log(1.2)))
        assign, choice, on = ob.search_by_projection(c["Cur"], q, 0, orbdist, 0.9, False, True)
        assert n == on and n > 60 and np.array_equal(res, np.where(assign >= 0, owners(assign, owner), -1))


@pytest.mark.parametrize("name", [k for k in __import__("proj_cases").KF_CASES if __import__("proj_cases").kf_inputs(k)["kind"] != "sim3"])
def test_fuse_on_real_keyframes(name):
    """Fuse(pKF, vpMapPoints, th, bRight) :1629 and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) :1823 with the reference's own map
    surgery (MapPoint::AddObservation / Replace, KeyFrame::AddMapPoint / ReplaceMapPointMatch): the feature every point ends up at."""
    import os
    import proj_cases as pc
    from common import GOLDEN
    from test_proj_oracle import run_kf
    c = pc.kf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    ores, on, _, _ = run_kf(c, ob.search_by_projection)
    sim3 = c["kind"] == "fuse_sim3"
    with ob.objects_matcher_library():
        n, res = ob.ref_fuse(c["KF"], c["pts"], None if sim3 else c["Tcw"], c["Scw"] if sim3 else None, c["th"], c["kind"] == "fuse_right",
                             pc.CAM, pc.LOG_SF)
    assert n == on == int(g["n"]) and np.array_equal(res, ores) and np.array_equal(res, g["feature_of_point"])


@pytest.mark.parametrize("name", list(mc.BOW_CASES))
def test_search_by_bow_on_real_objects(name):
    """Both SearchByBoW overloads (:269 keyframe vs frame, with the left / right split of a two-camera frame; :823 keyframe vs
    keyframe) on reference KeyFrames / Frame / MapPoints, against the oracle and the golden vectors."""
    import os
    from common import GOLDEN
    from test_match_oracle import extract
    c = mc.bow_inputs(name, extract)
    g = np.load(os.path.join(GOLDEN, f"match_{name}.npz"))
    ang1, ang2 = c["k1"]["angle"], c["k2"]["angle"]
    with ob.objects_matcher_library():
        a2b, b2a, n = ob.ref_search_bow(0, c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], None, ang2, 0.7, c["check_ori"], c["nleft"])
    oa2b, ob2a, on = ob.search_bow(c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], None, ang2, 0.7, c["check_ori"], strict=False,
                                   nleft=c["nleft"])
    assert n == on == int(g["n_kf_f"]) and np.array_equal(b2a, ob2a) and np.array_equal(b2a, g["kf_f"]) and n > 100
    with ob.objects_matcher_library():
        a2b, b2a, n = ob.ref_search_bow(1, c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], c["mp2"], ang2, 0.8, c["check_ori"])
    oa2b, ob2a, on = ob.search_bow(c["d1"], c["fv1"], c["mp1"], ang1, c["d2"], c["fv2"], c["mp2"], ang2, 0.8, c["check_ori"], strict=True)
    assert n == on == int(g["n_kf_kf"]) and np.array_equal(a2b, oa2b) and np.array_equal(a2b, g["kf_kf"])


@pytest.mark.parametrize("name", list(__import__("proj_cases").INIT_CASES))
def test_search_for_initialization_on_real_frames(name):
    import os
    import proj_cases as pc
    from common import GOLDEN
    c = pc.init_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    with ob.objects_matcher_library():
        n, m12, prev = ob.ref_search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    assert n == int(g["n"]) and np.array_equal(m12, g["vnMatches12"]) and np.array_equal(prev, g["prev_matched"])
