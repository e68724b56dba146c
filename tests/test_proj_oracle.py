"""CPU: the flattened window-search engine of the C++ oracle (oracle/cpu_ref/proj_ref.cc) fed by the host-side
flattening (orbfe/projection.py) against the golden results of the OBJECT-level Python restatement of
SearchByProjection (oracle/py/proj_py.py), plus GetFeaturesInArea known answers."""
import os
import zlib

import numpy as np
import pytest

import oracle_binding as ob
import proj_cases as pc
from common import GOLDEN
from orbfe import projection as pj


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c & 0xFFFFFFFF


def owners(assign, owner):
    out = np.full(len(assign), -1, np.int32)
    ok = assign >= 0
    out[ok] = owner[assign[ok]]
    return out


def run_mp(c, search):
    q, owner = pj.queries_from_mappoints(c["F"], c["mp"], c["th"], c["far"], c["th_far"])
    assign, choice, n = search(c["F"], q, 1, 100, c["nnratio"], True, False)
    return owners(assign, owner), n, q, choice


def run_lf(c, search):
    q, owner = pj.queries_from_lastframe(c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], c["project"], c["project_right"])
    assign, choice, n = search(c["Cur"], q, 0, 100, 0.9, False, c["ori"])
    return owners(assign, owner), n, q, choice


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_local_points_matches_golden(name):
    c = pc.mp_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["F"].kps, c["F"].desc, c["mp"]["desc"], c["mp"]["proj_x"]) == int(g["in_crc"])
    res, n, q, choice = run_mp(c, ob.search_by_projection)
    assert n == int(g["n"]) and np.array_equal(res, g["mappoint_of_feature"])
    assert n > 300 and (q["right"] > 0).any() == (c["F"].nleft != -1)
    # MapPoints with observations hold their feature exclusively; those without (obs == 0) can be overwritten
    held = choice[(choice >= 0) & (q["obs"] > 0)]
    assert len(set(held.tolist())) == len(held)
    assert len(set(choice[choice >= 0].tolist())) < (choice >= 0).sum()


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_track_last_frame_matches_golden(name):
    c = pc.lf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["Cur"].kps, c["Cur"].desc, c["last"]["desc"], c["last"]["world_pos"]) == int(g["in_crc"])
    res, n, q, choice = run_lf(c, ob.search_by_projection)
    assert n == int(g["n"]) and np.array_equal(res, g["lastframe_feature_of_feature"])
    assert n > 200


def test_features_in_area_known_answers():
    import orbfe
    k = np.zeros(6, orbfe.KEYPOINT_DTYPE)
    k["x"] = [100, 104, 96, 100, 400, 751.9]
    k["y"] = [100, 100, 100, 109.9, 300, 479.9]
    k["octave"] = [0, 1, 2, 0, 0, 0]
    F = pj.GridFrame(k, np.zeros((6, 32), np.uint8), (0, 0, 752, 480), pc.SCALE)
    assert ob.features_in_area(F, 100, 100, 5, -1, -1, False).tolist() == [2, 0, 1]       # cell order: x cell first
    assert ob.features_in_area(F, 100, 100, 4, -1, -1, False).tolist() == [0]             # |dx| < r is strict
    assert ob.features_in_area(F, 100, 100, 10, 0, 0, False).tolist() == [0, 3]
    assert ob.features_in_area(F, 100, 100, 10, 1, -1, False).tolist() == [2, 1]          # minLevel only
    assert ob.features_in_area(F, 100, 100, 10, -1, 1, False).tolist() == [0, 1, 3]       # cell (9,10) before (9,11)
    assert ob.features_in_area(F, 2000, 100, 10, -1, -1, False).tolist() == []
    # the last feature rounds to grid cell (64, 48): PosInGrid drops it (Frame.cc:714)
    assert ob.features_in_area(F, 751, 479, 5, -1, -1, False).tolist() == []


# ---- KeyFrame-side searches ----------------------------------------------------------------------------------------
def kf_windows(c):
    KF, pts, kind = c["KF"], c["pts"], c["kind"]
    if kind in ("sim3", "fuse_sim3"):
        q, owner = pj.queries_from_sim3_points(KF, pts, c["Scw"], c["th"], c["project"], pc.LOG_SF, claim=kind == "sim3")
        th = pj.hamming_threshold(50, c["ratio"]) if kind == "sim3" else 50
        return q, owner, th
    T = c["Tcw"]
    q, owner = pj.queries_for_fuse(KF, pts, T[:3, :3], T[:3, 3], pj.camera_centre(T[:3, :3], T[:3, 3]), c["th"], c["project"],
                                   pc.LOG_SF, kind == "fuse_right")
    return q, owner, 50


def run_kf(c, search):
    """Returns (what the golden stores, n, queries, choice)."""
    q, owner, th = kf_windows(c)
    assign, choice, n = search(c["KF"], q, 0, th, 0.9, False, False)
    if c["kind"] == "sim3":
        return owners(assign, owner), n, q, choice
    out = np.full(len(c["pts"]["desc"]), -1, np.int32)
    out[owner] = choice
    return out, n, q, choice


@pytest.mark.parametrize("name", list(pc.KF_CASES))
def test_keyframe_side_searches_match_golden(name):
    c = pc.kf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["KF"].kps, c["KF"].desc, c["pts"]["desc"], c["pts"]["world_pos"]) == int(g["in_crc"])
    res, n, q, choice = run_kf(c, ob.search_by_projection)
    key = "point_of_feature" if c["kind"] == "sim3" else "feature_of_point"
    assert n == int(g["n"]) and np.array_equal(res, g[key]) and n > 200
    if c["kind"] == "sim3":                  # claiming windows: a feature is taken once, never a pre-matched one
        taken = choice[choice >= 0]
        assert len(set(taken.tolist())) == len(taken) and not c["KF"].blocked[taken].any()
    else:                                    # Fuse never claims: several points may pick the same feature
        taken = choice[choice >= 0]
        assert len(set(taken.tolist())) < len(taken)
    if c["kind"] == "fuse_right":
        assert (choice[choice >= 0] >= c["KF"].nleft).all()


def test_fuse_chi_square_gate_matters():
    """Without the gate of ORBmatcher.cc:1744-1770 the same windows pick more (and different) features."""
    c = pc.kf_inputs("proj_kf_fuse")
    q, owner, th = kf_windows(c)
    _, with_gate, n1 = ob.search_by_projection(c["KF"], q, 0, th, 0.9, False, False)
    q2 = q.copy()
    q2["chi2"] = 0
    _, without, n2 = ob.search_by_projection(c["KF"], q2, 0, th, 0.9, False, False)
    assert n2 > n1 and (with_gate != without).sum() > 50
    # stereo features (mvuRight >= 0) use the 3-dof gate: moving `ur` far away rejects exactly those
    q3 = q.copy()
    q3["ur"] += 1000
    _, far, n3 = ob.search_by_projection(c["KF"], q3, 0, th, 0.9, False, False)
    picked = far[far >= 0]
    assert n3 < n1 and (c["KF"].u_right[picked] < 0).all()


def run_sim3(c, search):
    f = lambda F, q, mode, th: search(F, q, mode, th, 0.9, False, False)
    return pj.search_by_sim3(f, c["KF1"], c["KF2"], c["mps1"], c["mps2"], c["T1w"][:3], c["T2w"][:3], c["s12"], c["R12"], c["t12"],
                             c["th"], pc.CAM, pc.LOG_SF)


@pytest.mark.parametrize("name", list(pc.SIM3_CASES))
def test_search_by_sim3_matches_golden(name):
    c = pc.sim3_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["KF1"].kps, c["KF2"].kps, c["mps1"]["desc"], c["mps2"]["world_pos"]) == int(g["in_crc"])
    vn1, vn2, pairs, n = run_sim3(c, ob.search_by_projection)
    assert np.array_equal(vn1, g["vnMatch1"]) and np.array_equal(vn2, g["vnMatch2"]) and n == int(g["n"]) and n > 100
    assert n < (vn1 >= 0).sum()              # the agreement pass rejects some one-way matches


@pytest.mark.parametrize("name", list(pc.INIT_CASES))
def test_search_for_initialization_matches_golden(name):
    """oracle/cpu_ref searchForInitialization vs the object-level Python restatement of ORBmatcher.cc:706-821."""
    c = pc.init_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["F1"].kps, c["F1"].desc, c["F2"].kps, c["F2"].desc, c["prev"]) == int(g["in_crc"])
    n, m12, prev = ob.search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    assert n == int(g["n"]) and np.array_equal(m12, g["vnMatches12"]) and np.array_equal(prev, g["prev_matched"])
    assert n > 100 and n == (m12 >= 0).sum()
    # one-to-one, level 0 on both sides
    hit = m12[m12 >= 0]
    assert len(set(hit.tolist())) == len(hit)
    assert (c["F1"].kps["octave"][m12 >= 0] == 0).all() and (c["F2"].kps["octave"][hit] == 0).all()
