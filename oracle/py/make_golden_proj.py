#!/usr/bin/env python
"""Generates tests/golden/proj_*.npz with the OBJECT-level Python restatement of the projection-guided searches
(oracle/py/proj_py.py) on the scenarios of tests/proj_cases.py.  TEST INFRASTRUCTURE.
Run: python oracle/py/make_golden_proj.py.  A fixture stores, per Frame feature, which MapPoint (index into the
input list) ends up in mvpMapPoints (-1 untouched, -2 NULL-ed), the return value and a CRC of the inputs."""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [HERE, os.path.join(ROOT, "orb-slam3-noted_b200"), os.path.join(ROOT, "tests")]

import proj_cases as pc  # noqa: E402
import proj_py as pp  # noqa: E402
from orbfe import projection as pj  # noqa: E402


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return np.uint32(c & 0xFFFFFFFF)


class Old:
    """Stands for whatever MapPoint already sits in mvpMapPoints[i] with Observations() > 0."""
    def Observations(self):
        return 1


def frame_of(G):
    F = pp.Frame(G.kps, G.desc, G.scale_factors, (G.min_x, G.min_y, G.max_x, G.max_y), G.nleft, G.u_right, G.left_to_right,
                 G.right_to_left, G.mbf, G.mb)
    old = Old()
    for i in range(G.n):
        if G.blocked is not None and G.blocked[i]:
            F.mvpMapPoints[i] = old
    return F, old


def result_of(F, old, mps):
    idx = {id(m): i for i, m in enumerate(mps)}
    out = np.full(F.N, -1, np.int32)
    for i, m in enumerate(F.mvpMapPoints):
        if m is old:
            continue
        out[i] = idx[id(m)] if m is not None else -1
    return out


def kf_of(G):
    K = pp.KeyFrame(G.kps, G.desc, G.scale_factors, (G.min_x, G.min_y, G.max_x, G.max_y), G.nleft, G.u_right, G.left_to_right,
                    G.right_to_left, G.mbf, G.mb, inv_level_sigma2=G.inv_level_sigma2, log_scale_factor=pc.LOG_SF)
    return K


def kf_points(pts, K=None):
    out = []
    for i in range(len(pts["desc"])):
        if "null" in pts and pts["null"][i]:
            out.append(None)
            continue
        m = pp.MapPointKF(pts["desc"][i], pts["world_pos"][i], pts["normal"][i], pts["min_dist"][i], pts["max_dist"][i], 1, bool(pts["bad"][i]))
        if K is not None and "in_keyframe" in pts and pts["in_keyframe"][i]:
            m.in_kf.add(id(K))
        out.append(m)
    return out


def golden_kf(name):
    """Returns the arrays stored for one KeyFrame-side case (also used by nothing else: the tests read the file)."""
    c = pc.kf_inputs(name)
    K, kind, pts = kf_of(c["KF"]), c["kind"], c["pts"]
    mps = kf_points(pts, K)
    if kind in ("sim3", "fuse_sim3"):
        Rcw, tcw, Ow = pj.decompose_sim3(c["Scw"])
        T = np.concatenate([Rcw, tcw[:, None]], 1)
    else:
        T = c["Tcw"][:3]
        Ow = pj.camera_centre(T[:, :3], T[:, 3])
    to_cam = lambda p: pj.transform_point(T, p)
    if kind == "sim3":
        old = Old()
        vpMatched = [old if c["KF"].blocked[i] else None for i in range(K.N)]
        found = [m for i, m in enumerate(mps) if pts["already_found"][i]]
        vpMatched = _Matched(vpMatched, found)
        n = pp.search_by_projection_sim3(K, mps, vpMatched, c["th"], c["ratio"], to_cam, c["project"], Ow)
        idx = {id(m): i for i, m in enumerate(mps)}
        res = np.array([idx[id(m)] if (m is not None and m is not old) else -1 for m in list.__iter__(vpMatched)], np.int32)
        return dict(n=n, point_of_feature=res)
    if kind == "fuse_sim3":
        already = {id(m) for i, m in enumerate(mps) if pts["already_found"][i]}
        n, picks = pp.fuse_sim3(K, mps, c["th"], to_cam, c["project"], Ow, already)
    else:
        n, picks = pp.fuse(K, mps, c["th"], to_cam, c["project"], Ow, kind == "fuse_right")
    ch = np.full(len(mps), -1, np.int32)
    for i, b in picks:
        ch[i] = b
    return dict(n=n, feature_of_point=ch)


class _Matched(list):
    """vpMatched whose iteration (the spAlreadyFound set of :490) also yields the points flagged already-found,
    which in the reference sit in vpMatched at indices outside this test's keyframe features."""
    def __init__(self, items, extra):
        super().__init__(items)
        self.extra = extra

    def __iter__(self):
        yield from list.__iter__(self)
        yield from self.extra


def golden_sim3(name):
    c = pc.sim3_inputs(name)
    K1, K2 = kf_of(c["KF1"]), kf_of(c["KF2"])
    for K, P in ((K1, c["mps1"]), (K2, c["mps2"])):
        pts = kf_points(P)
        K.mvpMapPoints = [m if P["has_mp"][i] else None for i, m in enumerate(pts)]
    # vpMatches12: the already-matched KF1 features hold a point that KF2 does not observe (index -1)
    vpMatches12 = [Old() if c["mps1"]["already_matched"][i] else None for i in range(K1.N)]
    sR12, sR21, t21 = pj.sim3_relative(c["s12"], c["R12"], c["t12"])
    T12 = np.concatenate([sR12, c["t12"].reshape(3, 1)], 1)
    T21 = np.concatenate([sR21, t21.reshape(3, 1)], 1)
    # vbAlreadyMatched2 comes from GetIndexInKeyFrame in the reference; the scenario sets it directly
    already2 = c["mps2"]["already_matched"]
    n, v1, v2 = search_by_sim3_with_mask(K1, K2, vpMatches12, already2, c["th"], T12, T21, c)
    return dict(n=n, vnMatch1=np.array(v1, np.int32), vnMatch2=np.array(v2, np.int32))


def search_by_sim3_with_mask(K1, K2, vpMatches12, already2, th, T12, T21, c):
    marked = [i for i in range(K2.N) if already2[i]]
    holders = [m for m in vpMatches12 if m is not None]
    # each already-matched KF1 entry reports one of the marked KF2 indices (round robin), the rest report -1
    table = {id(m): (marked[k] if k < len(marked) else -1) for k, m in enumerate(holders)}
    assert len(holders) >= len(marked), "scenario: more marked KF2 features than already-matched KF1 entries"
    return pp.search_by_sim3(K1, K2, vpMatches12, lambda m: table[id(m)], th, pc.CAM,
                             lambda p: pj.transform_point(c["T1w"], p), lambda p: pj.transform_point(c["T2w"], p),
                             lambda p: pj.transform_point(T21, p), lambda p: pj.transform_point(T12, p))


def main():
    out = os.path.join(ROOT, "tests", "golden")
    for name in pc.MP_CASES:
        c = pc.mp_inputs(name)
        F, old = frame_of(c["F"])
        mp = c["mp"]
        mps = []
        for i in range(len(mp["desc"])):
            m = pp.MapPoint(mp["desc"][i], int(mp["obs"][i]), bool(mp["bad"][i]))
            m.mbTrackInView, m.mbTrackInViewR = bool(mp["in_view"][i]), bool(mp["in_view_r"][i])
            m.mTrackProjX, m.mTrackProjY, m.mTrackProjXR, m.mTrackProjYR = (np.float32(mp[k][i]) for k in
                                                                            ("proj_x", "proj_y", "proj_xr", "proj_yr"))
            m.mTrackDepth, m.mnTrackScaleLevel, m.mnTrackScaleLevelR = np.float32(mp["depth"][i]), int(mp["level"][i]), int(mp["level_r"][i])
            m.mTrackViewCos, m.mTrackViewCosR = np.float32(mp["view_cos"][i]), np.float32(mp["view_cos_r"][i])
            mps.append(m)
        n = pp.search_by_projection_mappoints(F, mps, c["th"], c["nnratio"], c["far"], c["th_far"])
        res = result_of(F, old, mps)
        np.savez_compressed(os.path.join(out, f"{name}.npz"), in_crc=crc(c["F"].kps, c["F"].desc, mp["desc"], mp["proj_x"]),
                            mappoint_of_feature=res, n=n)
        print(name, "nmatches", n, "assigned", int((res >= 0).sum()))
    for name in pc.LF_CASES:
        c = pc.lf_inputs(name)
        Cur, old = frame_of(c["Cur"])
        last = c["last"]
        m = len(last["has_mp"])
        Last = pp.Frame(np.zeros(m, c["Cur"].kps.dtype), np.zeros((m, 32), np.uint8), pc.SCALE, (0, 0, pc.W, pc.H))
        Last.keys["octave"], Last.keys["angle"] = last["octave"], last["angle"]
        mps = []
        for i in range(m):
            mpt = pp.MapPoint(last["desc"][i], int(last["obs"][i]), False, last["world_pos"][i])
            mps.append(mpt)
            Last.mvpMapPoints[i] = mpt if last["has_mp"][i] else None
            Last.mvbOutlier[i] = bool(last["outlier"][i])
        n = pp.search_by_projection_lastframe(Cur, Last, c["th"], False, c["ori"], c["project"],
                                              lambda mpt: pj.transform_point(c["Tcw"], mpt.mWorldPos), c["fb"], c["project_right"])
        idx = {id(x): i for i, x in enumerate(mps)}
        res = np.full(Cur.N, -1, np.int32)
        for i, x in enumerate(Cur.mvpMapPoints):
            if x is old:
                continue
            blocked0 = c["Cur"].blocked[i]
            res[i] = idx[id(x)] if x is not None else (-2 if not blocked0 and False else -1)
        np.savez_compressed(os.path.join(out, f"{name}.npz"), in_crc=crc(c["Cur"].kps, c["Cur"].desc, last["desc"], last["world_pos"]),
                            lastframe_feature_of_feature=res, n=n)
        print(name, "nmatches", n, "assigned", int((res >= 0).sum()))
    for name in pc.KF_CASES:
        g = golden_kf(name)
        c = pc.kf_inputs(name)
        np.savez_compressed(os.path.join(out, f"{name}.npz"), in_crc=crc(c["KF"].kps, c["KF"].desc, c["pts"]["desc"], c["pts"]["world_pos"]), **g)
        key = "point_of_feature" if "point_of_feature" in g else "feature_of_point"
        print(name, "n", g["n"], "assigned", int((g[key] >= 0).sum()))
    for name in pc.SIM3_CASES:
        g = golden_sim3(name)
        c = pc.sim3_inputs(name)
        np.savez_compressed(os.path.join(out, f"{name}.npz"), in_crc=crc(c["KF1"].kps, c["KF2"].kps, c["mps1"]["desc"], c["mps2"]["world_pos"]), **g)
        print(name, "nFound", g["n"], "one-way", int((g["vnMatch1"] >= 0).sum()), int((g["vnMatch2"] >= 0).sum()))


def golden_init(name):
    c = pc.init_inputs(name)
    fr = lambda G: pp.Frame(G.kps, G.desc, G.scale_factors, (G.min_x, G.min_y, G.max_x, G.max_y))
    prev = [list(p) for p in c["prev"]]
    n, v12 = pp.search_for_initialization(fr(c["F1"]), fr(c["F2"]), prev, c["win"], c["nnratio"], c["ori"])
    return dict(n=n, vnMatches12=np.array(v12, np.int32), prev_matched=np.array(prev, np.float32),
                in_crc=crc(c["F1"].kps, c["F1"].desc, c["F2"].kps, c["F2"].desc, c["prev"]))


def main_init():
    out = os.path.join(ROOT, "tests", "golden")
    for name in pc.INIT_CASES:
        g = golden_init(name)
        np.savez_compressed(os.path.join(out, f"{name}.npz"), **g)
        print(name, "nmatches", g["n"], "matched", int((g["vnMatches12"] >= 0).sum()))


if __name__ == "__main__":
    if "--init-only" in sys.argv:
        main_init()
        sys.exit(0)
    main()
    main_init()
