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


if __name__ == "__main__":
    main()
