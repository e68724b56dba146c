"""Host-side flattening of ORB-SLAM3's projection-guided searches onto orbfe_search_by_projection.

The reference's overloads differ only in how a MapPoint becomes a search window (which pose projects it, how the
scale level and the radius are predicted); what follows — Frame::GetFeaturesInArea, the best / second-best
descriptor scan, the ordered assignment into Frame::mvpMapPoints — is one loop, which runs on the GPU.  The
functions below restate the window construction of
  SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)   src/ORBmatcher.cc:44-236
  SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono)           src/ORBmatcher.cc:2173-2389
in float32 on plain arrays (in the C++ shim these lines are the reference's own) and emit `orbfe_proj_query`
records in the reference's loop order.
"""
import ctypes as C

import numpy as np

F32 = np.float32
FRAME_GRID_ROWS, FRAME_GRID_COLS = 48, 64

# include/orbfe.h: orbfe_proj_query (64 bytes)
PROJ_QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("ur", "<f4"), ("angle", "<f4"), ("min_level", "<i2"),
                             ("max_level", "<i2"), ("right", "u1"), ("obs", "u1"), ("check_ur", "u1"), ("valid", "u1"),
                             ("skip_if_prev_ratio_rejected", "u1"), ("pad", "u1", 3), ("desc", "u1", 32)])
assert PROJ_QUERY_DTYPE.itemsize == 64


class GridFrame:
    """The Frame being searched.  kps = mvKeysUn (nleft == -1) or mvKeys ++ mvKeysRight; bounds = (mnMinX, mnMinY,
    mnMaxX, mnMaxY); blocked[i] = mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0."""

    def __init__(self, kps, desc, bounds, scale_factors, nleft=-1, blocked=None, u_right=None, left_to_right=None,
                 right_to_left=None, mbf=0.0, mb=0.0):
        from ._lib import KEYPOINT_DTYPE
        self.kps = np.ascontiguousarray(kps, KEYPOINT_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.n, self.nleft = len(self.kps), int(nleft)
        self.min_x, self.min_y, self.max_x, self.max_y = (F32(b) for b in bounds)
        self.grid_w_inv = F32(F32(FRAME_GRID_COLS) / F32(self.max_x - self.min_x))        # Frame.cc:158-159
        self.grid_h_inv = F32(F32(FRAME_GRID_ROWS) / F32(self.max_y - self.min_y))
        self.scale_factors = np.asarray(scale_factors, F32)
        self.blocked = None if blocked is None else np.ascontiguousarray(blocked).astype(np.uint8)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, F32)
        self.left_to_right = None if left_to_right is None else np.ascontiguousarray(left_to_right, np.int32)
        self.right_to_left = None if right_to_left is None else np.ascontiguousarray(right_to_left, np.int32)
        self.mbf, self.mb = F32(mbf), F32(mb)


def radius_by_viewing_cos(view_cos):
    """ORBmatcher.cc:238-244"""
    return F32(2.5) if view_cos > F32(0.998) else F32(4.0)


def queries_from_mappoints(F, mp, th, bFarPoints=False, thFarPoints=50.0):
    """SearchByProjection(Frame&, vpMapPoints, th, ...) :44-236.  `mp` holds one array per MapPoint member the
    function reads: in_view, in_view_r, bad, depth, proj_x, proj_y, proj_xr, proj_yr, level, level_r, view_cos,
    view_cos_r, desc (n, 32), obs.  Returns (queries, mappoint index of each query)."""
    th = F32(th)
    b_factor = th != F32(1.0)
    q, owner = [], []
    for i in range(len(mp["desc"])):
        if not mp["in_view"][i] and not mp["in_view_r"][i]:
            continue
        if bFarPoints and mp["depth"][i] > F32(thFarPoints):
            continue
        if mp["bad"][i]:
            continue
        if mp["in_view"][i]:
            lev = int(mp["level"][i])
            r = radius_by_viewing_cos(F32(mp["view_cos"][i]))
            if b_factor:
                r = F32(r * th)
            e = np.zeros((), PROJ_QUERY_DTYPE)
            e["u"], e["v"], e["r"] = mp["proj_x"][i], mp["proj_y"][i], F32(r * F.scale_factors[lev])
            e["min_level"], e["max_level"] = lev - 1, lev
            e["ur"], e["check_ur"] = mp["proj_xr"][i], 1
            e["obs"], e["valid"], e["desc"] = mp["obs"][i] > 0, 1, mp["desc"][i]
            q.append(e); owner.append(i)
        if F.nleft != -1 and mp["in_view_r"][i]:
            lev = int(mp["level_r"][i])
            if lev != -1:
                r = radius_by_viewing_cos(F32(mp["view_cos_r"][i]))
                e = np.zeros((), PROJ_QUERY_DTYPE)
                e["u"], e["v"], e["r"] = mp["proj_xr"][i], mp["proj_yr"][i], F32(r * F.scale_factors[lev])
                e["min_level"], e["max_level"], e["right"] = lev - 1, lev, 1
                e["obs"], e["valid"], e["desc"] = mp["obs"][i] > 0, 1, mp["desc"][i]
                e["skip_if_prev_ratio_rejected"] = 1 if mp["in_view"][i] else 0
                q.append(e); owner.append(i)
    return (np.array(q, PROJ_QUERY_DTYPE) if q else np.zeros(0, PROJ_QUERY_DTYPE)), np.array(owner, np.int32)


def transform_point(T, p):
    """R * p + t with float32 products and sums in row order (T: 3x4 or 4x4 float32)."""
    T = np.asarray(T, F32)
    x, y, z = F32(p[0]), F32(p[1]), F32(p[2])
    return tuple(F32(F32(F32(F32(T[r, 0] * x) + F32(T[r, 1] * y)) + F32(T[r, 2] * z)) + T[r, 3]) for r in range(3))


def pinhole_project(cam, p):
    """Pinhole::project (src/CameraModels/Pinhole.cpp:34-42): fx * x / z + cx, fy * y / z + cy."""
    fx, fy, cx, cy = (F32(c) for c in cam[:4])
    return F32(F32(F32(fx * p[0]) / p[2]) + cx), F32(F32(F32(fy * p[1]) / p[2]) + cy)


def forward_backward(Tcw_cur, Tcw_last, mb, bMono):
    """:2184-2194: tlc = Rlw * twc + tlw;  bForward = tlc.z > mb && !bMono;  bBackward = -tlc.z > mb && !bMono."""
    Tc, Tl = np.asarray(Tcw_cur, F32), np.asarray(Tcw_last, F32)
    twc = -(Tc[:3, :3].T.astype(np.float64) @ Tc[:3, 3].astype(np.float64))
    tlc_z = F32(float(Tl[2, :3].astype(np.float64) @ twc) + float(Tl[2, 3]))
    return bool(tlc_z > F32(mb) and not bMono), bool(-tlc_z > F32(mb) and not bMono)


def queries_from_lastframe(Cur, last, Tcw, th, fwd_bwd, project=None, project_right=None):
    """SearchByProjection(CurrentFrame, LastFrame, th, bMono) :2173-2389.  `last`: arrays over LastFrame's
    features: has_mp, outlier, octave, angle, world_pos (n, 3), desc (n, 32: pMP->GetDescriptor()), obs.
    Returns (queries, LastFrame feature index of each query)."""
    th = F32(th)
    b_forward, b_backward = fwd_bwd
    q, owner = [], []
    for i in range(len(last["has_mp"])):
        if not last["has_mp"][i] or last["outlier"][i]:
            continue
        x3Dc = transform_point(Tcw, last["world_pos"][i])
        invzc = F32(1.0 / float(x3Dc[2]))
        if invzc < 0:
            continue
        u, v = project(x3Dc)
        if u < Cur.min_x or u > Cur.max_x or v < Cur.min_y or v > Cur.max_y:
            continue
        octave = int(last["octave"][i])
        radius = F32(th * Cur.scale_factors[octave])
        lo, hi = (octave, -1) if b_forward else (0, octave) if b_backward else (octave - 1, octave + 1)
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, radius, lo, hi
        e["ur"], e["check_ur"] = F32(u - F32(Cur.mbf * invzc)), 1
        e["angle"], e["obs"], e["valid"], e["desc"] = last["angle"][i], last["obs"][i] > 0, 1, last["desc"][i]
        q.append(e); owner.append(i)
        if Cur.nleft != -1:
            u, v = project_right(x3Dc)
            e = np.zeros((), PROJ_QUERY_DTYPE)
            e["u"], e["v"], e["r"], e["min_level"], e["max_level"], e["right"] = u, v, radius, lo, hi, 1
            e["angle"], e["obs"], e["valid"], e["desc"] = last["angle"][i], last["obs"][i] > 0, 1, last["desc"][i]
            q.append(e); owner.append(i)
    return (np.array(q, PROJ_QUERY_DTYPE) if q else np.zeros(0, PROJ_QUERY_DTYPE)), np.array(owner, np.int32)


def predict_scale(max_distance, current_dist, log_scale_factor, n_levels):
    """MapPoint::PredictScale (src/MapPoint.cc:551-566): log() of a float ratio in double, ceil, clamp."""
    import math
    ratio = F32(F32(max_distance) / F32(current_dist))
    n = int(math.ceil(math.log(float(ratio)) / float(F32(log_scale_factor))))
    return 0 if n < 0 else min(n, n_levels - 1)


def queries_from_keyframe(Cur, kf, Tcw, th, project, log_scale_factor):
    """SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist) :2391-2510 (relocalisation).
    `kf`: arrays over pKF->GetMapPointMatches(): has_mp, bad, already_found, world_pos (n, 3), min_dist, max_dist
    (Get{Min,Max}DistanceInvariance), angle (pKF->mvKeysUn[i].angle), desc.  In this overload ANY MapPoint already in
    CurrentFrame.mvpMapPoints blocks a feature (:2455), so the frame's `blocked` mask is "entry non-NULL" and every
    window claims (obs = 1); search with mode 0, th = ORBdist, rot_hist = 1."""
    th = F32(th)
    T = np.asarray(Tcw, F32)
    Ow = -(T[:3, :3].T.astype(np.float64) @ T[:3, 3].astype(np.float64))
    q, owner = [], []
    for i in range(len(kf["has_mp"])):
        if not kf["has_mp"][i] or kf["bad"][i] or kf["already_found"][i]:
            continue
        x3Dc = transform_point(T, kf["world_pos"][i])
        u, v = project(x3Dc)
        if u < Cur.min_x or u > Cur.max_x or v < Cur.min_y or v > Cur.max_y:
            continue
        PO = kf["world_pos"][i].astype(np.float64) - Ow
        dist3D = F32(np.sqrt((PO * PO).sum()))
        if dist3D < kf["min_dist"][i] or dist3D > kf["max_dist"][i]:
            continue
        lev = predict_scale(kf["max_dist"][i], dist3D, log_scale_factor, len(Cur.scale_factors))
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, F32(th * Cur.scale_factors[lev]), lev - 1, lev + 1
        e["angle"], e["obs"], e["valid"], e["desc"] = kf["angle"][i], 1, 1, kf["desc"][i]
        q.append(e); owner.append(i)
    return (np.array(q, PROJ_QUERY_DTYPE) if q else np.zeros(0, PROJ_QUERY_DTYPE)), np.array(owner, np.int32)
