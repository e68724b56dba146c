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
                             ("skip_if_prev_ratio_rejected", "u1"), ("chi2", "u1"), ("skip_if_prev_empty", "u1"), ("pad", "u1"),
                             ("desc", "u1", 32)])
assert PROJ_QUERY_DTYPE.itemsize == 64


class GridFrame:
    """The Frame being searched.  kps = mvKeysUn (nleft == -1) or mvKeys ++ mvKeysRight; bounds = (mnMinX, mnMinY,
    mnMaxX, mnMaxY); blocked[i] = mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0."""

    def __init__(self, kps, desc, bounds, scale_factors, nleft=-1, blocked=None, u_right=None, left_to_right=None,
                 right_to_left=None, mbf=0.0, mb=0.0, inv_level_sigma2=None):
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
        self.inv_level_sigma2 = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, F32)


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
            e["skip_if_prev_empty"] = 1         # :2236-2237: an empty left window `continue`s past the right camera too
            q.append(e); owner.append(i)
    return (np.array(q, PROJ_QUERY_DTYPE) if q else np.zeros(0, PROJ_QUERY_DTYPE)), np.array(owner, np.int32)


_libm_logf = None


def logf(x):
    """log() of a float argument as the reference's translation units see it: `using namespace std` puts std::log(float) in
    scope, so `log(ratio)` and `log(mfScaleFactor)` (src/Frame.cc:109) are the C library's logf, not the double log."""
    global _libm_logf
    if _libm_logf is None:
        import ctypes.util
        f = C.CDLL(ctypes.util.find_library("m") or "libm.so.6").logf
        f.restype, f.argtypes = C.c_float, [C.c_float]
        _libm_logf = f
    return F32(_libm_logf(float(F32(x))))


def predict_scale(max_distance, current_dist, log_scale_factor, n_levels):
    """MapPoint::PredictScale (src/MapPoint.cc:534-566): `ceil(log(ratio) / mfLogScaleFactor)` is evaluated entirely in
    float (logf, float division, ceilf) - checked against the reference's own object code in tests/test_reference_mappoint.py."""
    import math
    ratio = F32(F32(max_distance) / F32(current_dist))
    if not (ratio > 0 and np.isfinite(ratio)):          # log(0) = -inf, x / 0 = inf, nan: the int conversion gives INT_MIN on x86 -> level 0
        return 0
    n = int(math.ceil(float(F32(logf(ratio) / F32(log_scale_factor)))))
    return 0 if n < 0 else min(n, n_levels - 1)


def distance_invariance(mf_min_distance, mf_max_distance):
    """MapPoint::GetMinDistanceInvariance / GetMaxDistanceInvariance (src/MapPoint.cc:522-532): 0.8f * mfMinDistance,
    1.2f * mfMaxDistance."""
    return F32(F32(0.8) * F32(mf_min_distance)), F32(F32(1.2) * F32(mf_max_distance))


def camera_centre(Rcw, tcw):
    """Ow = -Rcw.t() * tcw (a cv::Mat product: float32 terms accumulated in double, stored as float32)."""
    R, t = np.asarray(Rcw, F32).astype(np.float64), np.asarray(tcw, F32).astype(np.float64)
    return (-(R.T @ t)).astype(F32)


def _empty_or(q):
    return np.array(q, PROJ_QUERY_DTYPE) if q else np.zeros(0, PROJ_QUERY_DTYPE)


def queries_from_keyframe(Cur, kf, Tcw, th, project, log_scale_factor):
    """SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist) :2391-2510 (relocalisation).
    `kf`: arrays over pKF->GetMapPointMatches(): has_mp, bad, already_found, world_pos (n, 3), min_dist, max_dist
    (mfMinDistance, mfMaxDistance), angle (pKF->mvKeysUn[i].angle), desc.  In this overload ANY MapPoint already in
    CurrentFrame.mvpMapPoints blocks a feature (:2455), so the frame's `blocked` mask is "entry non-NULL" and every
    window claims (obs = 1); search with mode 0, th = ORBdist, rot_hist = 1."""
    th = F32(th)
    T = np.asarray(Tcw, F32)
    Ow = camera_centre(T[:3, :3], T[:3, 3])
    q, owner = [], []
    for i in range(len(kf["has_mp"])):
        if not kf["has_mp"][i] or kf["bad"][i] or kf["already_found"][i]:
            continue
        x3Dc = transform_point(T, kf["world_pos"][i])
        u, v = project(x3Dc)
        if u < Cur.min_x or u > Cur.max_x or v < Cur.min_y or v > Cur.max_y:
            continue
        PO = (np.asarray(kf["world_pos"][i], F32) - Ow).astype(np.float64)
        dist3D = F32(np.sqrt((PO * PO).sum()))
        lo, hi = distance_invariance(kf["min_dist"][i], kf["max_dist"][i])
        if dist3D < lo or dist3D > hi:
            continue
        lev = predict_scale(kf["max_dist"][i], dist3D, log_scale_factor, len(Cur.scale_factors))
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, F32(th * Cur.scale_factors[lev]), lev - 1, lev + 1
        e["angle"], e["obs"], e["valid"], e["desc"] = kf["angle"][i], 1, 1, kf["desc"][i]
        q.append(e); owner.append(i)
    return _empty_or(q), np.array(owner, np.int32)


# ---- KeyFrame-side searches: loop closing / merging / local mapping ------------------------------------------------
# KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:737-781) walks the same 64 x 48 grid (copied from the Frame, :60) with
# the same floor / ceil cell range and |dx| < r, |dy| < r test; it has no level filter, the callers below apply
# nPredictedLevel - 1 <= octave <= nPredictedLevel themselves, which is min_level / max_level of the window.

def is_in_image(KF, u, v):
    """KeyFrame::IsInImage (src/KeyFrame.cc:783-786): half-open on the right / bottom."""
    return bool(u >= KF.min_x and u < KF.max_x and v >= KF.min_y and v < KF.max_y)


def decompose_sim3(Scw):
    """:483-487 / :1830-1835: scw = sqrt(sRcw.row(0).dot(sRcw.row(0))) (double dot, float result), Rcw = sRcw / scw and
    tcw = Scw.col(3) / scw (cv::Mat / double = a float32 multiply by (float)(1.0 / scw)), Ow = -Rcw.t() * tcw."""
    S = np.asarray(Scw, F32)
    scw = F32(np.sqrt((S[0, :3].astype(np.float64) ** 2).sum()))
    inv = F32(1.0 / float(scw))
    Rcw, tcw = (S[:3, :3] * inv).astype(F32), (S[:3, 3] * inv).astype(F32)
    return Rcw, tcw, camera_centre(Rcw, tcw)


def _viewing_window(KF, wp, normal, mf_min, mf_max, Ow, log_scale_factor):
    """The distance / viewing-angle gates and PredictScale shared by :523-541, :1706-1728, :1875-1892.  Returns the
    predicted level or None."""
    PO = (np.asarray(wp, F32) - Ow).astype(F32)
    dist = F32(np.sqrt((PO.astype(np.float64) ** 2).sum()))             # cv::norm -> double -> const float
    lo, hi = distance_invariance(mf_min, mf_max)
    if dist < lo or dist > hi:
        return None
    if float(PO.astype(np.float64) @ np.asarray(normal, F32).astype(np.float64)) < 0.5 * float(dist):
        return None
    return predict_scale(mf_max, dist, log_scale_factor, len(KF.scale_factors))


def queries_from_sim3_points(KF, pts, Scw, th, project, log_scale_factor, claim=True):
    """SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th, ratioHamming) :473-585, its twin with
    vpPointsKFs / vpMatchedKF :588-700 (claim=True: KF.blocked = vpMatched[idx] != NULL, every window claims, search
    with mode 0 and th = hamming_threshold(ratioHamming)) and Fuse(KeyFrame*, cv::Mat Scw, vpPoints, th,
    vpReplacePoint) :1823-1945 (claim=False: nothing blocks, read `choice`, th = TH_LOW).
    `pts`: arrays over vpPoints: bad, already_found (in spAlreadyFound), world_pos, normal, min_dist, max_dist
    (mfMinDistance, mfMaxDistance), desc.  Returns (queries, vpPoints index of each query)."""
    th = F32(th)
    Rcw, tcw, Ow = decompose_sim3(Scw)
    T = np.concatenate([Rcw, tcw[:, None]], 1)
    q, owner = [], []
    for i in range(len(pts["desc"])):
        if pts["bad"][i] or pts["already_found"][i]:
            continue
        p3Dc = transform_point(T, pts["world_pos"][i])
        if p3Dc[2] < 0.0:
            continue
        u, v = project(p3Dc)
        if not is_in_image(KF, u, v):
            continue
        lev = _viewing_window(KF, pts["world_pos"][i], pts["normal"][i], pts["min_dist"][i], pts["max_dist"][i], Ow, log_scale_factor)
        if lev is None:
            continue
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, F32(th * KF.scale_factors[lev]), lev - 1, lev
        e["obs"], e["valid"], e["desc"] = 1 if claim else 0, 1, pts["desc"][i]
        q.append(e); owner.append(i)
    return _empty_or(q), np.array(owner, np.int32)


def hamming_threshold(th_low, ratio_hamming):
    """`bestDist <= TH_LOW * ratioHamming` (:576, :694) compares an int with a float product: the largest integer that
    passes."""
    import math
    return int(math.floor(float(F32(F32(th_low) * F32(ratio_hamming)))))


def queries_for_fuse(KF, pts, Rcw, tcw, Ow, th, project, log_scale_factor, bRight=False):
    """Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, th, bRight) :1629-1821.  (Rcw, tcw, Ow) and `project`
    belong to the camera selected by bRight.  `pts`: null, bad, in_keyframe (pMP->IsInKeyFrame(pKF)), world_pos,
    normal, min_dist, max_dist, desc.  The windows never claim (obs = 0) and carry the chi-square gate; search with
    mode 0, th = TH_LOW, KF.inv_level_sigma2 set; the caller applies `choice` in order (:1787-1812)."""
    th = F32(th)
    T = np.concatenate([np.asarray(Rcw, F32), np.asarray(tcw, F32)[:, None]], 1)
    Ow = np.asarray(Ow, F32)
    q, owner = [], []
    for i in range(len(pts["desc"])):
        if pts["null"][i] or pts["bad"][i] or pts["in_keyframe"][i]:
            continue
        p3Dc = transform_point(T, pts["world_pos"][i])
        if p3Dc[2] < 0.0:
            continue
        invz = F32(F32(1.0) / p3Dc[2])
        u, v = project(p3Dc)
        if not is_in_image(KF, u, v):
            continue
        ur = F32(u - F32(KF.mbf * invz))
        lev = _viewing_window(KF, pts["world_pos"][i], pts["normal"][i], pts["min_dist"][i], pts["max_dist"][i], Ow, log_scale_factor)
        if lev is None:
            continue
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, F32(th * KF.scale_factors[lev]), lev - 1, lev
        e["ur"], e["chi2"], e["right"], e["valid"], e["desc"] = ur, 1, 1 if bRight else 0, 1, pts["desc"][i]
        q.append(e); owner.append(i)
    return _empty_or(q), np.array(owner, np.int32)


def sim3_relative(s12, R12, t12):
    """:1964-1967: sR12 = s12 * R12, sR21 = (1.0 / s12) * R12.t(), t21 = -sR21 * t12."""
    R12, t12 = np.asarray(R12, F32), np.asarray(t12, F32).reshape(3)
    sR12 = (R12 * F32(s12)).astype(F32)
    sR21 = (R12.T * F32(1.0 / float(F32(s12)))).astype(F32)
    t21 = (-(sR21.astype(np.float64) @ t12.astype(np.float64))).astype(F32)
    return sR12, sR21, t21


def queries_for_sim3_direction(KFdst, src, Tsw, sR, t, th, cam, log_scale_factor):
    """One direction of SearchBySim3 :1991-2067 (KF1 -> KF2) / :2069-2145 (KF2 -> KF1).  `src`: arrays over the source
    keyframe's GetMapPointMatches(): has_mp, already_matched (vbAlreadyMatched), bad, world_pos, min_dist, max_dist,
    desc.  Tsw = source camera from world (3x4), (sR, t) = destination camera from source camera.  Windows never
    claim (obs = 0); search with mode 0, th = TH_HIGH; `choice` of window k = vnMatch[owner[k]]."""
    th = F32(th)
    fx, fy, cx, cy = (F32(c) for c in cam[:4])
    Tds = np.concatenate([np.asarray(sR, F32), np.asarray(t, F32).reshape(3, 1)], 1)
    q, owner = [], []
    for i in range(len(src["has_mp"])):
        if not src["has_mp"][i] or src["already_matched"][i] or src["bad"][i]:
            continue
        pd = transform_point(Tds, transform_point(Tsw, src["world_pos"][i]))
        if pd[2] < 0.0:
            continue
        invz = F32(1.0 / float(pd[2]))
        x, y = F32(pd[0] * invz), F32(pd[1] * invz)
        u, v = F32(F32(fx * x) + cx), F32(F32(fy * y) + cy)
        if not is_in_image(KFdst, u, v):
            continue
        dist3D = F32(np.sqrt((np.array(pd, np.float64) ** 2).sum()))
        lo, hi = distance_invariance(src["min_dist"][i], src["max_dist"][i])
        if dist3D < lo or dist3D > hi:
            continue
        lev = predict_scale(src["max_dist"][i], dist3D, log_scale_factor, len(KFdst.scale_factors))
        e = np.zeros((), PROJ_QUERY_DTYPE)
        e["u"], e["v"], e["r"], e["min_level"], e["max_level"] = u, v, F32(th * KFdst.scale_factors[lev]), lev - 1, lev
        e["valid"], e["desc"] = 1, src["desc"][i]
        q.append(e); owner.append(i)
    return _empty_or(q), np.array(owner, np.int32)


def sim3_agreement(n1, n2, q12, q21):
    """:2147-2168: vnMatch1 / vnMatch2 from the two directions' (owner, choice) and the mutual-agreement pass.
    Returns (vnMatch1, vnMatch2, [(i1, idx2) accepted], nFound)."""
    vn1, vn2 = np.full(n1, -1, np.int32), np.full(n2, -1, np.int32)
    vn1[q12[0]] = q12[1]
    vn2[q21[0]] = q21[1]
    pairs = [(i1, int(vn1[i1])) for i1 in range(n1) if vn1[i1] >= 0 and vn2[vn1[i1]] == i1]
    return vn1, vn2, pairs, len(pairs)


def search_by_sim3(search, KF1, KF2, mps1, mps2, T1w, T2w, s12, R12, t12, th, cam, log_scale_factor, th_high=100):
    """ORBmatcher::SearchBySim3 :1947-2171 on top of a window-search callable
    `search(F, queries, mode, th) -> (assign, choice, n)` (ORBmatcher.search_by_projection): two launches and the
    agreement pass.  mps1 / mps2 as `src` of queries_for_sim3_direction."""
    sR12, sR21, t21 = sim3_relative(s12, R12, t12)
    q1, o1 = queries_for_sim3_direction(KF2, mps1, T1w, sR21, t21, th, cam, log_scale_factor)
    q2, o2 = queries_for_sim3_direction(KF1, mps2, T2w, sR12, np.asarray(t12, F32).reshape(3), th, cam, log_scale_factor)
    c1 = search(KF2, q1, 0, th_high)[1]
    c2 = search(KF1, q2, 0, th_high)[1]
    return sim3_agreement(len(mps1["has_mp"]), len(mps2["has_mp"]), (o1, c1), (o2, c2))
