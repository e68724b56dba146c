"""Synthetic projection-search scenarios shared by the golden generator (object-level Python restatement), the CPU
oracle tests and the GPU parity tests (TEST INFRASTRUCTURE).  Features are synthetic keypoints (positions snapped
to 1/16 px, octaves by the extractor's quota shape, random descriptors); MapPoints are noisy copies of a subset
of them, ordered at random, with deliberate competition (several MapPoints near the same feature), distractors,
pre-assigned features, stereo coordinates and points without observations."""
import numpy as np

import orbfe
from orbfe import projection as pj

SCALE = np.array([1.2 ** i for i in range(8)], np.float32)

# SearchByProjection(Frame&, vpMapPoints, th): name -> (seed, features, mappoints, th, nnratio, fisheye-stereo rig, far filter)
MP_CASES = {
    "proj_local_mono": (4100, 1500, 1200, 1.0, 0.8, False, False),
    "proj_local_th3": (4101, 1200, 1500, 3.0, 0.8, False, True),
    "proj_local_fisheye": (4102, 1600, 1100, 1.0, 0.8, True, False),
}
# SearchByProjection(CurrentFrame, LastFrame, th, bMono): name -> (seed, features, th, (bForward, bBackward), fisheye rig, check_ori)
LF_CASES = {
    "proj_last_mono": (4200, 1400, 15.0, (False, False), False, True),
    "proj_last_forward": (4201, 1400, 7.0, (True, False), False, True),
    "proj_last_backward_noori": (4202, 1000, 7.0, (False, True), False, False),
    "proj_last_fisheye": (4203, 1500, 15.0, (False, False), True, True),
}
W, H = 752, 480
CAM = np.array([458.654, 457.296, 367.215, 248.375], np.float32)          # EuRoC cam0 (Examples/Monocular/EuRoC.yaml:9-12)


def snap(a):
    return (np.round(np.asarray(a) * 16) / 16).astype(np.float32)


def make_frame(rng, n, rig):
    """Returns a projection.GridFrame without assignments.  rig: features are mvKeys ++ mvKeysRight with stereo links."""
    k = np.zeros(n, orbfe.KEYPOINT_DTYPE)
    k["x"], k["y"] = snap(rng.uniform(2, W - 2, n)), snap(rng.uniform(2, H - 2, n))
    k["octave"] = np.minimum(rng.geometric(0.3, n) - 1, 7)
    k["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    k["size"], k["class_id"] = 31, -1
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    if not rig:
        ur = np.where(rng.random(n) < 0.6, k["x"] - snap(rng.uniform(2, 40, n)), -1).astype(np.float32)
        return pj.GridFrame(k, d, (0, 0, W, H), SCALE, u_right=ur, mbf=47.9, mb=0.11)
    nleft = n * 55 // 100
    l2r, r2l = np.full(nleft, -1, np.int32), np.full(n - nleft, -1, np.int32)
    pairs = rng.permutation(min(nleft, n - nleft))[: nleft // 3]
    for a, b in zip(rng.permutation(nleft)[:len(pairs)], pairs):
        l2r[a], r2l[b] = b, a
    return pj.GridFrame(k, d, (0, 0, W, H), SCALE, nleft=nleft, left_to_right=l2r, right_to_left=r2l)


def noisy_desc(rng, d, maxflip):
    out = d.copy()
    for b in rng.choice(256, int(rng.integers(0, maxflip)), replace=False):
        out[b >> 3] ^= 1 << (b & 7)
    return out


def mp_inputs(name):
    seed, n, m, th, nnratio, rig, far = MP_CASES[name]
    rng = np.random.default_rng(seed)
    F = make_frame(rng, n, rig)
    blocked = rng.random(n) < 0.15
    F.blocked = blocked.astype(np.uint8)
    mp = dict(in_view=np.zeros(m, bool), in_view_r=np.zeros(m, bool), bad=rng.random(m) < 0.03,
              depth=rng.uniform(1, 80, m).astype(np.float32), proj_x=np.zeros(m, np.float32), proj_y=np.zeros(m, np.float32),
              proj_xr=np.zeros(m, np.float32), proj_yr=np.zeros(m, np.float32), level=np.zeros(m, np.int32),
              level_r=np.full(m, -1, np.int32), view_cos=np.where(rng.random(m) < 0.5, 0.9995, 0.99).astype(np.float32),
              view_cos_r=np.where(rng.random(m) < 0.5, 0.9995, 0.99).astype(np.float32),
              desc=np.zeros((m, 32), np.uint8), obs=(rng.random(m) < 0.97).astype(np.int32) * 3)
    nl = F.nleft if rig else n
    for i in range(m):
        kind = rng.random()
        src = int(rng.integers(0, nl)) if i % 5 else int(rng.integers(0, min(nl, 40)))     # every 5th competes for 40 features
        kp = F.kps[src]
        if kind < 0.85:
            mp["in_view"][i] = True
            mp["proj_x"][i], mp["proj_y"][i] = snap(kp["x"] + rng.normal(0, 1.5)), snap(kp["y"] + rng.normal(0, 1.5))
            mp["level"][i] = int(np.clip(kp["octave"] + rng.integers(0, 2), 0, 7))
            mp["desc"][i] = noisy_desc(rng, F.desc[src], 70)
            if not rig:
                mp["proj_xr"][i] = snap((F.u_right[src] if F.u_right[src] > 0 else kp["x"] - 10) + rng.normal(0, 2.0))
        else:
            mp["in_view"][i] = kind < 0.93
            mp["proj_x"][i], mp["proj_y"][i] = snap(rng.uniform(-20, W + 20)), snap(rng.uniform(-20, H + 20))
            mp["level"][i] = int(rng.integers(0, 8))
            mp["desc"][i] = rng.integers(0, 256, 32, dtype=np.uint8)
        if rig and rng.random() < 0.6:
            rs = int(rng.integers(nl, n))
            if F.left_to_right[src] != -1 and rng.random() < 0.7:
                rs = nl + int(F.left_to_right[src])
            kr = F.kps[rs]
            mp["in_view_r"][i] = True
            mp["proj_xr"][i], mp["proj_yr"][i] = snap(kr["x"] + rng.normal(0, 1.5)), snap(kr["y"] + rng.normal(0, 1.5))
            mp["level_r"][i] = int(np.clip(kr["octave"] + rng.integers(0, 2), 0, 7)) if rng.random() < 0.95 else -1
            if not mp["in_view"][i]:
                mp["desc"][i] = noisy_desc(rng, F.desc[rs], 70)
    return dict(F=F, mp=mp, th=th, nnratio=nnratio, far=far, th_far=40.0)


def lf_inputs(name):
    """LastFrame -> CurrentFrame: world points in front of the current camera project (pinhole) next to current
    features; LastFrame's MapPoint carries that feature's descriptor with noise."""
    seed, n, th, fb, rig, ori = LF_CASES[name]
    rng = np.random.default_rng(seed)
    Cur = make_frame(rng, n, rig)
    Cur.blocked = (rng.random(n) < 0.1).astype(np.uint8)
    Tcw = np.eye(4, dtype=np.float32)
    Tcw[:3, 3] = (0.03, -0.01, 0.02)
    Trl = np.eye(4, dtype=np.float32)
    Trl[:3, 3] = (-0.101, 0.0, 0.0)
    nl = Cur.nleft if rig else n
    m = n
    last = dict(has_mp=rng.random(m) < 0.8, outlier=rng.random(m) < 0.05, octave=np.zeros(m, np.int32),
                angle=np.zeros(m, np.float32), world_pos=np.zeros((m, 3), np.float32), desc=np.zeros((m, 32), np.uint8),
                obs=(rng.random(m) < 0.8).astype(np.int32))
    for i in range(m):
        src = int(rng.integers(0, nl)) if i % 6 else int(rng.integers(0, min(nl, 30)))
        kp = Cur.kps[src]
        z = float(rng.uniform(0.8, 12.0)) if rng.random() < 0.97 else -1.0          # a few behind the camera
        u, v = kp["x"] + rng.normal(0, 2.5), kp["y"] + rng.normal(0, 2.5)
        if rng.random() < 0.05:
            u += 900                                                                 # outside the image bounds
        pc = np.array([(u - CAM[2]) / CAM[0] * z, (v - CAM[3]) / CAM[1] * z, z])
        last["world_pos"][i] = (pc - Tcw[:3, 3].astype(np.float64)).astype(np.float32)   # R = I
        last["octave"][i] = int(np.clip(kp["octave"] + rng.integers(-1, 2), 0, 7))
        last["angle"][i] = np.float32((kp["angle"] + rng.normal(0, 6) + (0 if rng.random() < 0.9 else 90)) % 360)
        last["desc"][i] = noisy_desc(rng, Cur.desc[src], 80)
    project = lambda p: pj.pinhole_project(CAM, p)
    project_right = lambda p: pj.pinhole_project(CAM, pj.transform_point(Trl, p))
    return dict(Cur=Cur, last=last, Tcw=Tcw, th=th, fb=fb, ori=ori, project=project, project_right=project_right)


# ---- KeyFrame-side searches (SearchByProjection(KeyFrame*, Scw, ...), Fuse, SearchBySim3) -------------------------
LOG_SF = np.float32(np.log(np.float32(1.2)))
INV_SIGMA2 = (1.0 / (SCALE.astype(np.float64) ** 2)).astype(np.float32)
# name -> (seed, features, points, th, kind); kind: "sim3" :473, "fuse" :1629, "fuse_right" :1629 bRight, "fuse_sim3" :1823
KF_CASES = {
    "proj_kf_sim3": (4300, 1500, 1300, 4.0, "sim3"),
    "proj_kf_sim3_ratio": (4301, 1200, 1000, 8.0, "sim3"),
    "proj_kf_fuse": (4302, 1500, 1400, 3.0, "fuse"),
    "proj_kf_fuse_right": (4303, 1600, 1200, 3.0, "fuse_right"),
    "proj_kf_fuse_sim3": (4304, 1400, 1200, 4.0, "fuse_sim3"),
}
SIM3_CASES = {"proj_sim3_pair": (4400, 1300, 7.5)}


def rodrigues(w):
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.eye(3)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K


def pose(rng, rot=0.05, trans=0.3):
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] = rodrigues(rng.normal(0, rot, 3)).astype(np.float32)
    T[:3, 3] = rng.normal(0, trans, 3).astype(np.float32)
    return T


def points_seen_by(rng, KF, m, Tcw, lo_feat, hi_feat, compete=40, px_noise=1.5, flips=60):
    """m world points whose projections (pinhole CAM, pose Tcw) fall next to features lo_feat..hi_feat of KF; every 5th
    aims at one of `compete` features.  Returns the per-point arrays the flatteners read."""
    R, t = Tcw[:3, :3].astype(np.float64), Tcw[:3, 3].astype(np.float64)
    Ow = -R.T @ t
    P = dict(bad=rng.random(m) < 0.03, world_pos=np.zeros((m, 3), np.float32), normal=np.zeros((m, 3), np.float32),
             min_dist=np.zeros(m, np.float32), max_dist=np.zeros(m, np.float32), desc=np.zeros((m, 32), np.uint8),
             src=np.zeros(m, np.int32))
    for i in range(m):
        src = int(rng.integers(lo_feat, hi_feat)) if i % 5 else int(rng.integers(lo_feat, min(hi_feat, lo_feat + compete)))
        kp = KF.kps[src]
        z = float(rng.uniform(0.8, 15.0)) if rng.random() < 0.97 else -2.0
        u, v = kp["x"] + rng.normal(0, px_noise) * SCALE[kp["octave"]], kp["y"] + rng.normal(0, px_noise) * SCALE[kp["octave"]]
        if rng.random() < 0.04:
            u += 800
        pc = np.array([(u - CAM[2]) / CAM[0] * z, (v - CAM[3]) / CAM[1] * z, z])
        pw = R.T @ (pc - t)
        P["world_pos"][i] = pw.astype(np.float32)
        PO = pw - Ow
        d = np.linalg.norm(PO)
        nrm = PO / d if rng.random() < 0.9 else rodrigues(rng.normal(0, 1.0, 3)) @ (PO / d)     # some seen from the side
        P["normal"][i] = nrm.astype(np.float32)
        mx = d * 1.2 ** (kp["octave"] + rng.uniform(-0.9, 0.9)) if rng.random() < 0.95 else d * rng.uniform(0.3, 0.9)
        P["max_dist"][i] = mx
        P["min_dist"][i] = mx / 1.2 ** 7 if rng.random() < 0.95 else d * 1.5
        P["desc"][i] = noisy_desc(rng, KF.desc[src], flips)
        P["src"][i] = src
    return P


def kf_inputs(name):
    seed, n, m, th, kind = KF_CASES[name]
    rng = np.random.default_rng(seed)
    rig = kind == "fuse_right"
    KF = make_frame(rng, n, rig)
    KF.inv_level_sigma2 = INV_SIGMA2
    if rig:
        KF.u_right = np.full(n, -1, np.float32)
    Tcw = pose(rng)
    nl = KF.nleft if rig else n
    c = dict(KF=KF, th=th, kind=kind, Tcw=Tcw, project=lambda p: pj.pinhole_project(CAM, p))
    if kind in ("sim3", "fuse_sim3"):
        s = np.float32(1.07)
        Scw = Tcw.copy()
        Scw[:3, :] *= s
        c["Scw"] = Scw
        # the flatteners see Rcw, tcw after the decomposition; generate the points with exactly those
        Rcw, tcw, _ = pj.decompose_sim3(Scw)
        T = np.eye(4, dtype=np.float32)
        T[:3, :3], T[:3, 3] = Rcw, tcw
        pts = points_seen_by(rng, KF, m, T, 0, n)
        pts["already_found"] = rng.random(m) < 0.1
        if kind == "sim3":
            KF.blocked = (rng.random(n) < 0.2).astype(np.uint8)           # vpMatched[idx] != NULL
            c["ratio"] = 1.0 if name == "proj_kf_sim3" else 1.5
    else:
        lo, hi = (nl, n) if rig else (0, n)
        pts = points_seen_by(rng, KF, m, Tcw, lo, hi)
        pts["null"], pts["in_keyframe"] = rng.random(m) < 0.03, rng.random(m) < 0.1
    c["pts"] = pts
    return c


def sim3_inputs(name):
    """Two keyframes seeing a common set of points; KF2's camera = Sim3(s12, R12, t12)^-1 of KF1's."""
    seed, n, th = SIM3_CASES[name]
    rng = np.random.default_rng(seed)
    KF1, KF2 = make_frame(rng, n, False), make_frame(rng, n, False)
    T1w, T2w = pose(rng), pose(rng)
    # x1 = s12 R12 x2 + t12 with x_i = T_iw X  =>  for s12 = 1: R12 = R1 R2^T, t12 = t1 - R12 t2
    R1, R2 = T1w[:3, :3].astype(np.float64), T2w[:3, :3].astype(np.float64)
    R12 = R1 @ R2.T
    t12 = T1w[:3, 3].astype(np.float64) - R12 @ T2w[:3, 3].astype(np.float64)
    out = dict(KF1=KF1, KF2=KF2, T1w=T1w, T2w=T2w, s12=np.float32(1.0), R12=R12.astype(np.float32), t12=t12.astype(np.float32), th=th)
    # map points of KF1: half of them are also visible in KF2 next to a KF2 feature that carries a similar descriptor
    m1 = points_seen_by(rng, KF1, n, T1w, 0, n, compete=n, px_noise=0.0, flips=20)
    m2 = points_seen_by(rng, KF2, n, T2w, 0, n, compete=n, px_noise=0.0, flips=20)
    for P, KFa, KFb, Tb in ((m1, KF1, KF2, T2w), (m2, KF2, KF1, T1w)):
        # feature i of KFa observes P[i]: put the point of feature i where feature src[i] sat, then swap roles so that
        # point i belongs to feature i (a permutation of the generated points)
        order = np.argsort(P["src"], kind="stable")
        _, first = np.unique(P["src"][order], return_index=True)
        keep = order[first]                                   # one point per distinct feature
        has = np.zeros(n, bool)
        for k in list(P):
            if k == "src":
                continue
            full = np.zeros((n,) + P[k].shape[1:], P[k].dtype)
            full[P["src"][keep]] = P[k][keep]
            P[k] = full
        has[P.pop("src")[keep]] = True
        P["has_mp"] = has
        P["already_matched"] = rng.random(n) < 0.05
    # make a third of KF1's points visible in KF2 as well: move a KF2 feature onto their projection, similar descriptor
    proj = lambda T, p: pj.pinhole_project(CAM, pj.transform_point(T, p))
    idx1 = np.flatnonzero(m1["has_mp"])
    for i1 in idx1[: len(idx1) // 3 * 2]:
        u, v = proj(T2w, m1["world_pos"][i1])
        if not (4 < u < W - 4 and 4 < v < H - 4):
            continue
        i2 = int(rng.integers(0, n))
        KF2.kps["x"][i2], KF2.kps["y"][i2] = snap(u + rng.normal(0, 1.0)), snap(v + rng.normal(0, 1.0))
        KF2.kps["octave"][i2] = KF1.kps["octave"][i1]
        KF2.desc[i2] = noisy_desc(rng, m1["desc"][i1], 30)
        if rng.random() < 0.8:                                 # and KF2's own map point there is the same 3-D point
            for k in ("world_pos", "normal", "min_dist", "max_dist"):
                m2[k][i2] = m1[k][i1]
            m2["desc"][i2], m2["has_mp"][i2], m2["bad"][i2] = noisy_desc(rng, m1["desc"][i1], 10), True, False
    # vbAlreadyMatched2 is derived from vpMatches12 in the reference (:1977-1987): never more marks than in KF1
    m2["already_matched"] = np.zeros(n, bool)
    m2["already_matched"][rng.permutation(n)[: int(m1["already_matched"].sum()) // 2]] = True
    out["mps1"], out["mps2"] = m1, m2
    return out


# SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) :706: name -> (seed, n1, n2, windowSize, nnratio, check_ori)
INIT_CASES = {
    "proj_init_w100": (4500, 2000, 2200, 100, 0.9, True),
    "proj_init_w30_noori": (4501, 1500, 1500, 30, 0.8, False),
}


def init_inputs(name=None, spec=None):
    """F2's level-0 features are the sources; an F1 feature copies one (position within the window, descriptor with
    up to 45 flipped bits).  Every 3rd F1 feature competes for one of 60 sources, so matches are stolen by later,
    closer descriptors, later but farther ones see vMatchedDistance[i2] <= dist and fall back to their next
    candidate; some sources get a near-duplicate neighbour so the ratio test fires."""
    seed, n1, n2, win, nnratio, ori = spec if spec is not None else INIT_CASES[name]
    rng = np.random.default_rng(seed)
    F2 = make_frame(rng, n2, False)
    lvl0 = np.flatnonzero(F2.kps["octave"] == 0)
    for s in lvl0[:: 7]:                                       # near-duplicates: same place, descriptor a few bits away
        t = int(rng.integers(0, n2))
        F2.kps["x"][t], F2.kps["y"][t] = snap(F2.kps["x"][s] + rng.normal(0, 2.0)), snap(F2.kps["y"][s] + rng.normal(0, 2.0))
        F2.kps["octave"][t] = 0
        F2.desc[t] = noisy_desc(rng, F2.desc[s], 12)
    lvl0 = np.flatnonzero(F2.kps["octave"] == 0)
    F1 = make_frame(rng, n1, False)
    prev = np.zeros((n1, 2), np.float32)
    for i in range(n1):
        if rng.random() < 0.75:
            s = int(lvl0[int(rng.integers(0, len(lvl0))) if i % 3 else int(rng.integers(0, min(60, len(lvl0))))])
            F1.kps["x"][i] = snap(np.clip(F2.kps["x"][s] + rng.normal(0, win / 4), 1, W - 1))
            F1.kps["y"][i] = snap(np.clip(F2.kps["y"][s] + rng.normal(0, win / 4), 1, H - 1))
            F1.desc[i] = noisy_desc(rng, F2.desc[s], 46)
            F1.kps["angle"][i] = np.float32((F2.kps["angle"][s] + (rng.normal(0, 4) if rng.random() < 0.85 else rng.uniform(0, 360))) % 360)
        prev[i] = (F1.kps["x"][i], F1.kps["y"][i]) if rng.random() < 0.8 else (snap(rng.uniform(-30, W + 30)), snap(rng.uniform(-30, H + 30)))
    return dict(F1=F1, F2=F2, prev=prev, win=win, nnratio=nnratio, ori=ori)
