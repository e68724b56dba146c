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
