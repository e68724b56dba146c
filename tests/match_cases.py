"""Synthetic matcher problems shared by the golden-vector generator (oracle/py/make_golden_match.py),
the CPU oracle tests and the GPU parity tests (TEST INFRASTRUCTURE).

A case is described by seeds only; its inputs (keypoints, descriptors, FeatureVectors, MapPoint masks,
two-view geometry) are rebuilt from the synthetic frame generator by running an extractor.  Which
extractor is a parameter: the generator uses the cv2-backed Python restatement, the tests use the C++
oracle (proven identical to it by tests/test_oracle_golden.py), the GPU tests the CUDA extractor.
"""
import numpy as np

import orbfe
from orbfe import FeatureVectorCSR

ORB = (1.2, 8, 20, 7)
# EuRoC (Examples/Stereo/EuRoC.yaml:9-12,28), KITTI 00-02 (Examples/Stereo/KITTI00-02.yaml:10-13,29)
EUROC = dict(w=752, h=480, fx=435.2047, fy=435.2047, cx=367.4517, cy=252.2008, bf=47.9064)
KITTI = dict(w=1241, h=376, fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448)

STEREO_CASES = {
    # name: (camera, nfeatures, seed)
    "stereo_euroc": (EUROC, 1200, 1100),
    "stereo_euroc_b": (EUROC, 1200, 1101),
    "stereo_kitti": (KITTI, 2000, 1300),
}

BOW_CASES = {
    # name: (camera, nfeatures, seed, shift of the second view, check_ori, nleft, mask %)
    "bow_euroc": (EUROC, 1000, 1000, 6, True, -1, 60),
    "bow_euroc_noori": (EUROC, 1000, 1002, 9, False, -1, 60),
    "bow_euroc_nleft": (EUROC, 1000, 1003, 5, True, 500, 80),
    "bow_kitti": (KITTI, 2000, 1301, 7, True, -1, 60),
}

TRI_CASES = {
    # name: (camera, nfeatures, seed, shift, check_ori, only_stereo, coarse, epipole or None (= far away), existing-MapPoint %)
    "tri_euroc": (EUROC, 1000, 1005, 6, True, False, False, None, 40),
    "tri_euroc_epi_inside": (EUROC, 1000, 1006, 8, False, False, False, (400.0, 250.0), 40),
    "tri_euroc_stereo_only": (EUROC, 1000, 1007, 6, True, True, False, None, 30),
    "tri_euroc_coarse": (EUROC, 1000, 1008, 10, True, False, True, (300.0, 200.0), 40),
}

LAP_CASES = {
    # name: (w, h, nfeatures, seed, lapping left, lapping right)
    "lap_tumvi": (512, 512, 1500, 1200, (128, 511), (0, 383)),
}


def fv_of(desc):
    return FeatureVectorCSR.from_node_of_feature(orbfe.synth.node_of_feature(desc))


def stereo_inputs(name, extract):
    """extract(img, nfeatures, lapping) -> (mono, kps, desc, handle-for-pyramids)."""
    cam, nf, seed = STEREO_CASES[name]
    imL, imR = orbfe.synth.stereo_pair(seed, cam["w"], cam["h"])
    L = extract(imL, nf, (0, 0))
    R = extract(imR, nf, (0, 0))
    bf = np.float32(cam["bf"])
    b = np.float32(bf / np.float32(cam["fx"]))       # mb = mbf / fx (Frame.cc:151)
    return dict(L=L, R=R, bf=float(bf), b=float(b), imL=imL, imR=imR, nf=nf, cam=cam)


def bow_inputs(name, extract):
    cam, nf, seed, shift, ori, nleft, pct = BOW_CASES[name]
    im1 = orbfe.synth.frame(seed, cam["w"], cam["h"], 0, 0)
    im2 = orbfe.synth.frame(seed, cam["w"], cam["h"], 0, shift)
    _, k1, d1 = extract(im1, nf, (0, 0))[:3]
    _, k2, d2 = extract(im2, nf, (0, 0))[:3]
    return dict(k1=k1, d1=d1, fv1=fv_of(d1), mp1=orbfe.synth.mask(seed * 3 + 1, len(d1), pct), k2=k2, d2=d2, fv2=fv_of(d2),
                mp2=orbfe.synth.mask(seed * 3 + 2, len(d2), pct), check_ori=ori, nleft=nleft)


def tri_inputs(name, extract, scale, sigma2):
    cam, nf, seed, shift, ori, only_stereo, coarse, ep, pct = TRI_CASES[name]
    im1 = orbfe.synth.frame(seed, cam["w"], cam["h"], 0, 0)
    im2 = orbfe.synth.frame(seed, cam["w"], cam["h"], 0, shift)
    _, k1, d1 = extract(im1, nf, (0, 0))[:3]
    _, k2, d2 = extract(im2, nf, (0, 0))[:3]
    # a pure x translation: image 2 content sits `shift` px to the left, epipolar lines are the rows
    F12 = orbfe.synth.fundamental_x_shift(cam["fx"], cam["fy"], cam["cx"], cam["cy"], 0.1)
    ur1 = np.where(orbfe.synth.mask(seed * 5 + 1, len(d1), 50) > 0, k1["x"] - 3.0, -1.0).astype(np.float32)
    ur2 = np.where(orbfe.synth.mask(seed * 5 + 2, len(d2), 50) > 0, k2["x"] - 3.0, -1.0).astype(np.float32)
    return dict(k1=k1, d1=d1, fv1=fv_of(d1), mp1=orbfe.synth.mask(seed * 7 + 1, len(d1), pct), ur1=ur1, k2=k2, d2=d2,
                fv2=fv_of(d2), mp2=orbfe.synth.mask(seed * 7 + 2, len(d2), pct), ur2=ur2, F12=F12,
                ep=ep if ep is not None else (1e9, 1e9), check_ori=ori, only_stereo=only_stereo, coarse=coarse,
                scale=np.asarray(scale, np.float32), sigma2=np.asarray(sigma2, np.float32))


def lap_inputs(name, extract):
    w, h, nf, seed, lapL, lapR = LAP_CASES[name]
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    monoL, kL, dL = extract(imL, nf, lapL)[:3]
    monoR, kR, dR = extract(imR, nf, lapR)[:3]
    return dict(q=dL[monoL:], t=dR[monoR:], monoL=monoL, monoR=monoR)


def knn_inputs(seed=5, nq=700, nt=3000):
    q = orbfe.synth.descriptors(seed, nq)
    t = orbfe.synth.descriptors(seed + 1, nt)
    t[100] = t[7]
    t[2900] = t[7]
    q[0] = t[7]
    q[1] = t[2999]
    t[1500:1510] = q[5]          # ten-way tie at distance 0
    return q, t


# TUM-VI 512 calibration (Examples/Stereo/TUM_512.yaml:9-46): KB8 intrinsics of both cameras and Tlr
KB8_LEFT = np.array([190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504, 0.0034823894022493434,
                     0.0007150348452162257, -0.0020532361418706202, 0.00020293673591811182], np.float32)
KB8_RIGHT = np.array([190.44236969414825, 190.4344384721956, 252.59949716835982, 254.91723064636983, 0.0034003170790442797,
                      0.001766278153469831, -0.00266312569781606, 0.0003299517423931039], np.float32)
TLR = np.array([[0.999999445773493, 0.000791687752817, 0.000694034010224, 0.101063427414194],
                [-0.000823363992158, 0.998899461915674, 0.046895490788700, 0.001946204678584],
                [-0.000656143613644, -0.046896036240590, 0.998899560146304, 0.001015350132563]], np.float32)

FISHEYE_CASES = {"fisheye_tumvi": (2100, 900), "fisheye_tumvi_b": (2101, 1500)}   # name: (seed, stereo points)


def _kb8_project64(cam, P):
    cam = cam.astype(np.float64)
    th = np.arctan2(np.hypot(P[:, 0], P[:, 1]), P[:, 2])
    psi = np.arctan2(P[:, 1], P[:, 0])
    r = th + cam[4] * th ** 3 + cam[5] * th ** 5 + cam[6] * th ** 7 + cam[7] * th ** 9
    return np.stack([cam[0] * r * np.cos(psi) + cam[2], cam[1] * r * np.sin(psi) + cam[3]], axis=1)


def fisheye_inputs(name):
    """Synthetic fisheye stereo frame: 3-D points seen by both KB8 cameras (pixel noise, a few gross outliers,
    depths from 0.4 m to 12 m so that the parallax test fires), descriptors of the right view = left +
    bit flips, plus unmatched mono features in front of both lists.  Coordinates are snapped to 1/16 px
    so the float32 inputs are identical on every machine."""
    seed, n = FISHEYE_CASES[name]
    rng = np.random.default_rng(seed)
    nlev = 8
    P1 = np.stack([rng.uniform(-2.5, 2.5, n), rng.uniform(-1.8, 1.8, n), rng.uniform(0.4, 12.0, n)], axis=1)
    R12, t12 = TLR[:, :3].astype(np.float64), TLR[:, 3].astype(np.float64)
    P2 = (P1 - t12) @ R12            # R12^T (P1 - t12)
    uvL = _kb8_project64(KB8_LEFT, P1) + rng.normal(0, 0.4, (n, 2))
    uvR = _kb8_project64(KB8_RIGHT, P2) + rng.normal(0, 0.4, (n, 2))
    bad = rng.random(n) < 0.08
    uvR[bad] += rng.uniform(-12, 12, (int(bad.sum()), 2))
    mono_l, mono_r = 150, 210
    nL, nR = n + mono_l, n + mono_r
    kL, kR = np.zeros(nL, orbfe.KEYPOINT_DTYPE), np.zeros(nR, orbfe.KEYPOINT_DTYPE)
    perm = rng.permutation(n)        # right list order differs from the left one
    snap = lambda a: (np.round(a * 16) / 16).astype(np.float32)
    kL["x"][mono_l:], kL["y"][mono_l:] = snap(uvL[:, 0]), snap(uvL[:, 1])
    kR["x"][mono_r:][perm], kR["y"][mono_r:][perm] = snap(uvR[:, 0]), snap(uvR[:, 1])
    kL["x"][:mono_l], kL["y"][:mono_l] = snap(rng.uniform(20, 120, mono_l)), snap(rng.uniform(20, 490, mono_l))
    kR["x"][:mono_r], kR["y"][:mono_r] = snap(rng.uniform(390, 490, mono_r)), snap(rng.uniform(20, 490, mono_r))
    kL["octave"], kR["octave"] = rng.integers(0, nlev, nL), rng.integers(0, nlev, nR)
    kL["class_id"] = kR["class_id"] = -1
    dL = rng.integers(0, 256, (nL, 32), dtype=np.uint8)
    dR = rng.integers(0, 256, (nR, 32), dtype=np.uint8)
    flips = rng.integers(0, 60, n)
    src = dL[mono_l:].copy()
    for i in range(n):
        bits = rng.choice(256, int(flips[i]), replace=False)
        for b in bits:
            src[i, b >> 3] ^= 1 << (b & 7)
    dR[mono_r:][perm] = src
    return dict(cam1=KB8_LEFT, cam2=KB8_RIGHT, R12=TLR[:, :3].copy(), t12=TLR[:, 3].copy(), kL=kL, dL=dL, kR=kR, dR=dR,
                mono_l=mono_l, mono_r=mono_r, nlevels=nlev)


# ---- SearchForTriangulation with Kannala-Brandt cameras --------------------------------------------------
TRI_KB8_CASES = {
    # name: (seed, points, dual rig, check_ori, coarse, translation of keyframe 2 in keyframe 1's left camera)
    "tri_kb8_dual": (2200, 900, True, True, False, (0.22, 0.03, 0.05)),
    "tri_kb8_mono": (2201, 1400, False, False, False, (0.06, 0.03, -0.30)),     # epipole inside the image
    "tri_kb8_coarse": (2202, 700, True, True, True, (0.18, -0.02, 0.02)),
}


# more keyframe pairs for the runs against the reference's matcher + camera classes (no golden files: reference vs oracle vs GPU)
TRI_KB8_EXTRA = {
    "tri_kb8_dual_sideways": (2210, 1200, True, True, False, (-0.30, 0.05, 0.02)),
    "tri_kb8_mono_forward": (2211, 1000, False, True, False, (0.02, -0.01, 0.35)),
    "tri_kb8_dual_no_ori": (2212, 800, True, False, False, (0.12, 0.10, -0.08)),
    "tri_kb8_mono_coarse": (2213, 900, False, True, True, (0.25, 0.0, 0.0)),
}


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def tri_kb8_inputs(name):
    """Two keyframes of a (stereo-)fisheye rig looking at the same 3-D points.  Feature lists are
    [left-camera features..., right-camera features...] (KeyFrame::NLeft splits them); every observation of
    a point carries the point's descriptor with a few bits flipped outside bytes 0-1 (which define the
    stand-in vocabulary node), pixel noise, and 6 % gross outliers so the geometric test rejects some."""
    seed, n, dual, ori, coarse, trans = (TRI_KB8_CASES.get(name) or TRI_KB8_EXTRA[name])
    rng = np.random.default_rng(seed)
    nlev = 8
    snap = lambda a: (np.round(a * 16) / 16).astype(np.float32)
    Pw = np.stack([rng.uniform(-2.5, 2.5, n), rng.uniform(-1.8, 1.8, n), rng.uniform(0.6, 9.0, n)], axis=1)
    base = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    ang = rng.uniform(0, 360, n)
    Tlr = np.eye(4)
    Tlr[:3] = TLR.astype(np.float64)
    T_w_l1 = np.eye(4)
    T_w_l2 = np.eye(4)
    T_w_l2[:3, :3] = _rot(0.02, -0.03, 0.015)
    T_w_l2[:3, 3] = trans
    cams = [KB8_LEFT, KB8_RIGHT]

    def keyframe(T_w_l, salt):
        r = np.random.default_rng(seed * 10 + salt)
        ks, ds = [], []
        nleft = -1
        for ci in range(2 if dual else 1):
            T_w_c = T_w_l @ Tlr if ci else T_w_l
            T_c_w = np.linalg.inv(T_w_c)
            see = np.nonzero(r.random(n) < 0.7)[0]
            Pc = Pw[see] @ T_c_w[:3, :3].T + T_c_w[:3, 3]
            uv = _kb8_project64(cams[ci], Pc) + r.normal(0, 0.35, (len(see), 2))
            bad = r.random(len(see)) < 0.06
            uv[bad] += r.uniform(-10, 10, (int(bad.sum()), 2))
            inside = (uv[:, 0] > 20) & (uv[:, 0] < 492) & (uv[:, 1] > 20) & (uv[:, 1] < 492) & (Pc[:, 2] > 0.1)
            see, uv = see[inside], uv[inside]
            nz = 80                                     # features of no common point
            k = np.zeros(len(see) + nz, orbfe.KEYPOINT_DTYPE)
            k["x"][:len(see)], k["y"][:len(see)] = snap(uv[:, 0]), snap(uv[:, 1])
            k["x"][len(see):], k["y"][len(see):] = snap(r.uniform(20, 490, nz)), snap(r.uniform(20, 490, nz))
            k["angle"][:len(see)] = ((ang[see] + r.normal(0, 4, len(see))) % 360).astype(np.float32)
            k["angle"][len(see):] = r.uniform(0, 360, nz).astype(np.float32)
            k["octave"] = r.integers(0, nlev, len(k))
            k["class_id"] = -1
            d = np.empty((len(k), 32), np.uint8)
            d[:len(see)] = base[see]
            d[len(see):] = r.integers(0, 256, (nz, 32), dtype=np.uint8)
            for i in range(len(see)):
                for b in r.choice(np.arange(16, 256), int(r.integers(0, 40)), replace=False):
                    d[i, b >> 3] ^= 1 << (b & 7)
            perm = r.permutation(len(k))
            ks.append(k[perm]); ds.append(d[perm])
            if ci == 0 and dual:
                nleft = len(k)
        return np.concatenate(ks), np.concatenate(ds), nleft

    k1, d1, nleft1 = keyframe(T_w_l1, 1)
    k2, d2, nleft2 = keyframe(T_w_l2, 2)
    R12, t12 = np.zeros((4, 9), np.float32), np.zeros((4, 3), np.float32)
    for c in range(4):                                   # ll, lr, rl, rr (ORBmatcher.cc:1248-1262)
        T1 = T_w_l1 @ Tlr if c >> 1 else T_w_l1
        T2 = T_w_l2 @ Tlr if c & 1 else T_w_l2
        T12 = np.linalg.inv(T1) @ T2                     # x1 = R12 x2 + t12
        R12[c], t12[c] = T12[:3, :3].reshape(9), T12[:3, 3]
    T21 = np.linalg.inv(np.linalg.inv(T_w_l1) @ T_w_l2)
    ep = _kb8_project64(KB8_LEFT, T21[:3, 3][None, :])[0]          # camera-1 centre seen by camera 2 (:1213-1218)
    cam = np.stack([KB8_LEFT, KB8_RIGHT]).astype(np.float32)
    return dict(k1=k1, d1=d1, fv1=fv_of(d1), mp1=orbfe.synth.mask(seed * 7 + 1, len(d1), 40), k2=k2, d2=d2, fv2=fv_of(d2),
                mp2=orbfe.synth.mask(seed * 7 + 2, len(d2), 40), cam_a=cam, cam_b=cam, nleft_a=nleft1, nleft_b=nleft2, R12=R12,
                t12=t12, ep=(float(np.float32(ep[0])), float(np.float32(ep[1]))), check_ori=ori, only_stereo=False,
                coarse=coarse, nlevels=nlev,
                # the same geometry as keyframe poses (what the reference's SearchForTriangulation_ starts from): Tcw of the
                # two left cameras and KeyFrame::mTrl (right camera from left)
                Tcw1=np.linalg.inv(T_w_l1).astype(np.float32), Tcw2=np.linalg.inv(T_w_l2).astype(np.float32),
                Trl=np.linalg.inv(Tlr).astype(np.float32))


def distinctive_inputs(seed=5200, npoints=400):
    """Observations of `npoints` map points: 0..70 descriptors each (noisy copies of a base descriptor, a few
    outliers, exact duplicates so that medians tie).  Returns (desc (total, 32), offsets (npoints + 1))."""
    rng = np.random.default_rng(seed)
    counts = rng.integers(0, 12, npoints)
    counts[rng.random(npoints) < 0.15] = rng.integers(12, 71, int((rng.random(npoints) < 0.15).sum()) or 1)[0]
    counts[:6] = [0, 1, 2, 3, 64, 33]
    rows = []
    for c in counts:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        for k in range(int(c)):
            d = base.copy()
            nb = int(rng.integers(0, 40)) if rng.random() < 0.85 else 128
            for b in rng.choice(256, nb, replace=False):
                d[b >> 3] ^= 1 << (b & 7)
            if k and rng.random() < 0.1:
                d = rows[-1].copy()
            rows.append(d)
    desc = np.stack(rows) if rows else np.zeros((0, 32), np.uint8)
    return desc, np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
