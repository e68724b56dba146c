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
