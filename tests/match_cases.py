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
