#!/usr/bin/env python
"""Drives the matcher kernels that the extraction bench does not reach (k_stereo_search / k_stereo_filter, k_bow_nodes,
k_tri_nodes, k_match_finish) at BASELINE shapes, for `ncu -k regex:...` (profiles/r02*_matchers_*):
    ncu --set full --clock-control none --import-source on -k regex:'k_stereo|k_bow_nodes|k_tri_nodes|k_match_finish' -o gpurun_out/m python tools/profile_matchers.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "orb-slam3-noted_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import match_cases as mc  # noqa: E402
import orbfe  # noqa: E402


def main():
    # config 1: 64 EuRoC stereo pairs, batched
    cam, W, H, P = mc.EUROC, 752, 480, 64
    frames = np.stack([orbfe.synth.stereo_pair(1100 + p % 8, W, H)[e] for p in range(P) for e in (0, 1)])
    ex = orbfe.ORBextractor(1200, *mc.ORB, max_width=W, max_height=H, max_batch=2 * P)
    m = orbfe.ORBmatcher(0.6, True)
    dev = torch.from_numpy(frames).cuda()
    ur = torch.empty((P, ex.cap), dtype=torch.float32, device="cuda")
    dp = torch.empty((P, ex.cap), dtype=torch.float32, device="cuda")
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    for _ in range(2):
        ex.extract_batch_device(dev.data_ptr(), 2 * P, W, H)
        m.stereo_match_batch_device(ex, P, bf, b, ur.data_ptr(), dp.data_ptr())
    m.synchronize()
    # config 3: one KITTI frame against 50 keyframes
    cam = mc.KITTI
    W, H = cam["w"], cam["h"]
    kf = np.stack([orbfe.synth.frame(1300 + s // 25, W, H, 0, s % 25) for s in range(51)])
    ex3 = orbfe.ORBextractor(2000, *mc.ORB, max_width=W, max_height=H, max_batch=51)
    res = ex3.extract_batch(kf, (0, 0))
    _, fk, fd = res[0]
    kfs = [(d, mc.fv_of(d), orbfe.synth.mask(3900 + i, len(d), 60), k) for i, (_, k, d) in enumerate(res[1:])]
    m7 = orbfe.ORBmatcher(0.7, True)
    for _ in range(2):
        m7.SearchByBoW_KFs_F(kfs, fd, mc.fv_of(fd), fk)
    # SearchForTriangulation_ (pinhole) between the frame and a keyframe
    sc, s2 = ex3.GetScaleFactors(), ex3.GetScaleSigmaSquares()
    F12 = orbfe.synth.fundamental_x_shift(cam["fx"], cam["fy"], cam["cx"], cam["cy"], 0.1)
    _, k1, d1 = res[0]
    _, k2, d2 = res[6]
    ur1 = np.full(len(d1), -1, np.float32)
    ur2 = np.full(len(d2), -1, np.float32)
    m6 = orbfe.ORBmatcher(0.6, True)
    for _ in range(2):
        m6.SearchForTriangulation(d1, mc.fv_of(d1), orbfe.synth.mask(1, len(d1), 40), k1, ur1, d2, mc.fv_of(d2), orbfe.synth.mask(2, len(d2), 40),
                                  k2, ur2, F12, (1e9, 1e9), s2, sc, False, False)
    print("ok")


if __name__ == "__main__":
    main()
