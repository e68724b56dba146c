#!/usr/bin/env python
"""Generates tests/golden/match_*.npz with the Python restatement of the matchers (oracle/py/match_py.py,
cv2 for BFMatcher / norm) on inputs produced by the cv2-backed extractor (oracle/py/orb_cv2.py).

TEST INFRASTRUCTURE.  Run in the build container:   python oracle/py/make_golden_match.py

The cases are defined by seeds in tests/match_cases.py; a fixture stores only the expected outputs and
a CRC of the inputs (so that a drifting generator is noticed), never the inputs themselves.
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [HERE, os.path.join(ROOT, "orb-slam3-noted_b200"), os.path.join(ROOT, "tests")]

import match_cases as mc  # noqa: E402
import match_py as mp  # noqa: E402
import orbfe  # noqa: E402
from orb_cv2 import ORBextractorCv2  # noqa: E402

KP = orbfe.KEYPOINT_DTYPE


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return np.uint32(c & 0xFFFFFFFF)


def extract(img, nf, lap):
    ex = ORBextractorCv2(nf, *mc.ORB)
    kp, desc, mono = ex(img, lap)
    k = np.zeros(len(kp), KP)
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        k[f] = kp[:, i].astype(np.float32)
    k["octave"] = kp[:, 5].astype(np.int32)
    k["class_id"] = -1
    return mono, k, desc, ex


def main():
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    tab = ORBextractorCv2(1000, *mc.ORB)
    scale, inv_scale, sigma2 = tab.mvScaleFactor, tab.mvInvScaleFactor, tab.mvLevelSigma2

    for name in mc.STEREO_CASES:
        c = mc.stereo_inputs(name, extract)
        (_, kL, dL, eL), (_, kR, dR, eR) = c["L"], c["R"]
        u, d, sad = mp.compute_stereo_matches(eL.pyramid, eR.pyramid, eL.level_size, scale, inv_scale, kL, dL, kR, dR, c["bf"],
                                              c["b"])
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(kL, dL, kR, dR), u_right=u, depth=d, sad=sad)
        print(name, int((d > 0).sum()), "of", len(kL), "left keypoints have depth")

    for name in mc.BOW_CASES:
        c = mc.bow_inputs(name, extract)
        fv1, fv2 = c["fv1"].to_dict(), c["fv2"].to_dict()
        m_kf_f, n_kf_f = mp.search_by_bow_kf_frame(c["d1"], fv1, c["mp1"], c["k1"]["angle"], c["d2"], fv2, c["k2"]["angle"], 0.7,
                                                   c["check_ori"], c["nleft"])
        m_kk, n_kk = mp.search_by_bow_kf_kf(c["d1"], fv1, c["mp1"], c["k1"]["angle"], c["d2"], fv2, c["mp2"], c["k2"]["angle"],
                                            0.8, c["check_ori"])
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(c["d1"], c["d2"], c["mp1"], c["mp2"]),
                            kf_f=m_kf_f, n_kf_f=n_kf_f, kf_kf=m_kk, n_kf_kf=n_kk)
        print(name, "KF-F", n_kf_f, "KF-KF", n_kk)

    for name in mc.TRI_CASES:
        c = mc.tri_inputs(name, extract, scale, sigma2)
        m12, n = mp.search_for_triangulation(c["k1"], c["d1"], c["fv1"].to_dict(), c["mp1"], c["ur1"], c["k2"], c["d2"],
                                             c["fv2"].to_dict(), c["mp2"], c["ur2"], sigma2, scale, c["F12"], c["ep"],
                                             c["check_ori"], c["only_stereo"], c["coarse"])
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(c["d1"], c["d2"], c["mp1"], c["mp2"], c["F12"]),
                            m12=m12, n=n)
        print(name, n, "pairs")

    for name in mc.LAP_CASES:
        c = mc.lap_inputs(name, extract)
        match, idx, dist = mp.fisheye_lapping_match(c["q"], c["t"])
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(c["q"], c["t"]), match=match, idx=idx, dist=dist)
        print(name, len(c["q"]), "x", len(c["t"]), "->", int((match >= 0).sum()), "matches")

    for name in mc.FISHEYE_CASES:
        c = mc.fisheye_inputs(name)
        l2r, r2l, depth, p3d, margins, n = mp.compute_stereo_fisheye_matches(c["cam1"], c["cam2"], c["kL"], c["dL"], c["mono_l"],
                                                                            c["kR"], c["dR"], c["mono_r"], c["R12"], c["t12"],
                                                                            sigma2)
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(c["kL"], c["dL"], c["kR"], c["dR"]), l2r=l2r,
                            r2l=r2l, depth=depth, p3d=p3d, margins=margins, n=n)
        tried = int((margins < 1e29).sum())
        print(name, tried, "descriptor matches ->", n, "triangulated; knife-edge (margin < 1e-3):", int((margins < 1e-3).sum()))

    for name in mc.TRI_KB8_CASES:
        c = mc.tri_kb8_inputs(name)
        m12, n, pts, margins = mp.search_for_triangulation_kb8(c["k1"], c["d1"], c["fv1"].to_dict(), c["mp1"], sigma2, c["k2"],
                                                               c["d2"], c["fv2"].to_dict(), c["mp2"], sigma2, scale, c["cam_a"],
                                                               c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"],
                                                               c["check_ori"], c["only_stereo"], c["coarse"])
        np.savez_compressed(os.path.join(out, f"match_{name}.npz"), in_crc=crc(c["k1"], c["d1"], c["k2"], c["d2"], c["R12"]),
                            m12=m12, n=n, points=pts, margins=margins)
        print(name, n, "pairs; knife-edge rows:", int((margins < 1e-3).sum()))

    d, off = mc.distinctive_inputs()
    best = np.array([mp.compute_distinctive_descriptor(d[off[i]:off[i + 1]]) for i in range(len(off) - 1)], np.int32)
    np.savez_compressed(os.path.join(out, "match_distinctive.npz"), in_crc=crc(d, off), best=best)
    print("distinctive", len(best), "map points,", len(d), "observations")

    q, t = mc.knn_inputs()
    match, idx, dist = mp.fisheye_lapping_match(q, t)
    np.savez_compressed(os.path.join(out, "match_knn_ties.npz"), in_crc=crc(q, t), match=match, idx=idx, dist=dist)
    print("knn_ties", int((match >= 0).sum()), "matches")


if __name__ == "__main__":
    main()
