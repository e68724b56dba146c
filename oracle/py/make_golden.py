#!/usr/bin/env python
"""Generates tests/golden/extract_*.npz with the cv2-backed restatement (oracle/py/orb_cv2.py).

TEST INFRASTRUCTURE.  Run in the build container (cv2 4.13.0 is the pinned OpenCV):

    python oracle/py/make_golden.py

One fixture per BASELINE.json shape (one seed each, plus a low-texture 752x480 frame that forces the
minThFAST retry).  A fixture holds everything the parity tests compare, small enough to commit:
the generator arguments of the input frame (the frame itself is re-rendered by csrc/synth.c, its
CRC32 is stored), CRC32 of every bordered pyramid level, per-level FAST candidate count + CRC32 of
the (x, y, response) int32 triples in reference push order, the selected keypoints (cv::KeyPoint
fields), monoIndex and the descriptors.  The GPU box has no /root/reference and needs no cv2: the
tests there compare against these files.
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [HERE, os.path.join(ROOT, "orb-slam3-noted_b200")]

from orb_cv2 import ORBextractorCv2  # noqa: E402

# name, seed, w, h, nfeatures, lapping, low_texture
CASES = [
    ("c1_euroc_mono", 1000, 752, 480, 1000, (0, 1000), False),
    ("c1_euroc_mono_nolap", 1001, 752, 480, 1000, (0, 0), False),
    ("c2_euroc_stereo", 1100, 752, 480, 1200, (0, 0), False),
    ("c3_tumvi_fisheye", 1200, 512, 512, 1500, (128, 511), False),
    ("c4_kitti", 1300, 1241, 376, 2000, (0, 0), False),
    ("c5_hd", 1400, 1920, 1080, 4000, (0, 1000), False),
    ("lowtex_retry", 77, 752, 480, 1000, (0, 0), True),
]
ORB = (1.2, 8, 20, 7)


def make_input(seed, w, h, low_texture):
    import orbfe
    img = orbfe.synth.frame(seed, w, h)
    if low_texture:   # same transformation as tests/common.py:low_texture()
        img = img.copy()
        img[:, : w // 2] = (img[:, : w // 2].astype(np.int32) // 8 + 112).astype(np.uint8)
        img[: h // 3, : w // 4] = 90
    return img


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, seed, w, h, nf, lap, low in CASES:
        img = make_input(seed, w, h, low)
        ex = ORBextractorCv2(nf, *ORB)
        kp, desc, mono = ex(img, lap)
        nl = ex.nlevels
        pyr_crc = np.array([crc(ex.pyramid[l]) for l in range(nl)], np.uint32)
        lvl_size = np.array(ex.level_size, np.int32)
        cand_n = np.array([len(c) for c in ex.candidates], np.int32)
        cand_crc = np.array([crc(np.array(c, np.float64).reshape(-1, 3).astype(np.int32)) for c in ex.candidates], np.uint32)
        np.savez_compressed(
            os.path.join(out_dir, f"extract_{name}.npz"),
            seed=seed, w=w, h=h, nfeatures=nf, lapping=np.array(lap, np.int32), low_texture=low,
            orb=np.array(ORB, np.float64), image_crc=np.uint32(crc(img)), level_size=lvl_size, pyramid_crc=pyr_crc,
            cand_n=cand_n, cand_crc=cand_crc, mono=mono,
            x=kp[:, 0].astype(np.float32), y=kp[:, 1].astype(np.float32), size=kp[:, 2].astype(np.float32),
            angle=kp[:, 3].astype(np.float32), response=kp[:, 4].astype(np.float32), octave=kp[:, 5].astype(np.int32),
            desc=desc)
        print(name, len(kp), "keypoints, monoIndex", mono)


if __name__ == "__main__":
    main()
