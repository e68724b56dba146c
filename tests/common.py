"""Shared helpers of the parity tests (TEST INFRASTRUCTURE)."""
import glob
import os
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
ANGLE_TOL_DEG = 1e-4 * 180.0 / np.pi   # north_star: keypoint angles within 1e-4 rad
DESC_IDENTICAL_MIN = 0.9999            # north_star: >= 99.99 % of descriptors bit-identical


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def low_texture(img):
    """Left half at 1/8 contrast and a fully flat block: forces the minThFAST retry (ORBextractor.cc:825-828)."""
    h, w = img.shape
    img = img.copy()
    img[:, : w // 2] = (img[:, : w // 2].astype(np.int32) // 8 + 112).astype(np.uint8)
    img[: h // 3, : w // 4] = 90
    return img


def golden_cases(prefix="extract_"):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def golden_input(g):
    import orbfe
    img = orbfe.synth.frame(int(g["seed"]), int(g["w"]), int(g["h"]))
    if bool(g["low_texture"]):
        img = low_texture(img)
    assert crc(img) == int(g["image_crc"]), "synthetic generator drifted from the committed fixture"
    return img


def check_against_golden(g, mono, kps, desc, pyramid_level=None, candidates=None):
    """Compares one extraction result with a golden fixture.  pyramid_level(l) -> bordered level,
    candidates(l) -> (n, 3) int32 in reference push order; both optional."""
    nl = len(g["pyramid_crc"])
    if pyramid_level is not None:
        for l in range(nl):
            assert crc(pyramid_level(l)) == int(g["pyramid_crc"][l]), f"pyramid level {l}"
    if candidates is not None:
        for l in range(nl):
            c = candidates(l)
            assert len(c) == int(g["cand_n"][l]), f"candidate count level {l}"
            assert crc(c.astype(np.int32)) == int(g["cand_crc"][l]), f"candidates level {l}"
    assert mono == int(g["mono"])
    assert len(kps) == len(g["x"])
    for f in ("x", "y", "size", "response", "octave"):
        assert np.array_equal(kps[f], g[f]), f
    assert np.all(kps["class_id"] == -1)
    assert np.max(np.abs(kps["angle"] - g["angle"]), initial=0) <= ANGLE_TOL_DEG
    same = np.all(desc == g["desc"], axis=1)
    return int(same.sum()), len(same)
