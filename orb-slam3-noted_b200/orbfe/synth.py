"""Deterministic synthetic inputs (csrc/synth.c): textured frames, stereo pairs, random descriptors."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(os.path.dirname(_HERE), "csrc", "liborbfe_synth.so")
_lib = None


def _l():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            raise ImportError(f"{_PATH} is missing: run __graft_entry__.build()")
        L = C.CDLL(_PATH)
        L.orbfe_synth_frame.restype = C.c_int
        L.orbfe_synth_frame.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orbfe_synth_disparity.restype = C.c_int
        L.orbfe_synth_disparity.argtypes = [C.c_int, C.c_int]
        L.orbfe_synth_descriptors.restype = None
        L.orbfe_synth_descriptors.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def frame(seed, w, h, view=0, shift=0):
    out = np.empty((h, w), np.uint8)
    if _l().orbfe_synth_frame(int(seed), w, h, int(view), int(shift), out.ctypes.data) != 0:
        raise ValueError("orbfe_synth_frame: bad arguments")
    return out


def frames(seed0, n, w, h):
    return np.stack([frame(seed0 + i, w, h) for i in range(n)])


def stereo_pair(seed, w, h):
    return frame(seed, w, h, 0), frame(seed, w, h, 1)


def disparity(y, h):
    return _l().orbfe_synth_disparity(int(y), int(h))


def descriptors(seed, n):
    out = np.empty((n, 32), np.uint8)
    _l().orbfe_synth_descriptors(int(seed), n, out.ctypes.data)
    return out


# ---- synthetic matcher inputs: the vocabulary file is not available, so the node of a
# feature is a deterministic function of its descriptor; MapPoint masks and two-view geometry are seeded.
def node_of_feature(desc):
    """<= 100 level-2-like node ids; descriptors that agree in their first two bytes mod 10 share a node."""
    desc = np.asarray(desc, np.uint8).reshape(-1, 32)
    return (1 + 10 * (desc[:, 0].astype(np.int64) % 10) + (desc[:, 1].astype(np.int64) % 10)).astype(np.uint32)


def _splitmix(seed, n):
    with np.errstate(over="ignore"):
        x = np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def mask(seed, n, percent):
    """n bytes, 1 with probability percent/100 (integer hash, machine independent)."""
    return (_splitmix(seed, n) % np.uint64(100) < np.uint64(percent)).astype(np.uint8)


def fundamental_x_shift(fx, fy, cx, cy, tx):
    """F12 = K^-T [t]x R K^-1 (Pinhole.cpp:161-164) for R = I, t12 = (tx, 0, 0), in float32 like cv::Matx33f."""
    f = np.float32
    Kinv = np.array([[f(1) / f(fx), 0, -f(cx) / f(fx)], [0, f(1) / f(fy), -f(cy) / f(fy)], [0, 0, 1]], np.float32)
    tx_ = np.array([[0, 0, 0], [0, 0, -tx], [0, tx, 0]], np.float32)
    return (Kinv.T @ tx_ @ Kinv).astype(np.float32)


def vocabulary(seed, k=10, L=6, stop_pct=1):
    """A complete random k-ary vocabulary tree of depth L in breadth-first node order (ORBvoc.txt is not shipped
    with the reference; k = 10, L = 6 gives its size: 1.1 M nodes, 10^6 words).  Returns the flattened arrays
    the ABI takes: (parent, is_leaf, descriptors, weights)."""
    rng = np.random.default_rng(seed)
    n = (k ** (L + 1) - 1) // (k - 1)
    parent = np.zeros(n, np.int32)
    start = 1
    prev_start = 0
    for lev in range(1, L + 1):
        cnt = k ** lev
        parent[start:start + cnt] = prev_start + np.arange(cnt) // k
        prev_start, start = start, start + cnt
    leaf = np.zeros(n, np.uint8)
    leaf[prev_start:] = 1
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    weight = np.where(leaf > 0, rng.uniform(0.5, 9.0, n), 0.0)
    weight[(rng.random(n) < stop_pct / 100.0) & (leaf > 0)] = 0.0
    return parent, leaf, desc, weight
