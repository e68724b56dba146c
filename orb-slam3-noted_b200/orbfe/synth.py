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
