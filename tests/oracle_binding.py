"""ctypes view of the CPU oracle (oracle/_build/liborbref.so).  TEST INFRASTRUCTURE: the checker only."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_build", "liborbref.so")

KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                           ("octave", "<i4"), ("class_id", "<i4")])
_lib = None
_P, _I, _F, _L = C.c_void_p, C.c_int, C.c_float, C.c_long


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(PATH)
        sig = {
            "orbref_create": (_P, [_I, _F, _I, _I, _I]),
            "orbref_destroy": (None, [_P]),
            "orbref_set_tie_rule": (None, [_P, _I]),
            "orbref_tables": (None, [_P] * 7),
            "orbref_extract": (_I, [_P, _P, _I, _I, _L, _I, _I, _P, _P, _I, C.POINTER(_I)]),
            "orbref_level_size": (_I, [_P, _I, C.POINTER(_I), C.POINTER(_I)]),
            "orbref_get_level": (_I, [_P, _I, _P]),
            "orbref_get_candidates": (_I, [_P, _I, _P, _I]),
            "orbref_resize": (None, [_P, _I, _I, _L, _P, _I, _I, _L]),
            "orbref_border101": (None, [_P, _I, _I, _I, _L]),
            "orbref_fast": (_I, [_P, _I, _I, _L, _I, _P, _I]),
            "orbref_blur": (None, [_P, _I, _I, _L, _P, _L]),
            "orbref_fast_atan2": (_F, [_F, _F]),
            "orbref_knn2": (None, [_P, _I, _P, _I, _P, _P]),
            "orbref_hamming": (_I, [_P, _P]),
            "orbref_extract_many": (_L, [_I, _F, _I, _I, _I, _P, _I, _I, _I, _I, _I, _I]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(_P)


class OrbRef:
    """CPU oracle of ORBextractor (oracle/cpu_ref/orb_ref.cc)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST):
        self.L = lib()
        self.h = self.L.orbref_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self.nlevels = nlevels
        self.cap = nfeatures + 64 * nlevels + 4096

    def __del__(self):
        try:
            self.L.orbref_destroy(self.h)
        except Exception:
            pass

    def tables(self):
        nl = self.nlevels
        sc, isc, s2, is2 = (np.zeros(nl, np.float32) for _ in range(4))
        per = np.zeros(nl, np.int32)
        umax = np.zeros(16, np.int32)
        self.L.orbref_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(per), _p(umax))
        return sc, isc, s2, is2, per, umax

    def __call__(self, image, lapping=(0, 0)):
        image = np.ascontiguousarray(image)
        h, w = image.shape
        kps = np.zeros(self.cap, KEYPOINT_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self.L.orbref_extract(self.h, _p(image), w, h, w, int(lapping[0]), int(lapping[1]), _p(kps), _p(desc),
                                     self.cap, C.byref(n))
        assert mono >= -1, "oracle capacity"
        if mono < 0:
            return -1, kps[:0], desc[:0]
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        assert self.L.orbref_level_size(self.h, l, C.byref(w), C.byref(h)) == 0
        buf = np.zeros((h.value + 38, w.value + 38), np.uint8)
        self.L.orbref_get_level(self.h, l, _p(buf))
        return buf

    def candidates(self, l):
        n = self.L.orbref_get_candidates(self.h, l, None, 0)
        out = np.zeros((max(n, 1), 3), np.float32)
        self.L.orbref_get_candidates(self.h, l, _p(out), out.shape[0])
        return out[:n].astype(np.int32)


def knn2(q, t):
    q, t = np.ascontiguousarray(q), np.ascontiguousarray(t)
    idx = np.zeros((len(q), 2), np.int32)
    dist = np.zeros((len(q), 2), np.int32)
    lib().orbref_knn2(_p(q), len(q), _p(t), len(t), _p(idx), _p(dist))
    return idx, dist


def extract_many(params, frames, lap, nthreads):
    frames = np.ascontiguousarray(frames)
    n, h, w = frames.shape
    nf, sf, nl, ini, mn = params
    return lib().orbref_extract_many(nf, sf, nl, ini, mn, _p(frames), n, w, h, int(lap[0]), int(lap[1]), nthreads)
