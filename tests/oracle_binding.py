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
            "orbref_descriptor_distance": (_I, [_P, _P]),
            "orbref_three_maxima": (None, [_P, _I, _P]),
            "orbref_stereo_match": (None, [_P, _P, _P, _P, _I, _P, _P, _I, _F, _F, _P, _P, _P]),
            "orbref_search_bow": (_I, [_P, _I, _I, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P, _F, _I, _I, _I, _I,
                                       _P, _P]),
            "orbref_search_triangulation": (_I, [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P, _P,
                                                 _P, _P, _F, _F, _I, _I, _I, _P]),
            "orbref_knn2_ratio": (None, [_P, _I, _P, _I, C.c_double, _P]),
            "orbref_knn2_many": (_L, [_P, _I, _P, _I, _I, _L, _L, _I]),
            "orbref_kb8_project": (None, [_P, _P, _P]),
            "orbref_kb8_unproject": (None, [_P, _F, _F, _P]),
            "orbref_search_triangulation_kb8": (_I, [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P, _P,
                                                     _P, _P, _I, _I, _P, _P, _F, _F, _I, _I, _I, _P, _P, _P]),
            "orbref_fisheye_stereo": (_I, [_P, _P, _P, _P, _I, _I, _P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P]),
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


# ---- matchers (oracle/cpu_ref/match_ref.cc) ----------------------------------------------------------
def _pn(a):
    return None if a is None else a.ctypes.data_as(_P)


def descriptor_distance(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib().orbref_descriptor_distance(_p(a), _p(b))


def three_maxima(counts):
    c = np.ascontiguousarray(counts, np.int32)
    out = np.zeros(3, np.int32)
    lib().orbref_three_maxima(_p(c), len(c), _p(out))
    return tuple(int(v) for v in out)


def stereo_match(ref_left, ref_right, kpL, descL, kpR, descR, bf, b):
    """ref_left / ref_right: OrbRef objects that have just extracted the two views (their pyramids are read)."""
    kpL, kpR = np.ascontiguousarray(kpL), np.ascontiguousarray(kpR)
    descL, descR = np.ascontiguousarray(descL), np.ascontiguousarray(descR)
    u = np.zeros(len(kpL), np.float32)
    d = np.zeros(len(kpL), np.float32)
    sad = np.zeros(len(kpL), np.int32)
    lib().orbref_stereo_match(ref_left.h, ref_right.h, _p(kpL), _p(descL), len(kpL), _p(kpR), _p(descR), len(kpR), bf, b, _p(u),
                              _p(d), _p(sad))
    return u, d, sad


def _fv(fv):
    return len(fv.node_ids), _p(fv.node_ids), _p(fv.offsets), _p(fv.indices)


def search_bow(descA, fvA, maskA, angA, descB, fvB, maskB, angB, ratio, check_ori, strict, nleft=-1, th=50):
    descA, descB = np.ascontiguousarray(descA, np.uint8), np.ascontiguousarray(descB, np.uint8)
    maskA = np.ascontiguousarray(maskA).astype(np.uint8)
    maskB = None if maskB is None else np.ascontiguousarray(maskB).astype(np.uint8)
    angA, angB = np.ascontiguousarray(angA, np.float32), np.ascontiguousarray(angB, np.float32)
    a2b = np.zeros(max(len(descA), 1), np.int32)
    b2a = np.zeros(max(len(descB), 1), np.int32)
    n = lib().orbref_search_bow(_p(descA), len(descA), *_fv(fvA), _p(maskA), _p(angA), _p(descB), len(descB), *_fv(fvB),
                                _pn(maskB), _p(angB), ratio, int(check_ori), th, int(strict), nleft, _p(a2b), _p(b2a))
    return a2b[:len(descA)], b2a[:len(descB)], n


def search_triangulation(kpA, descA, fvA, mpA, urA, kpB, descB, fvB, mpB, urB, sigma2, scale, F12, ep, check_ori, only_stereo,
                         coarse):
    kpA, kpB = np.ascontiguousarray(kpA), np.ascontiguousarray(kpB)
    descA, descB = np.ascontiguousarray(descA, np.uint8), np.ascontiguousarray(descB, np.uint8)
    mpA, mpB = np.ascontiguousarray(mpA).astype(np.uint8), np.ascontiguousarray(mpB).astype(np.uint8)
    urA, urB = np.ascontiguousarray(urA, np.float32), np.ascontiguousarray(urB, np.float32)
    sigma2, scale = np.ascontiguousarray(sigma2, np.float32), np.ascontiguousarray(scale, np.float32)
    F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
    m12 = np.zeros(max(len(descA), 1), np.int32)
    n = lib().orbref_search_triangulation(_p(kpA), _p(descA), len(descA), *_fv(fvA), _p(mpA), _p(urA), _p(kpB), _p(descB),
                                          len(descB), *_fv(fvB), _p(mpB), _p(urB), _p(sigma2), _p(scale), _p(F12), float(ep[0]),
                                          float(ep[1]), int(check_ori), int(only_stereo), int(coarse), _p(m12))
    return m12[:len(descA)], n


def knn2_ratio(q, t, ratio=0.7):
    q, t = np.ascontiguousarray(q, np.uint8), np.ascontiguousarray(t, np.uint8)
    m = np.zeros(max(len(q), 1), np.int32)
    lib().orbref_knn2_ratio(_p(q), len(q), _p(t), len(t), ratio, _p(m))
    return m[:len(q)]


def knn2_many(q, t, npairs, qstride, tstride, nthreads):
    """q, t: arrays holding `npairs` problems (nq x 32, nt x 32 each) at the given byte strides."""
    return lib().orbref_knn2_many(_p(q), q.shape[-2], _p(t), t.shape[-2], npairs, qstride, tstride, nthreads)


def fisheye_stereo(cam1, cam2, kpL, descL, mono_left, kpR, descR, mono_right, R12, t12, sigma2):
    """Frame::ComputeStereoFishEyeMatches (oracle/cpu_ref/kb8_ref.cc).  Returns (l2r, r2l, depth, p3d, margins, n)."""
    cam1, cam2 = np.ascontiguousarray(cam1, np.float32), np.ascontiguousarray(cam2, np.float32)
    kpL, kpR = np.ascontiguousarray(kpL), np.ascontiguousarray(kpR)
    descL, descR = np.ascontiguousarray(descL, np.uint8), np.ascontiguousarray(descR, np.uint8)
    R12, t12 = np.ascontiguousarray(R12, np.float32).reshape(9), np.ascontiguousarray(t12, np.float32).reshape(3)
    sigma2 = np.ascontiguousarray(sigma2, np.float32)
    nL, nR = len(kpL), len(kpR)
    l2r, r2l = np.zeros(max(nL, 1), np.int32), np.zeros(max(nR, 1), np.int32)
    depth, p3d, mg = np.zeros(max(nL, 1), np.float32), np.zeros((max(nL, 1), 3), np.float32), np.zeros(max(nL, 1), np.float32)
    n = lib().orbref_fisheye_stereo(_p(cam1), _p(cam2), _p(kpL), _p(descL), nL, mono_left, _p(kpR), _p(descR), nR, mono_right,
                                    _p(R12), _p(t12), _p(sigma2), _p(l2r), _p(r2l), _p(depth), _p(p3d), _p(mg))
    return l2r[:nL], r2l[:nR], depth[:nL], p3d[:nL], mg[:nL], n


def search_triangulation_kb8(kpA, descA, fvA, mpA, sigma2A, kpB, descB, fvB, mpB, sigma2B, scaleB, camA, camB, nleftA, nleftB,
                             R12, t12, ep, check_ori, only_stereo, coarse):
    """SearchForTriangulation_ with Kannala-Brandt cameras.  Returns (m12, n, points (nA, 3), margins (nA))."""
    kpA, kpB = np.ascontiguousarray(kpA), np.ascontiguousarray(kpB)
    descA, descB = np.ascontiguousarray(descA, np.uint8), np.ascontiguousarray(descB, np.uint8)
    mpA, mpB = np.ascontiguousarray(mpA).astype(np.uint8), np.ascontiguousarray(mpB).astype(np.uint8)
    f = lambda a: np.ascontiguousarray(a, np.float32)
    sigma2A, sigma2B, scaleB, camA, camB, R12, t12 = map(f, (sigma2A, sigma2B, scaleB, camA, camB, R12, t12))
    assert camA.size == 16 and camB.size == 16 and R12.size == 36 and t12.size == 12
    nA = len(descA)
    m12, pts, mg = np.zeros(max(nA, 1), np.int32), np.zeros((max(nA, 1), 3), np.float32), np.zeros(max(nA, 1), np.float32)
    n = lib().orbref_search_triangulation_kb8(_p(kpA), _p(descA), nA, *_fv(fvA), _p(mpA), _p(sigma2A), _p(kpB), _p(descB),
                                              len(descB), *_fv(fvB), _p(mpB), _p(sigma2B), _p(scaleB), _p(camA), _p(camB),
                                              int(nleftA), int(nleftB), _p(R12), _p(t12), float(ep[0]), float(ep[1]),
                                              int(check_ori), int(only_stereo), int(coarse), _p(m12), _p(pts), _p(mg))
    return m12[:nA], n, pts[:nA], mg[:nA]


class VocRef:
    """oracle/cpu_ref/bow_ref.cc: the DBoW2 vocabulary tree and its transform."""

    def __init__(self, k, L, weighting, norm, parent, leaf, desc, weight):
        L_ = lib()
        L_.orbref_voc_create.restype = C.c_void_p
        L_.orbref_voc_create.argtypes = [_I, _I, _I, _I, _I, _P, _P, _P, _P]
        L_.orbref_voc_destroy.argtypes = [_P]
        L_.orbref_voc_words.argtypes = [_P]
        L_.orbref_bow_transform.argtypes = [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P]
        L_.orbref_bow_descend.argtypes = [_P, _P, _I, _I, _P, _P, _P]
        L_.orbref_bow_transform_many.restype = _L
        L_.orbref_bow_transform_many.argtypes = [_P, _P, _I, _I, _L, _I, _I]
        parent, leaf = np.ascontiguousarray(parent, np.int32), np.ascontiguousarray(leaf, np.uint8)
        desc, weight = np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(weight, np.float64)
        self._v = L_.orbref_voc_create(k, L, weighting, norm, len(parent), _p(parent), _p(leaf), _p(desc), _p(weight))
        if not self._v:
            raise ValueError("inconsistent vocabulary arrays")
        self.nwords = L_.orbref_voc_words(self._v)

    def __del__(self):
        if getattr(self, "_v", None):
            lib().orbref_voc_destroy(self._v)
            self._v = None

    def transform(self, desc, levelsup):
        """Returns (words, values, node_ids, offsets, indices)."""
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        words, values = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        nodes, off, idx = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        cnt = np.zeros(2, np.int32)
        lib().orbref_bow_transform(self._v, _p(desc), n, levelsup, _p(words), _p(values), _p(nodes), _p(off), _p(idx), _p(cnt))
        nw, nn = int(cnt[0]), int(cnt[1])
        return words[:nw], values[:nw], nodes[:nn], off[:nn + 1], idx[:off[nn]]

    def descend(self, desc, levelsup):
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        word, w, node = np.zeros(n, np.uint32), np.zeros(n, np.float64), np.zeros(n, np.uint32)
        lib().orbref_bow_descend(self._v, _p(desc), n, levelsup, _p(word), _p(w), _p(node))
        return word, w, node

    def transform_many(self, desc, nframes, stride, levelsup, nthreads):
        return lib().orbref_bow_transform_many(self._v, _p(desc), desc.shape[-2], nframes, stride, levelsup, nthreads)


def search_by_projection(F, queries, mode, th, nnratio, link_stereo, rot_hist):
    """oracle/cpu_ref/proj_ref.cc on a projection.GridFrame and an array of PROJ_QUERY_DTYPE records.
    Returns (assign, choice, nmatches)."""
    L_ = lib()
    L_.orbref_search_by_projection.argtypes = [_P, _P, _I, _I, _P, _P, _P, _P, _F, _F, _F, _F, _P, _I, _I, _I, _F, _I, _I, _P, _P, _P]
    q = np.ascontiguousarray(queries)
    assign, choice = np.zeros(max(F.n, 1), np.int32), np.zeros(max(len(q), 1), np.int32)
    opt = lambda a: None if a is None else _p(a)
    n = L_.orbref_search_by_projection(_p(F.kps), _p(F.desc), F.n, F.nleft, opt(F.blocked), opt(F.u_right), opt(F.left_to_right),
                                       opt(F.right_to_left), float(F.min_x), float(F.min_y), float(F.grid_w_inv),
                                       float(F.grid_h_inv), _p(q), len(q), mode, th, float(nnratio), int(link_stereo),
                                       int(rot_hist), opt(F.inv_level_sigma2), _p(assign), _p(choice))
    return assign[:F.n], choice[:len(q)], n


def search_for_initialization(kps1, desc1, F2, prev_matched, window_size, nnratio, check_ori):
    """oracle/cpu_ref/proj_ref.cc searchForInitialization.  Returns (nmatches, vnMatches12, updated vbPrevMatched)."""
    L_ = lib()
    L_.orbref_search_for_initialization.argtypes = [_P, _P, _I, _P, _P, _I, _F, _F, _F, _F, _P, _I, _F, _I, _P]
    kps1, desc1 = np.ascontiguousarray(kps1), np.ascontiguousarray(desc1, np.uint8)
    prev = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(max(len(kps1), 1), np.int32)
    n = L_.orbref_search_for_initialization(_p(kps1), _p(desc1), len(kps1), _p(F2.kps), _p(F2.desc), F2.n, float(F2.min_x),
                                            float(F2.min_y), float(F2.grid_w_inv), float(F2.grid_h_inv), _p(prev), int(window_size),
                                            float(nnratio), int(check_ori), _p(m12))
    return n, m12[:len(kps1)], prev


def features_in_area(F, x, y, r, min_level, max_level, right):
    L_ = lib()
    L_.orbref_features_in_area.argtypes = [_P, _I, _I, _F, _F, _F, _F, _F, _F, _F, _I, _I, _I, _P]
    out = np.zeros(max(F.n, 1), np.int32)
    n = L_.orbref_features_in_area(_p(F.kps), F.n, F.nleft, float(F.min_x), float(F.min_y), float(F.grid_w_inv),
                                   float(F.grid_h_inv), float(x), float(y), float(r), min_level, max_level, int(right), _p(out))
    return out[:n]


def distinctive_descriptors(desc, offsets):
    desc, offsets = np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(offsets, np.int32)
    best = np.zeros(max(len(offsets) - 1, 1), np.int32)
    L_ = lib()
    L_.orbref_distinctive_descriptors.argtypes = [_P, _P, _I, _P]
    L_.orbref_distinctive_descriptors(_p(desc), _p(offsets), len(offsets) - 1, _p(best))
    return best[:len(offsets) - 1]


def undistort_keypoints(kps, K4, dist):
    kps = np.ascontiguousarray(kps)
    K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(kps)
    L_ = lib()
    L_.orbref_undistort_keypoints.argtypes = [_P, _I, _P, _P, _I, _P]
    L_.orbref_undistort_keypoints(_p(kps), len(kps), _p(K4), _p(dist), len(dist), _p(out))
    return out


# ---- the reference's own ORBextractor.cc, compiled against the OpenCV stand-in (oracle/_ref, see oracle/Makefile) ----
REF_LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "liborbslam3_orbextractor.so")
_ref_lib = None


def ref_available():
    return os.path.exists(REF_LIB)


def ref_lib():
    global _ref_lib
    if _ref_lib is None:
        L = C.CDLL(REF_LIB)
        L.orbslam3_extractor_create.restype = C.c_void_p
        L.orbslam3_extractor_create.argtypes = [_I, _F, _I, _I, _I]
        L.orbslam3_extractor_destroy.argtypes = [C.c_void_p]
        L.orbslam3_extractor_extract.restype = _I
        L.orbslam3_extractor_extract.argtypes = [C.c_void_p, _P, _I, _I, C.c_long, _I, _I, _P, _P, _I, _P]
        L.orbslam3_extractor_pyramid.restype = _I
        L.orbslam3_extractor_pyramid.argtypes = [C.c_void_p, _I, _P, _I, _P, _P]
        L.orbslam3_extractor_tables.argtypes = [C.c_void_p, _P, _P, _P, _P]
        L.orbslam3_extractor_pyramid_bordered.restype = _I
        L.orbslam3_extractor_pyramid_bordered.argtypes = [C.c_void_p, _I, _P, _I, _P, _P]
        _ref_lib = L
    return _ref_lib


class OrbSlam3Extractor:
    """ORB_SLAM3::ORBextractor of /root/reference/src/ORBextractor.cc itself (its OpenCV calls go to the stand-in
    in oracle/refshim, whose primitives are oracle/cpu_ref/cvprim.cc)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST):
        self.L = ref_lib()
        self.h = self.L.orbslam3_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self.nlevels = nlevels
        self.cap = nfeatures + 64 * nlevels + 4096

    def __del__(self):
        try:
            self.L.orbslam3_extractor_destroy(self.h)
        except Exception:
            pass

    def __call__(self, image, lapping=(0, 0)):
        image = np.ascontiguousarray(image)
        h, w = image.shape
        kps = np.zeros(self.cap, KEYPOINT_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        mono = C.c_int(0)
        n = self.L.orbslam3_extractor_extract(self.h, _p(image), w, h, w, int(lapping[0]), int(lapping[1]), _p(kps), _p(desc),
                                              self.cap, C.byref(mono))
        assert n >= 0, "reference capacity"
        return mono.value, kps[:n].copy(), desc[:n].copy()

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        buf = np.zeros(4096 * 4096, np.uint8)
        assert self.L.orbslam3_extractor_pyramid(self.h, l, _p(buf), buf.size, C.byref(w), C.byref(h)) == 0
        return buf[:w.value * h.value].reshape(h.value, w.value).copy()

    def level_bordered(self, l):
        w, h = C.c_int(), C.c_int()
        buf = np.zeros(4096 * 4096, np.uint8)
        assert self.L.orbslam3_extractor_pyramid_bordered(self.h, l, _p(buf), buf.size, C.byref(w), C.byref(h)) == 0
        return buf[:w.value * h.value].reshape(h.value, w.value).copy()

    def tables(self):
        t = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        self.L.orbslam3_extractor_tables(self.h, *[_p(a) for a in t])
        return t


REF_MT_LIB = REF_LIB.replace(".so", "_mt.so")
_ref_mt = None


def ref_extract_many(orb_params, frames, lapping, nthreads):
    """Frame-parallel run of the reference's ORBextractor (process allocator build), one instance per thread."""
    global _ref_mt
    if _ref_mt is None:
        _ref_mt = C.CDLL(REF_MT_LIB)
        _ref_mt.orbslam3_extract_many.restype = C.c_long
        _ref_mt.orbslam3_extract_many.argtypes = [_I, _F, _I, _I, _I, _P, _I, _I, _I, _I, _I, _I]
    frames = np.ascontiguousarray(frames)
    n, h, w = frames.shape
    nf, sf, nl, ini, mn = orb_params
    return _ref_mt.orbslam3_extract_many(nf, sf, nl, ini, mn, _p(frames), n, w, h, int(lapping[0]), int(lapping[1]), int(nthreads))


# ---- the reference's own vendored DBoW2 (oracle/_ref/libdbow2_orbvoc.so, see oracle/Makefile) ----
REF_VOC_LIB = os.path.join(os.path.dirname(REF_LIB), "libdbow2_orbvoc.so")
_ref_voc = None


def ref_voc_available():
    return os.path.exists(REF_VOC_LIB)


class DBoW2Vocabulary:
    """ORBVocabulary (TemplatedVocabulary<FORB::TDescriptor, FORB>) of Thirdparty/DBoW2 itself, loaded with its own
    loadFromTextFile from an ORBvoc-format text file."""

    def __init__(self, path):
        global _ref_voc
        if _ref_voc is None:
            L = C.CDLL(REF_VOC_LIB)
            L.orbslam3_voc_load_text.restype = C.c_void_p
            L.orbslam3_voc_load_text.argtypes = [C.c_char_p]
            L.orbslam3_voc_destroy.argtypes = [C.c_void_p]
            L.orbslam3_voc_words.restype = _I
            L.orbslam3_voc_words.argtypes = [C.c_void_p]
            L.orbslam3_voc_transform.argtypes = [C.c_void_p, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P]
            _ref_voc = L
        self.L = _ref_voc
        self.h = self.L.orbslam3_voc_load_text(str(path).encode())
        assert self.h, "DBoW2 loadFromTextFile failed"
        self.nwords = self.L.orbslam3_voc_words(self.h)

    def __del__(self):
        try:
            self.L.orbslam3_voc_destroy(self.h)
        except Exception:
            pass

    def transform(self, desc, levelsup):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        words, values = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        nodes, off, idx = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nw, nn = C.c_int(0), C.c_int(0)
        self.L.orbslam3_voc_transform(self.h, _p(desc), n, int(levelsup), _p(words), _p(values), C.byref(nw), _p(nodes), _p(off),
                                      _p(idx), C.byref(nn))
        return words[:nw.value], values[:nw.value], nodes[:nn.value], off[:nn.value + 1], idx[:off[nn.value]]


# ---- the reference's own ORBmatcher.cc (oracle/_ref/liborbslam3_orbmatcher.so, see oracle/Makefile) ----
REF_MATCHER_LIB = os.path.join(os.path.dirname(REF_LIB), "liborbslam3_orbmatcher.so")
_ref_matcher = None


def ref_matcher_available():
    return os.path.exists(REF_MATCHER_LIB)


def _load_matcher(path):
    return _set_matcher_argtypes(C.CDLL(path))


def _set_matcher_argtypes(L):
    L.orbslam3_search_bow.argtypes = [_I, _F, _I, _P, _P, _I, _P, _P, _P, _P, _I, _P, _P, _I, _P, _P, _P, _P, _I, _I, _P, _P]
    L.orbslam3_search_for_initialization.argtypes = [_F, _I, _P, _P, _I, _P, _P, _I, _F, _F, _F, _F, _P, _I, _P]
    L.orbslam3_search_by_projection_mappoints.argtypes = [_F, _P, _P, _I, _I, _P, _P, _P, _P, _F, _F, _F, _F, _P, _I, _I] + [_P] * 14 + \
        [_F, _I, _F, _P]
    return L


def ref_matcher():
    global _ref_matcher
    if _ref_matcher is None:
        _ref_matcher = _load_matcher(REF_MATCHER_LIB)
    return _ref_matcher


# ---- the reference's ORBmatcher.cc TOGETHER WITH its own camera models (oracle/_ref/liborbslam3_orbmatcher_cam.so) ----
REF_MATCHER_CAM_LIB = os.path.join(os.path.dirname(REF_LIB), "liborbslam3_orbmatcher_cam.so")
_ref_matcher_cam = None


def ref_matcher_cam_available():
    return os.path.exists(REF_MATCHER_CAM_LIB)


class _KfArrays(C.Structure):
    _fields_ = [("kps", _P), ("desc", _P), ("n", _I), ("nleft", _I), ("hasMp", _P), ("uRight", _P), ("nodes", _P), ("off", _P),
                ("idx", _P), ("nnodes", _I), ("sigma2", _P), ("scale", _P), ("nlevels", _I), ("cam", _P), ("Tcw", _P), ("Trl", _P)]


SHIM_MATCHER_CAM_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "liborbfe_shim_orbmatcher_cam.so")
_shim_matcher_cam = None


def shim_matcher_cam_available():
    return os.path.exists(SHIM_MATCHER_CAM_LIB)


def ref_search_for_triangulation_cam(fisheye, A, B, nnratio, check_ori, only_stereo, coarse, shim=False):
    """SearchForTriangulation_ of the reference on its own Pinhole / KannalaBrandt8 classes, from keyframe POSES.
    A, B: dicts with kps, desc, fv, mp, nleft, u_right (or None), sigma2, scale, cam (2 x 8), Tcw (4 x 4), Trl (4 x 4).
    shim=True: the same call on ORB_SLAM3::ORBmatcher of include/orbfe_shim (GPU-backed), same objects, same real cameras.
    Returns (matches12, nmatches)."""
    global _ref_matcher_cam, _shim_matcher_cam
    if shim and _shim_matcher_cam is None:
        _shim_matcher_cam = C.CDLL(SHIM_MATCHER_CAM_LIB)
        _shim_matcher_cam.orbslam3_search_for_triangulation_cam.argtypes = [_F, _I, _I, _P, _P, _I, _I, _P]
    if not shim and _ref_matcher_cam is None:
        _ref_matcher_cam = C.CDLL(REF_MATCHER_CAM_LIB)
        _ref_matcher_cam.orbslam3_search_for_triangulation_cam.argtypes = [_F, _I, _I, _P, _P, _I, _I, _P]
    library = _shim_matcher_cam if shim else _ref_matcher_cam
    keep = []

    def side(d):
        f = lambda a: np.ascontiguousarray(a, np.float32)
        arrs = dict(kps=np.ascontiguousarray(d["kps"]), desc=np.ascontiguousarray(d["desc"], np.uint8),
                    hasMp=np.ascontiguousarray(d["mp"]).astype(np.uint8),
                    uRight=None if d.get("u_right") is None else f(d["u_right"]), nodes=np.ascontiguousarray(d["fv"].node_ids),
                    off=np.ascontiguousarray(d["fv"].offsets), idx=np.ascontiguousarray(d["fv"].indices), sigma2=f(d["sigma2"]),
                    scale=f(d["scale"]), cam=f(d["cam"]), Tcw=f(d["Tcw"]), Trl=f(d["Trl"]))
        assert arrs["cam"].size == 16 and arrs["Tcw"].size == 16 and arrs["Trl"].size == 16
        keep.append(arrs)
        ptr = {k: (_pn(v) if v is None else _p(v)) for k, v in arrs.items()}
        return _KfArrays(n=len(arrs["desc"]), nleft=int(d["nleft"]), nnodes=len(arrs["nodes"]), nlevels=len(arrs["sigma2"]), **ptr)

    a, b = side(A), side(B)
    m12 = np.zeros(max(a.n, 1), np.int32)
    n = library.orbslam3_search_for_triangulation_cam(float(nnratio), int(check_ori), int(fisheye), C.byref(a), C.byref(b),
                                                               int(only_stereo), int(coarse), _p(m12))
    return m12[:a.n], n


# ---- the reference's matcher on its OWN Frame / KeyFrame / MapPoint / camera classes (oracle/_ref/liborbslam3_objects.so) ----
REF_OBJECTS_LIB = os.path.join(os.path.dirname(REF_LIB), "liborbslam3_objects.so")
_ref_objects = None


def ref_objects_available():
    return os.path.exists(REF_OBJECTS_LIB)


def _ref_objects_lib():
    global _ref_objects
    if _ref_objects is None:
        _ref_objects = C.CDLL(REF_OBJECTS_LIB)
        _ref_objects.orbslam3_objects_search_for_triangulation.argtypes = [_F, _I, _I, _P, _P, _I, _I, _P]
        _ref_objects.orbslam3_objects_search_by_bow_kf_kf.argtypes = [_F, _I, _P, _P, _P]
    return _ref_objects


class objects_matcher_library:
    """with objects_matcher_library(): the ref_search_by_projection_* calls below run on liborbslam3_objects.so - the same C names,
    but real reference Frames (their own grid) and real MapPoints instead of the stand-ins of orbslam3_standin.h."""

    def __enter__(self):
        global _ref_matcher
        L = _ref_objects_lib()
        if not getattr(L, "_matcher_types_set", False):
            _set_matcher_argtypes(L)
            L._matcher_types_set = True
        self.saved = ref_matcher() if ref_matcher_available() else None
        _ref_matcher = L
        return self

    def __exit__(self, *exc):
        global _ref_matcher
        _ref_matcher = self.saved
        return False


def _kf_arrays(d, keep):
    f = lambda a: np.ascontiguousarray(a, np.float32)
    arrs = dict(kps=np.ascontiguousarray(d["kps"]), desc=np.ascontiguousarray(d["desc"], np.uint8),
                hasMp=np.ascontiguousarray(d["mp"]).astype(np.uint8),
                uRight=None if d.get("u_right") is None else f(d["u_right"]), nodes=np.ascontiguousarray(d["fv"].node_ids),
                off=np.ascontiguousarray(d["fv"].offsets), idx=np.ascontiguousarray(d["fv"].indices), sigma2=f(d["sigma2"]),
                scale=f(d["scale"]), cam=f(d["cam"]), Tcw=f(d["Tcw"]), Trl=f(d["Trl"]))
    assert arrs["cam"].size == 16 and arrs["Tcw"].size == 16 and arrs["Trl"].size == 16
    keep.append(arrs)
    ptr = {k: (_pn(v) if v is None else _p(v)) for k, v in arrs.items()}
    return _KfArrays(n=len(arrs["desc"]), nleft=int(d["nleft"]), nnodes=len(arrs["nodes"]), nlevels=len(arrs["sigma2"]), **ptr)


def ref_objects_search_for_triangulation(fisheye, A, B, nnratio, check_ori, only_stereo, coarse):
    """SearchForTriangulation_ of the reference on two KeyFrames made by the reference's own KeyFrame(Frame&, Map*, KeyFrameDatabase*)
    from reference Frames; cameras are the reference's classes.  A, B as for ref_search_for_triangulation_cam."""
    keep = []
    a, b = _kf_arrays(A, keep), _kf_arrays(B, keep)
    m12 = np.zeros(max(a.n, 1), np.int32)
    n = _ref_objects_lib().orbslam3_objects_search_for_triangulation(float(nnratio), int(check_ori), int(fisheye), C.byref(a), C.byref(b),
                                                                    int(only_stereo), int(coarse), _p(m12))
    return m12[:a.n], n


def ref_objects_search_by_bow_kf_kf(A, B, nnratio, check_ori):
    """SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) on real KeyFrames holding real MapPoints.  Returns (a2b, nmatches)."""
    keep = []
    a, b = _kf_arrays(A, keep), _kf_arrays(B, keep)
    a2b = np.zeros(max(a.n, 1), np.int32)
    n = _ref_objects_lib().orbslam3_objects_search_by_bow_kf_kf(float(nnratio), int(check_ori), C.byref(a), C.byref(b), _p(a2b))
    return a2b[:a.n], n


# The orbfe shim's ORBmatcher object layer behind the same C entry points (make -C oracle shim): GPU-backed.
SHIM_MATCHER_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "liborbfe_shim_orbmatcher.so")
_shim_matcher = None


def shim_matcher_available():
    return os.path.exists(SHIM_MATCHER_LIB)


class shim_matcher_library:
    """with shim_matcher_library(): every ref_* matcher call below runs ORB_SLAM3::ORBmatcher of include/orbfe_shim instead of the
    reference's ORBmatcher.cc (same stand-in Frame / KeyFrame / MapPoint objects, same C wrappers)."""

    def __enter__(self):
        global _ref_matcher, _shim_matcher
        if _shim_matcher is None:
            _shim_matcher = _load_matcher(SHIM_MATCHER_LIB)
        self.saved = ref_matcher() if ref_matcher_available() else None
        _ref_matcher = _shim_matcher
        return self

    def __exit__(self, *exc):
        global _ref_matcher
        _ref_matcher = self.saved
        return False


def _kp_with_angle(angles):
    k = np.zeros(len(angles), KEYPOINT_DTYPE)
    k["angle"] = angles
    return k


def ref_descriptor_distance(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return ref_matcher().orbslam3_descriptor_distance(_p(a), _p(b))


def ref_three_maxima(counts):
    c = np.ascontiguousarray(counts, np.int32)
    out = np.zeros(3, np.int32)
    ref_matcher().orbslam3_three_maxima(_p(c), len(c), _p(out))
    return tuple(int(v) for v in out)


def ref_search_bow(form, descA, fvA, maskA, angA, descB, fvB, maskB, angB, ratio, check_ori, nleft=-1):
    """ORBmatcher::SearchByBoW of the reference itself.  form 0: (KeyFrame*, Frame&), form 1: (KeyFrame*, KeyFrame*)."""
    descA, descB = np.ascontiguousarray(descA, np.uint8), np.ascontiguousarray(descB, np.uint8)
    maskA = np.ascontiguousarray(maskA).astype(np.uint8)
    maskB = None if maskB is None else np.ascontiguousarray(maskB).astype(np.uint8)
    kA, kB = _kp_with_angle(angA), _kp_with_angle(angB)
    a2b, b2a = np.zeros(max(len(descA), 1), np.int32), np.zeros(max(len(descB), 1), np.int32)
    n = ref_matcher().orbslam3_search_bow(form, ratio, int(check_ori), _p(kA), _p(descA), len(descA), _p(maskA), _p(fvA.node_ids),
                                          _p(fvA.offsets), _p(fvA.indices), len(fvA.node_ids), _p(kB), _p(descB), len(descB), _pn(maskB),
                                          _p(fvB.node_ids), _p(fvB.offsets), _p(fvB.indices), len(fvB.node_ids), nleft, _p(a2b), _p(b2a))
    return a2b[:len(descA)], b2a[:len(descB)], n


def ref_search_for_initialization(kps1, desc1, F2, prev_matched, window_size, nnratio, check_ori):
    kps1, desc1 = np.ascontiguousarray(kps1), np.ascontiguousarray(desc1, np.uint8)
    prev = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(max(len(kps1), 1), np.int32)
    n = ref_matcher().orbslam3_search_for_initialization(float(nnratio), int(check_ori), _p(kps1), _p(desc1), len(kps1), _p(F2.kps),
                                                         _p(F2.desc), F2.n, float(F2.min_x), float(F2.min_y), float(F2.max_x),
                                                         float(F2.max_y), _p(prev), int(window_size), _p(m12))
    return n, m12[:len(kps1)], prev


def ref_search_by_projection_mappoints(F, mp, th, nnratio, far, th_far):
    """ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th, bFarPoints, thFarPoints) of the reference itself on a
    projection.GridFrame and the MapPoint member arrays of tests/proj_cases.mp_inputs."""
    m = len(mp["desc"])
    u8 = lambda a: np.ascontiguousarray(a).astype(np.uint8)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    i32 = lambda a: np.ascontiguousarray(a, np.int32)
    arrs = [u8(mp["in_view"]), u8(mp["in_view_r"]), u8(mp["bad"]), f32(mp["depth"]), f32(mp["proj_x"]), f32(mp["proj_y"]),
            f32(mp["proj_xr"]), f32(mp["proj_yr"]), i32(mp["level"]), i32(mp["level_r"]), f32(mp["view_cos"]), f32(mp["view_cos_r"]),
            np.ascontiguousarray(mp["desc"], np.uint8), i32(mp["obs"])]
    out = np.zeros(max(F.n, 1), np.int32)
    sf = f32(F.scale_factors)
    opt = lambda a: None if a is None else _p(a)
    n = ref_matcher().orbslam3_search_by_projection_mappoints(
        float(nnratio), _p(F.kps), _p(F.desc), F.n, F.nleft, opt(F.blocked), opt(F.u_right), opt(F.left_to_right), opt(F.right_to_left),
        float(F.min_x), float(F.min_y), float(F.max_x), float(F.max_y), _p(sf), len(sf), m, *[_p(a) for a in arrs], float(th), int(far),
        float(th_far), _p(out))
    return n, out[:F.n]


def ref_search_triangulation(kpA, descA, fvA, mpA, urA, kpB, descB, fvB, mpB, urB, sigma2, scale, F12, ep, nnratio, check_ori,
                             only_stereo, coarse):
    """ORBmatcher::SearchForTriangulation_ of the reference itself (pinhole; F12 and the epipole as the C ABI receives them)."""
    L = ref_matcher()
    L.orbslam3_search_for_triangulation.argtypes = [_F, _I, _P, _P, _I, _P, _P, _P, _P, _P, _I, _P, _P, _I, _P, _P, _P, _P, _P, _I, _P, _P,
                                                    _I, _P, _F, _F, _I, _I, _P]
    kpA, kpB = np.ascontiguousarray(kpA), np.ascontiguousarray(kpB)
    descA, descB = np.ascontiguousarray(descA, np.uint8), np.ascontiguousarray(descB, np.uint8)
    mpA, mpB = np.ascontiguousarray(mpA).astype(np.uint8), np.ascontiguousarray(mpB).astype(np.uint8)
    urA, urB = np.ascontiguousarray(urA, np.float32), np.ascontiguousarray(urB, np.float32)
    sigma2, scale = np.ascontiguousarray(sigma2, np.float32), np.ascontiguousarray(scale, np.float32)
    F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
    m12 = np.zeros(max(len(descA), 1), np.int32)
    n = L.orbslam3_search_for_triangulation(float(nnratio), int(check_ori), _p(kpA), _p(descA), len(descA), _p(mpA), _p(urA),
                                            _p(fvA.node_ids), _p(fvA.offsets), _p(fvA.indices), len(fvA.node_ids), _p(kpB), _p(descB),
                                            len(descB), _p(mpB), _p(urB), _p(fvB.node_ids), _p(fvB.offsets), _p(fvB.indices),
                                            len(fvB.node_ids), _p(sigma2), _p(scale), len(sigma2), _p(F12), float(ep[0]), float(ep[1]),
                                            int(only_stereo), int(coarse), _p(m12))
    return m12[:len(descA)], n


def ref_search_by_projection_lastframe(Cur, last, Tcw, th, fb, cam, Trl, check_ori):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) of the reference itself.  fb = (bForward,
    bBackward): LastFrame's pose is the current one moved +-1 m along z (or not at all) so that the reference derives
    those flags itself (:2184-2194) with mb = 0.11."""
    L = ref_matcher()
    L.orbslam3_search_by_projection_lastframe.argtypes = [_I, _P, _P, _I, _I, _P, _P, _F, _F, _F, _F, _P, _I, _F, _F, _P, _P, _P, _P, _I,
                                                          _P, _P, _P, _P, _P, _P, _P, _F, _I, _P]
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: np.ascontiguousarray(a).astype(np.uint8)
    Tcw = f32(Tcw).reshape(4, 4)
    Tlw = Tcw.copy()
    Tlw[2, 3] += np.float32(1.0 if fb[0] else -1.0 if fb[1] else 0.0)
    sf = f32(Cur.scale_factors)
    cam = f32(cam[:4])
    m = len(last["has_mp"])
    arrs = [u8(last["has_mp"]), u8(last["outlier"]), np.ascontiguousarray(last["octave"], np.int32), f32(last["angle"]),
            f32(last["world_pos"]), np.ascontiguousarray(last["desc"], np.uint8), np.ascontiguousarray(last["obs"], np.int32)]
    out = np.zeros(max(Cur.n, 1), np.int32)
    opt = lambda a: None if a is None else _p(a)
    trl = None if Trl is None else f32(Trl).reshape(4, 4)
    n = L.orbslam3_search_by_projection_lastframe(int(check_ori), _p(Cur.kps), _p(Cur.desc), Cur.n, Cur.nleft, opt(Cur.blocked),
                                                  opt(Cur.u_right), float(Cur.min_x), float(Cur.min_y), float(Cur.max_x), float(Cur.max_y),
                                                  _p(sf), len(sf), float(Cur.mbf), 0.11, _p(cam), _p(Tcw), _p(Tlw), opt(trl), m,
                                                  *[_p(a) for a in arrs], float(th), 0, _p(out))
    return n, out[:Cur.n]


class _KfInput(C.Structure):
    _fields_ = [("kps", _P), ("desc", _P), ("n", _I), ("nleft", _I), ("blocked", _P), ("uRight", _P), ("minX", _F), ("minY", _F),
                ("maxX", _F), ("maxY", _F), ("scaleFactors", _P), ("invLevelSigma2", _P), ("nlevels", _I), ("cam", _F * 4), ("mbf", _F),
                ("logScaleFactor", _F), ("Tcw", _P)]


class _KfPoints(C.Structure):
    _fields_ = [("m", _I), ("null", _P), ("bad", _P), ("flagged", _P), ("worldPos", _P), ("normal", _P), ("minDist", _P),
                ("maxDist", _P), ("desc", _P)]


def _kf_structs(KF, pts, cam, Tcw, log_sf, flagged_key):
    keep = []                                           # arrays must outlive the call

    def arr(a, dt):
        if a is None:
            return None
        a = np.ascontiguousarray(np.asarray(a).astype(dt) if dt == np.uint8 else a, dt)
        keep.append(a)
        return C.cast(_p(a), _P)
    k = _KfInput()
    k.kps, k.desc, k.n, k.nleft = arr(KF.kps, KF.kps.dtype), arr(KF.desc, np.uint8), KF.n, KF.nleft
    k.blocked, k.uRight = arr(KF.blocked, np.uint8), arr(KF.u_right, np.float32)
    k.minX, k.minY, k.maxX, k.maxY = float(KF.min_x), float(KF.min_y), float(KF.max_x), float(KF.max_y)
    k.scaleFactors, k.invLevelSigma2 = arr(KF.scale_factors, np.float32), arr(getattr(KF, "inv_level_sigma2", None), np.float32)
    k.nlevels = len(KF.scale_factors)
    k.cam = (_F * 4)(*[float(c) for c in cam[:4]])
    k.mbf, k.logScaleFactor = float(KF.mbf), float(log_sf)
    T = np.eye(4, dtype=np.float32)
    T[:np.asarray(Tcw).shape[0]] = np.asarray(Tcw, np.float32)
    k.Tcw = arr(T, np.float32)
    p = _KfPoints()
    p.m = len(pts["desc"])
    p.null, p.bad = arr(pts.get("null"), np.uint8), arr(pts.get("bad"), np.uint8)
    p.flagged = arr(pts.get(flagged_key), np.uint8)
    p.worldPos, p.normal = arr(pts["world_pos"], np.float32), arr(pts["normal"], np.float32)
    p.minDist, p.maxDist, p.desc = arr(pts["min_dist"], np.float32), arr(pts["max_dist"], np.float32), arr(pts["desc"], np.uint8)
    return k, p, keep


def ref_search_by_projection_sim3(KF, pts, Scw, th, ratio, cam, log_sf):
    """ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming) of the reference itself."""
    L = ref_matcher()
    L.orbslam3_search_by_projection_sim3.argtypes = [_P, _P, _P, _I, _F, _P]
    k, p, keep = _kf_structs(KF, pts, cam, np.eye(4), log_sf, "already_found")
    S = np.ascontiguousarray(Scw, np.float32)
    out = np.zeros(max(KF.n, 1), np.int32)
    n = L.orbslam3_search_by_projection_sim3(C.byref(k), _p(S), C.byref(p), int(th), float(ratio), _p(out))
    return n, out[:KF.n]


def ref_fuse(KF, pts, Tcw, Scw, th, right, cam, log_sf):
    """ORBmatcher::Fuse of the reference itself: (pKF, vpMapPoints, th, bRight) when Scw is None, else (pKF, Scw, vpPoints, th, ...)."""
    L = ref_matcher()
    L.orbslam3_fuse.argtypes = [_P, _P, _P, _F, _I, _P]
    k, p, keep = _kf_structs(KF, pts, cam, np.eye(4) if Tcw is None else Tcw, log_sf, "in_keyframe" if Scw is None else "already_found")
    S = None if Scw is None else np.ascontiguousarray(Scw, np.float32)
    out = np.zeros(max(len(pts["desc"]), 1), np.int32)
    n = L.orbslam3_fuse(C.byref(k), None if S is None else _p(S), C.byref(p), float(th), int(right), _p(out))
    return n, out[:len(pts["desc"])]


def ref_search_by_sim3(c, cam, log_sf):
    """ORBmatcher::SearchBySim3 of the reference itself on a tests/proj_cases.sim3_inputs scenario.
    Returns (nFound, vpMatches12 as keyframe-2 feature indices)."""
    L = ref_matcher()
    L.orbslam3_search_by_sim3.argtypes = [_P, _P, _P, _P, _F, _P, _P, _F, _P]
    keep = []

    def side(KF, mps, T):
        pts = dict(mps)
        pts["null"] = ~np.asarray(mps["has_mp"], bool)
        k, p, kp = _kf_structs(KF, pts, cam, T, log_sf, "already_matched")
        keep.append(kp)
        return k, p
    k1, p1 = side(c["KF1"], c["mps1"], c["T1w"])
    k2, p2 = side(c["KF2"], c["mps2"], c["T2w"])
    R12, t12 = np.ascontiguousarray(c["R12"], np.float32), np.ascontiguousarray(c["t12"], np.float32).reshape(3)
    out = np.zeros(max(c["KF1"].n, 1), np.int32)
    n = L.orbslam3_search_by_sim3(C.byref(k1), C.byref(k2), C.byref(p1), C.byref(p2), float(c["s12"]), _p(R12), _p(t12), float(c["th"]),
                                  _p(out))
    return n, out[:c["KF1"].n]


def ref_search_by_projection_reloc(Cur, kf, Tcw, th, orb_dist, cam, log_sf, check_ori=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) of the reference itself; `kf` as for
    projection.queries_from_keyframe.  Returns (nmatches, keyframe feature per frame feature)."""
    L = ref_matcher()
    L.orbslam3_search_by_projection_reloc.argtypes = [_I, _P, _P, _P, _F, _I, _P]
    pts = dict(kf)
    m = len(kf["has_mp"])
    pts["null"] = ~np.asarray(kf["has_mp"], bool)
    pts["normal"] = np.zeros((m, 3), np.float32)
    k, p, keep = _kf_structs(Cur, pts, cam, Tcw, log_sf, "already_found")
    ang = np.ascontiguousarray(kf["angle"], np.float32)
    out = np.zeros(max(Cur.n, 1), np.int32)
    n = L.orbslam3_search_by_projection_reloc(int(check_ori), C.byref(k), C.byref(p), _p(ang), float(th), int(orb_dist), _p(out))
    return n, out[:Cur.n]


# ---- the reference's own src/MapPoint.cc (oracle/_ref/liborbslam3_mappoint.so, see oracle/Makefile) ----
REF_MAPPOINT_LIB = os.path.join(os.path.dirname(REF_LIB), "liborbslam3_mappoint.so")
_ref_mappoint = None


def ref_mappoint_available():
    return os.path.exists(REF_MAPPOINT_LIB) and ref_matcher_available()


def ref_mappoint():
    global _ref_mappoint
    if _ref_mappoint is None:
        L = C.CDLL(REF_MAPPOINT_LIB)
        L.orbslam3_distinctive_descriptors.argtypes = [_P, _P, _I, _I, _P]
        L.orbslam3_mappoint_nobs.argtypes = [_I, _I]
        L.orbslam3_mappoint_predict_scale.argtypes = [_P, _P, _I, _F, _I, _I, _P]
        _ref_mappoint = L
    return _ref_mappoint


def ref_distinctive_descriptors(desc, offsets, stereo=False):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423), the reference's own object code."""
    desc = np.ascontiguousarray(desc, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.int32)
    best = np.full(len(offsets) - 1, -2, np.int32)
    ref_mappoint().orbslam3_distinctive_descriptors(_p(desc), _p(offsets), len(offsets) - 1, int(stereo), _p(best))
    return best


def ref_predict_scale(max_distance, current_dist, log_scale_factor, n_levels, via_frame=False):
    """MapPoint::PredictScale (src/MapPoint.cc:534-566), the reference's own object code."""
    md = np.ascontiguousarray(max_distance, np.float32)
    cd = np.ascontiguousarray(current_dist, np.float32)
    out = np.empty(len(md), np.int32)
    ref_mappoint().orbslam3_mappoint_predict_scale(_p(md), _p(cd), len(md), float(log_scale_factor), int(n_levels), int(via_frame), _p(out))
    return out


# ---- the reference's own src/Frame.cc + camera models (oracle/_ref/liborbslam3_frame.so, see oracle/Makefile) ----
REF_FRAME_LIB = os.path.join(os.path.dirname(REF_LIB), "liborbslam3_frame.so")
_ref_frame = None


def ref_frame_available():
    return os.path.exists(REF_FRAME_LIB) and ref_matcher_available()


def _load_frame(path):
    L = C.CDLL(path)
    L.orbslam3_frame_stereo_matches.argtypes = [_I, _F, _I, _I, _I, _P, _P, _P, _P, _P, _P, _I, _P, _P, _I, _F, _F, _P, _P]
    L.orbslam3_frame_fisheye_matches.argtypes = [_P, _P, _P, _P, _I, _I, _P, _P, _I, _I, _P, _P, _P, _I, _P, _P, _P, _P]
    L.orbslam3_frame_undistort_keypoints.argtypes = [_P, _I, _P, _P, _I, _P]
    L.orbslam3_frame_features_in_area.argtypes = [_P, _I, _I, _F, _F, _F, _F, _P, _I, _P, _I, _P]
    L.orbslam3_kb8_project.argtypes = [_P, _P, _I, _P]
    L.orbslam3_pinhole_project.argtypes = [_P, _P, _I, _P]
    L.orbslam3_frame_fisheye_construct.argtypes = [_P, _P, _I, _I, _I, _F, _I, _I, _I, _P, _P, _P, _P, _P, _F, _F, _I] + [_P] * 7
    L.orbslam3_frame_mono_construct.argtypes = [_P, _I, _I, _I, _F, _I, _I, _I, _P, _P, _I, C.c_char_p, _I] + [_P] * 9
    L.orbslam3_kb8_epipolar_constrain_.argtypes = [_P] * 4 + [_I] + [_P] * 7
    L.orbslam3_frame_stereo_construct.argtypes = [_P, _P, _I, _I, _I, _F, _I, _I, _I, _P, _P, _I, _F, _F, _I] + [_P] * 9
    L.orbslam3_kb8_unproject.argtypes = [_P, _P, _I, _P]
    return L


def ref_frame():
    global _ref_frame
    if _ref_frame is None:
        _ref_frame = _load_frame(REF_FRAME_LIB)
    return _ref_frame


# The reference's own Frame.cc compiled against the orbfe shim headers + liborbfe.so (make -C oracle shim): its Frame constructors
# extract on the GPU.
SHIM_FRAME_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "liborbfe_shim_frame.so")
_shim_frame = None


def shim_frame_available():
    return os.path.exists(SHIM_FRAME_LIB)


class shim_frame_library:
    """with shim_frame_library(): ref_stereo_frame / ref_fisheye_frame / ref_mono_frame / ref_stereo_matches run the reference's
    Frame.cc whose ORBextractor is the orbfe shim class (GPU) instead of the reference's ORBextractor.cc."""

    def __enter__(self):
        global _ref_frame, _shim_frame
        if _shim_frame is None:
            _shim_frame = _load_frame(SHIM_FRAME_LIB)
        self.saved = ref_frame() if ref_frame_available() else None
        _ref_frame = _shim_frame
        return self

    def __exit__(self, *exc):
        global _ref_frame
        _ref_frame = self.saved
        return False


def ref_stereo_matches(params, levelsL, levelsR, kpL, descL, kpR, descR, bf, b):
    """Frame::ComputeStereoMatches (src/Frame.cc:794-964), the reference's own object code.  levelsL / levelsR: the
    bordered pyramid levels ((h + 38) x (w + 38) uint8 arrays) of the two extractors."""
    nf, sf, nl, ini, mn = params
    levelsL = [np.ascontiguousarray(a, np.uint8) for a in levelsL]
    levelsR = [np.ascontiguousarray(a, np.uint8) for a in levelsR]
    w = np.array([a.shape[1] - 38 for a in levelsL], np.int32)
    h = np.array([a.shape[0] - 38 for a in levelsL], np.int32)
    pL = (C.c_void_p * nl)(*[a.ctypes.data for a in levelsL])
    pR = (C.c_void_p * nl)(*[a.ctypes.data for a in levelsR])
    kpL, kpR = np.ascontiguousarray(kpL), np.ascontiguousarray(kpR)
    descL, descR = np.ascontiguousarray(descL, np.uint8), np.ascontiguousarray(descR, np.uint8)
    u, d = np.zeros(max(len(kpL), 1), np.float32), np.zeros(max(len(kpL), 1), np.float32)
    ref_frame().orbslam3_frame_stereo_matches(nf, sf, nl, ini, mn, pL, pR, _p(w), _p(h), _p(kpL), _p(descL), len(kpL), _p(kpR), _p(descR),
                                              len(kpR), bf, b, _p(u), _p(d))
    return u[:len(kpL)], d[:len(kpL)]


def ref_fisheye_stereo(cam1, cam2, kpL, descL, mono_left, kpR, descR, mono_right, R12, t12, sigma2):
    """Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1116-1156) + KannalaBrandt8::TriangulateMatches, the reference's own
    object code (its cv::SVD is the stand-in's double-precision Jacobi).  Returns (l2r, r2l, depth, p3d, n)."""
    cam1, cam2 = np.ascontiguousarray(cam1, np.float32), np.ascontiguousarray(cam2, np.float32)
    kpL, kpR = np.ascontiguousarray(kpL), np.ascontiguousarray(kpR)
    descL, descR = np.ascontiguousarray(descL, np.uint8), np.ascontiguousarray(descR, np.uint8)
    R12, t12 = np.ascontiguousarray(R12, np.float32).reshape(9), np.ascontiguousarray(t12, np.float32).reshape(3)
    sigma2 = np.ascontiguousarray(sigma2, np.float32)
    nL, nR = len(kpL), len(kpR)
    l2r, r2l = np.zeros(max(nL, 1), np.int32), np.zeros(max(nR, 1), np.int32)
    depth, p3d = np.zeros(max(nL, 1), np.float32), np.zeros((max(nL, 1), 3), np.float32)
    n = ref_frame().orbslam3_frame_fisheye_matches(_p(cam1), _p(cam2), _p(kpL), _p(descL), nL, mono_left, _p(kpR), _p(descR), nR,
                                                   mono_right, _p(R12), _p(t12), _p(sigma2), len(sigma2), _p(l2r), _p(r2l), _p(depth), _p(p3d))
    return l2r[:nL], r2l[:nR], depth[:nL], p3d[:nL], n


def ref_undistort_keypoints(kps, K4, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:730-763), the reference's own object code (cv::undistortPoints = the restatement
    that tests/test_oracle_vs_cv2.py holds to cv2)."""
    kps = np.ascontiguousarray(kps)
    K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(kps)
    ref_frame().orbslam3_frame_undistort_keypoints(_p(kps), len(kps), _p(K4), _p(dist), len(dist), _p(out))
    return out


def ref_kb8_project(cam, p3d):
    cam, p3d = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(p3d, np.float32)
    uv = np.zeros((len(p3d), 2), np.float32)
    ref_frame().orbslam3_kb8_project(_p(cam), _p(p3d), len(p3d), _p(uv))
    return uv


def ref_kb8_unproject(cam, uv):
    cam, uv = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(uv, np.float32)
    ray = np.zeros((len(uv), 3), np.float32)
    ref_frame().orbslam3_kb8_unproject(_p(cam), _p(uv), len(uv), _p(ray))
    return ray


def ref_features_in_area(F, windows):
    """Frame::AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:380-411, 640-719), the reference's own object code.
    windows: (nq, 6) float32 rows (x, y, r, minLevel, maxLevel, bRight).  Returns a list of index arrays; right-camera
    indices are global (+ nleft) like the oracle's."""
    win = np.ascontiguousarray(windows, np.float32).reshape(-1, 6)
    cap = max(1, len(win) * max(F.n, 1))
    out, counts = np.zeros(cap, np.int32), np.zeros(max(len(win), 1), np.int32)
    total = ref_frame().orbslam3_frame_features_in_area(_p(F.kps), F.n, F.nleft, float(F.min_x), float(F.min_y), float(F.grid_w_inv),
                                                        float(F.grid_h_inv), _p(win), len(win), _p(out), cap, _p(counts))
    assert total <= cap
    res, o = [], 0
    for q in range(len(win)):
        idx = out[o:o + counts[q]].copy()
        if win[q, 5] != 0:
            idx += F.nleft
        res.append(idx)
        o += counts[q]
    return res


def ref_pinhole_project(K4, p3d):
    K4, p3d = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(p3d, np.float32)
    uv = np.zeros((len(p3d), 2), np.float32)
    ref_frame().orbslam3_pinhole_project(_p(K4), _p(p3d), len(p3d), _p(uv))
    return uv


def ref_stereo_frame(im_left, im_right, params, K4, dist, bf, th_depth=40.0):
    """The reference's stereo Frame constructor (src/Frame.cc:98-200) on two images, its own object code end to end: two
    extraction threads, UndistortKeyPoints, ComputeStereoMatches, ComputeImageBounds.  Returns a dict."""
    import orbfe
    nf, sf, nl, ini, mn = params
    imL, imR = np.ascontiguousarray(im_left, np.uint8), np.ascontiguousarray(im_right, np.uint8)
    h, w = imL.shape
    K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
    cap = nf + 64
    kps, kun, kr = (np.zeros(cap, orbfe.KEYPOINT_DTYPE) for _ in range(3))
    desc, descr = np.zeros((cap, 32), np.uint8), np.zeros((cap, 32), np.uint8)
    u, d = np.zeros(cap, np.float32), np.zeros(cap, np.float32)
    n, bounds = np.zeros(2, np.int32), np.zeros(4, np.float32)
    rc = ref_frame().orbslam3_frame_stereo_construct(_p(imL), _p(imR), w, h, nf, sf, nl, ini, mn, _p(K4), _p(dist), len(dist), bf, th_depth,
                                                     cap, _p(kps), _p(kun), _p(desc), _p(u), _p(d), _p(kr), _p(descr), _p(n), _p(bounds))
    assert rc == 0, (rc, n)
    N, NR = int(n[0]), int(n[1])
    return dict(kps=kps[:N], kps_un=kun[:N], desc=desc[:N], u_right=u[:N], depth=d[:N], kps_r=kr[:NR], desc_r=descr[:NR], bounds=bounds)


def ref_fisheye_frame(im_left, im_right, params, cam1, cam2, lap_left, lap_right, Tlr, bf=0.0, th_depth=40.0):
    """The reference's stereo-fisheye Frame constructor (src/Frame.cc:1017-1114) on two images, its own object code."""
    import orbfe
    nf, sf, nl, ini, mn = params
    imL, imR = np.ascontiguousarray(im_left, np.uint8), np.ascontiguousarray(im_right, np.uint8)
    h, w = imL.shape
    cam1, cam2 = np.ascontiguousarray(cam1, np.float32), np.ascontiguousarray(cam2, np.float32)
    lapL, lapR = np.ascontiguousarray(lap_left, np.int32), np.ascontiguousarray(lap_right, np.int32)
    Tlr = np.ascontiguousarray(Tlr, np.float32).reshape(3, 4)
    cap = 2 * (nf + 64)
    kps, desc = np.zeros(cap, orbfe.KEYPOINT_DTYPE), np.zeros((cap, 32), np.uint8)
    n = np.zeros(4, np.int32)
    l2r, r2l = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
    depth, p3d = np.zeros(cap, np.float32), np.zeros((cap, 3), np.float32)
    rc = ref_frame().orbslam3_frame_fisheye_construct(_p(imL), _p(imR), w, h, nf, sf, nl, ini, mn, _p(cam1), _p(cam2), _p(lapL), _p(lapR),
                                                      _p(Tlr), bf, th_depth, cap, _p(kps), _p(desc), _p(n), _p(l2r), _p(r2l), _p(depth), _p(p3d))
    assert rc == 0, (rc, n)
    NL, NR = int(n[0]), int(n[1])
    return dict(kps=kps[:NL + NR], desc=desc[:NL + NR], nleft=NL, nright=NR, mono_left=int(n[2]), mono_right=int(n[3]),
                l2r=l2r[:NL], r2l=r2l[:NR], depth=depth[:NL], p3d=p3d[:NL])


def ref_mono_frame(image, params, K4, dist, voc_path=None):
    """The reference's monocular Frame constructor (src/Frame.cc:284-378) + Frame::ComputeBoW (:721-728), its own object code
    (ORBextractor.cc, Frame.cc, Pinhole.cpp, DBoW2).  Returns a dict; bow = (words, values, node ids, offsets, indices)."""
    import orbfe
    nf, sf, nl, ini, mn = params
    im = np.ascontiguousarray(image, np.uint8)
    h, w = im.shape
    K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
    cap = nf + 64
    kps, kun, desc = np.zeros(cap, orbfe.KEYPOINT_DTYPE), np.zeros(cap, orbfe.KEYPOINT_DTYPE), np.zeros((cap, 32), np.uint8)
    n = np.zeros(4, np.int32)
    words, values = np.zeros(cap, np.uint32), np.zeros(cap, np.float64)
    nodes, off, idx = np.zeros(cap, np.uint32), np.zeros(cap + 1, np.int32), np.zeros(cap, np.int32)
    rc = ref_frame().orbslam3_frame_mono_construct(_p(im), w, h, nf, sf, nl, ini, mn, _p(K4), _p(dist), len(dist),
                                                   None if voc_path is None else str(voc_path).encode(), cap, _p(kps), _p(kun), _p(desc),
                                                   _p(n), _p(words), _p(values), _p(nodes), _p(off), _p(idx))
    assert rc == 0, (rc, n)
    N, nw, nn = int(n[0]), int(n[2]), int(n[3])
    return dict(kps=kps[:N], kps_un=kun[:N], desc=desc[:N], mono=int(n[1]),
                bow=(words[:nw], values[:nw], nodes[:nn], off[:nn + 1], idx[:off[nn]]))


def _pair_args(cam1, cam2, kp1, kp2, R12, t12, sigma, unc):
    f = lambda a: np.ascontiguousarray(a, np.float32)
    return f(cam1), f(cam2), np.ascontiguousarray(kp1), np.ascontiguousarray(kp2), f(R12).reshape(9), f(t12).reshape(3), f(sigma), f(unc)


def ref_kb8_epipolar_constrain(cam1, cam2, kp1, kp2, R12, t12, sigma, unc):
    """KannalaBrandt8::epipolarConstrain_ / TriangulateMatches_ (KannalaBrandt8.cpp:409-480), the reference's own object code.
    Returns (ok, depth, p3d)."""
    cam1, cam2, kp1, kp2, R12, t12, sigma, unc = _pair_args(cam1, cam2, kp1, kp2, R12, t12, sigma, unc)
    n = len(kp1)
    ok, depth, p3d = np.zeros(n, np.uint8), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
    ref_frame().orbslam3_kb8_epipolar_constrain_(_p(cam1), _p(cam2), _p(kp1), _p(kp2), n, _p(R12), _p(t12), _p(sigma), _p(unc), _p(ok),
                                                 _p(depth), _p(p3d))
    return ok.astype(bool), depth, p3d


def kb8_triangulate_pairs(cam1, cam2, kp1, kp2, R12, t12, sigma, unc):
    """The oracle's per-pair KB8 predicate (oracle/cpu_ref/kb8_ref.cc).  Returns (depth, p3d, margins)."""
    cam1, cam2, kp1, kp2, R12, t12, sigma, unc = _pair_args(cam1, cam2, kp1, kp2, R12, t12, sigma, unc)
    n = len(kp1)
    depth, p3d, mg = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
    L_ = lib()
    L_.orbref_kb8_triangulate_pairs.argtypes = [_P] * 4 + [_I] + [_P] * 7
    L_.orbref_kb8_triangulate_pairs(_p(cam1), _p(cam2), _p(kp1), _p(kp2), n, _p(R12), _p(t12), _p(sigma), _p(unc), _p(depth), _p(p3d), _p(mg))
    return depth, p3d, mg
