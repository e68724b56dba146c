"""ORBmatcher — Python mirror of ORB_SLAM3::ORBmatcher (include/ORBmatcher.h:35-111 in the ORB-SLAM3
tree) and of Frame's two stereo matchers (Frame.cc:794-964, 1116-1156), over the C ABI.

Objects the reference reaches through KeyFrame / Frame / MapPoint pointers are passed as arrays:
descriptors (n, 32) uint8, keypoints as the cv::KeyPoint-layout structured array the extractor
returns, a DBoW2::FeatureVector as `FeatureVectorCSR`, "has a MapPoint" as a bool/uint8 mask.
Every method runs on the GPU; there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from ._lib import KEYPOINT_DTYPE, OrbfeError, lib
from ._matcher_proto import BowSide, Epipolar, GridFrameC, Kb8Epipolar, FeatureVector


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class FeatureVectorCSR:
    """DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened to CSR."""

    def __init__(self, node_ids, offsets, indices):
        self.node_ids = np.ascontiguousarray(node_ids, np.uint32)
        self.offsets = np.ascontiguousarray(offsets, np.int32)
        self.indices = np.ascontiguousarray(indices, np.int32)
        assert len(self.offsets) == len(self.node_ids) + 1

    @classmethod
    def from_node_of_feature(cls, node_of_feature):
        """node_of_feature[i] = vocabulary node (at levelsup) of feature i; features are appended to their
        node in ascending index order, as DBoW2::FeatureVector::addFeature does (FeatureVector.cpp:31-45)."""
        node_of_feature = np.asarray(node_of_feature)
        order = np.argsort(node_of_feature, kind="stable")
        ids, counts = np.unique(node_of_feature, return_counts=True)
        return cls(ids, np.concatenate([[0], np.cumsum(counts)]), order)

    @classmethod
    def from_dict(cls, d):
        ids = sorted(d)
        off = np.concatenate([[0], np.cumsum([len(d[k]) for k in ids])]) if ids else np.zeros(1)
        idx = np.concatenate([np.asarray(d[k], np.int32) for k in ids]) if ids else np.zeros(0, np.int32)
        return cls(ids, off, idx)

    def to_dict(self):
        return {int(k): self.indices[self.offsets[i]:self.offsets[i + 1]].tolist() for i, k in enumerate(self.node_ids)}

    def c_struct(self):
        return FeatureVector(len(self.node_ids), _p(self.node_ids), _p(self.offsets), _p(self.indices))


def measure_popc_peak(device=0):
    """Sustained POPC instructions/s of the device (the matching roofline's denominator is this / 8)."""
    v = C.c_double(0)
    st = lib().orbfe_measure_popc_peak(int(device), C.byref(v))
    if st < 0:
        raise OrbfeError(st, "orbfe_measure_popc_peak")
    return v.value


class _Side:
    """Keeps the arrays behind an orbfe_bow_side alive."""

    def __init__(self, desc, fv, mask, kps):
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.fv = fv
        self.mask = None if mask is None else np.ascontiguousarray(mask).astype(np.uint8)
        self.kps = None if kps is None else np.ascontiguousarray(kps, KEYPOINT_DTYPE)
        assert self.mask is None or len(self.mask) == len(self.desc)
        assert self.kps is None or len(self.kps) == len(self.desc)

    def c_struct(self):
        return BowSide(_p(self.desc), len(self.desc), self.fv.c_struct(), _p(self.mask), _p(self.kps))


class ORBmatcher:
    TH_LOW = 50
    TH_HIGH = 100
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self._lib = lib()
        self._m = C.c_void_p()
        st = self._lib.orbfe_matcher_create(float(nnratio), int(bool(checkOri)), int(device), C.byref(self._m))
        if st != 0:
            self._m = C.c_void_p()
            raise OrbfeError(st, "orbfe_matcher_create")

    @property
    def handle(self):
        return self._m

    def close(self):
        if getattr(self, "_m", None) and self._m.value:
            self._lib.orbfe_matcher_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self):
        return self._lib.orbfe_matcher_stream(self._m)

    def synchronize(self):
        st = self._lib.orbfe_matcher_synchronize(self._m)
        if st < 0:
            raise OrbfeError(st, "orbfe_matcher_synchronize")

    def launch_count(self):
        return int(self._lib.orbfe_matcher_launch_count(self._m))

    @staticmethod
    def DescriptorDistance(a, b):
        """ORBmatcher::DescriptorDistance (ORBmatcher.cc:2561-2577) for one pair of host descriptors: a host
        function in the reference too (orbfe_descriptor_distance in include/orbfe.h is its C form)."""
        x = np.bitwise_xor(np.asarray(a, np.uint8).reshape(32), np.asarray(b, np.uint8).reshape(32))
        return int(np.unpackbits(x).sum())

    # ---- brute force (cv::BFMatcher(NORM_HAMMING).knnMatch(k=2), Frame.cc:1134)
    KNN2_POPC, KNN2_TENSOR = 0, 1

    def set_knn2_backend(self, backend):
        """KNN2_POPC (integer pipe) or KNN2_TENSOR (tcgen05 kind::i8): identical results."""
        st = self._lib.orbfe_matcher_set_knn2_backend(self._m, int(backend))
        if st < 0:
            raise OrbfeError(st, "orbfe_matcher_set_knn2_backend")

    def knn2(self, q, t):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
        idx = np.full((len(q), 2), -1, np.int32)
        dist = np.full((len(q), 2), -1, np.int32)
        st = self._lib.orbfe_hamming_knn2(self._m, _p(q), len(q), _p(t), len(t), _p(idx), _p(dist))
        if st < 0:
            raise OrbfeError(st, "orbfe_hamming_knn2")
        return idx, dist

    def knn2_batch_device(self, d_q, q_stride, d_nq, nq_max, d_t, t_stride, d_nt, nt_max, batch, d_idx2, d_dist2):
        """Raw device pointers (ints); asynchronous on the matcher's stream."""
        st = self._lib.orbfe_hamming_knn2_batch_device(self._m, C.c_void_p(d_q), q_stride, C.c_void_p(d_nq) if d_nq else None,
                                                       nq_max, C.c_void_p(d_t), t_stride, C.c_void_p(d_nt) if d_nt else None,
                                                       nt_max, batch, C.c_void_p(d_idx2), C.c_void_p(d_dist2))
        if st < 0:
            raise OrbfeError(st, "orbfe_hamming_knn2_batch_device")

    def match_lapping(self, desc_left, desc_right, ratio=0.7):
        """Frame::ComputeStereoFishEyeMatches up to the Lowe-ratio test (Frame.cc:1134-1141)."""
        q = np.ascontiguousarray(desc_left, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(desc_right, np.uint8).reshape(-1, 32)
        match = np.full(len(q), -1, np.int32)
        n = self._lib.orbfe_match_lapping(self._m, _p(q), len(q), _p(t), len(t), float(ratio), _p(match))
        if n < 0:
            raise OrbfeError(n, "orbfe_match_lapping")
        return match, n

    def ComputeStereoFishEyeMatches(self, kps_left, desc_left, mono_left, kps_right, desc_right, mono_right, cam1, cam2, R12,
                                    t12, sigma2):
        """Frame::ComputeStereoFishEyeMatches (Frame.cc:1116-1156).  Returns (mvLeftToRightMatch,
        mvRightToLeftMatch, mvDepth, mvStereo3Dpoints (n, 3), nMatches)."""
        kL, kR = np.ascontiguousarray(kps_left, KEYPOINT_DTYPE), np.ascontiguousarray(kps_right, KEYPOINT_DTYPE)
        dL = np.ascontiguousarray(desc_left, np.uint8).reshape(-1, 32)
        dR = np.ascontiguousarray(desc_right, np.uint8).reshape(-1, 32)
        f = lambda a, n: np.ascontiguousarray(a, np.float32).reshape(n)
        cam1, cam2, R12, t12 = f(cam1, 8), f(cam2, 8), f(R12, 9), f(t12, 3)
        sigma2 = np.ascontiguousarray(sigma2, np.float32)
        l2r, r2l = np.full(max(len(kL), 1), -1, np.int32), np.full(max(len(kR), 1), -1, np.int32)
        depth, p3d = np.full(max(len(kL), 1), -1, np.float32), np.zeros((max(len(kL), 1), 3), np.float32)
        n = self._lib.orbfe_fisheye_stereo_match(self._m, _p(kL), _p(dL), len(kL), int(mono_left), _p(kR), _p(dR), len(kR),
                                                 int(mono_right), _p(cam1), _p(cam2), _p(R12), _p(t12), _p(sigma2), len(sigma2),
                                                 _p(l2r), _p(r2l), _p(depth), _p(p3d))
        if n < 0:
            raise OrbfeError(n, "orbfe_fisheye_stereo_match")
        return l2r[:len(kL)], r2l[:len(kR)], depth[:len(kL)], p3d[:len(kL)], n

    # ---- Frame::ComputeStereoMatches (Frame.cc:794-964)
    def stereo_match(self, left, right, bf, b, frame_l=0, frame_r=0):
        """left / right: ORBextractor objects whose last extraction holds the two views (the same object
        with frame_l=0, frame_r=1 after a batch extract of the pair).  Returns (mvuRight, mvDepth, n)."""
        cap = left.cap
        u = np.full(cap, -1, np.float32)
        d = np.full(cap, -1, np.float32)
        n = self._lib.orbfe_stereo_match(self._m, left.handle, frame_l, right.handle, frame_r, float(bf), float(b), _p(u), _p(d),
                                         cap)
        if n < 0:
            raise OrbfeError(n, "orbfe_stereo_match")
        return u, d, n

    def stereo_match_batch_device(self, extractor, npairs, bf, b, d_u_right, d_depth):
        st = self._lib.orbfe_stereo_match_batch_device(self._m, extractor.handle, npairs, float(bf), float(b),
                                                       C.c_void_p(d_u_right), C.c_void_p(d_depth))
        if st < 0:
            raise OrbfeError(st, "orbfe_stereo_match_batch_device")

    # ---- SearchByBoW
    def _search_bow(self, form, sidesA, sidesB, shared, nleft):
        n = len(sidesA)
        arrA = (BowSide * n)(*[s.c_struct() for s in sidesA])
        arrB = (BowSide * len(sidesB))(*[s.c_struct() for s in sidesB])
        a2b = [np.full(len(s.desc), -1, np.int32) for s in sidesA]
        b2a = [np.full(len(sidesB[0 if shared else i].desc), -1, np.int32) for i in range(n)]
        pa = (C.c_void_p * n)(*[a.ctypes.data for a in a2b])
        pb = (C.c_void_p * n)(*[a.ctypes.data for a in b2a])
        cnt = np.zeros(n, np.int32)
        st = self._lib.orbfe_search_bow(self._m, form, arrA, arrB, n, int(shared), int(nleft), pa, pb, _p(cnt))
        if st < 0:
            raise OrbfeError(st, "orbfe_search_bow")
        return a2b, b2a, cnt

    def SearchByBoW_KF_F(self, kf_desc, kf_fv, kf_has_mp, kf_kps, f_desc, f_fv, f_kps, nleft=-1):
        """SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (ORBmatcher.cc:269-471).  Returns
        (matches per Frame feature = keyframe feature index or -1, nmatches)."""
        a2b, b2a, cnt = self._search_bow(0, [_Side(kf_desc, kf_fv, kf_has_mp, kf_kps)], [_Side(f_desc, f_fv, None, f_kps)], False,
                                         nleft)
        return b2a[0], int(cnt[0])

    def SearchByBoW_KFs_F(self, kfs, f_desc, f_fv, f_kps, nleft=-1):
        """The same search for many keyframes against one frame in one call (Tracking::Relocalization's
        loop, Tracking.cc:3495): kfs = [(desc, fv, has_mp, kps), ...].  Returns ([matches], [nmatches])."""
        a2b, b2a, cnt = self._search_bow(0, [_Side(*k) for k in kfs], [_Side(f_desc, f_fv, None, f_kps)], True, nleft)
        return b2a, cnt.tolist()

    def SearchByBoW_KF_KF(self, desc1, fv1, has_mp1, kps1, desc2, fv2, has_mp2, kps2):
        """SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (ORBmatcher.cc:823-963).  Returns
        (vpMatches12 as index into KF2 or -1, nmatches)."""
        a2b, b2a, cnt = self._search_bow(1, [_Side(desc1, fv1, has_mp1, kps1)], [_Side(desc2, fv2, has_mp2, kps2)], False, -1)
        return a2b[0], int(cnt[0])

    # ---- SearchForTriangulation (pinhole)
    def SearchForTriangulation(self, desc1, fv1, has_mp1, kps1, u_right1, desc2, fv2, has_mp2, kps2, u_right2, F12, epipole,
                               sigma2, scale, bOnlyStereo=False, bCoarse=False):
        """ORBmatcher::SearchForTriangulation_ (ORBmatcher.cc:1208-1432) for pinhole cameras.  Returns
        (vMatchedPairs as an (n, 2) array ascending in idx1, nmatches)."""
        A, B = _Side(desc1, fv1, has_mp1, kps1), _Side(desc2, fv2, has_mp2, kps2)
        sa, sb = A.c_struct(), B.c_struct()
        sigma2 = np.ascontiguousarray(sigma2, np.float32)
        scale = np.ascontiguousarray(scale, np.float32)
        geo = Epipolar()
        F12 = np.asarray(F12, np.float32).reshape(9)
        for i in range(9):
            geo.F12[i] = float(F12[i])
        geo.ex, geo.ey = float(epipole[0]), float(epipole[1])
        geo.sigma2_b, geo.scale_b, geo.nlevels = _p(sigma2), _p(scale), len(sigma2)
        u1 = None if u_right1 is None else np.ascontiguousarray(u_right1, np.float32)
        u2 = None if u_right2 is None else np.ascontiguousarray(u_right2, np.float32)
        m12 = np.full(len(A.desc), -1, np.int32)
        n = self._lib.orbfe_search_triangulation(self._m, C.byref(sa), _p(u1), C.byref(sb), _p(u2), C.byref(geo),
                                                 int(bool(bOnlyStereo)), int(bool(bCoarse)), _p(m12))
        if n < 0:
            raise OrbfeError(n, "orbfe_search_triangulation")
        i1 = np.nonzero(m12 >= 0)[0]
        return np.stack([i1, m12[i1]], axis=1).astype(np.int64), n, m12

    def UndistortKeyPoints(self, kps, K4, dist):
        """Frame::UndistortKeyPoints (Frame.cc:730-763).  K4 = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
        k = np.ascontiguousarray(kps, KEYPOINT_DTYPE)
        K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
        out = np.zeros_like(k)
        st = self._lib.orbfe_undistort_keypoints(self._m, _p(k), len(k), _p(K4), _p(dist), len(dist), _p(out))
        if st < 0:
            raise OrbfeError(st, "orbfe_undistort_keypoints")
        return out

    def ComputeDistinctiveDescriptors(self, desc, offsets):
        """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:349-423) for many map points: rows
        offsets[p]..offsets[p+1] of desc are point p's observed descriptors.  Returns the chosen row per point
        (relative to offsets[p]; -1 without observations)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        off = np.ascontiguousarray(offsets, np.int32)
        best = np.full(max(len(off) - 1, 1), -1, np.int32)
        st = self._lib.orbfe_distinctive_descriptors(self._m, _p(d), _p(off), len(off) - 1, _p(best))
        if st < 0:
            raise OrbfeError(st, "orbfe_distinctive_descriptors")
        return best[:len(off) - 1]

    def search_by_projection(self, F, queries, mode, th=TH_HIGH, link_stereo=False, rot_hist=False):
        """The window-search loop of ORBmatcher::SearchByProjection (ORBmatcher.cc:44-236, :2173-2389, :2391-2510) on a
        projection.GridFrame and an array of projection.PROJ_QUERY_DTYPE records (see orbfe/projection.py for how
        each overload becomes windows).  Returns (assign [F.n], choice [len(queries)], nmatches)."""
        from .projection import PROJ_QUERY_DTYPE
        q = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE)
        g = GridFrameC()
        g.kps, g.desc, g.n, g.nleft = _p(F.kps), _p(F.desc), F.n, F.nleft
        g.blocked, g.u_right = _p(F.blocked), _p(F.u_right)
        g.left_to_right, g.right_to_left = _p(F.left_to_right), _p(F.right_to_left)
        g.min_x, g.min_y, g.grid_w_inv, g.grid_h_inv = float(F.min_x), float(F.min_y), float(F.grid_w_inv), float(F.grid_h_inv)
        sig = getattr(F, "inv_level_sigma2", None)
        g.inv_level_sigma2, g.n_levels = _p(sig), 0 if sig is None else len(sig)
        assign, choice = np.full(max(F.n, 1), -1, np.int32), np.full(max(len(q), 1), -1, np.int32)
        n = self._lib.orbfe_search_by_projection(self._m, C.byref(g), _p(q), len(q), int(mode), int(th), int(bool(link_stereo)),
                                                 int(bool(rot_hist)), _p(assign), _p(choice))
        if n < 0:
            raise OrbfeError(n, "orbfe_search_by_projection")
        return assign[:F.n], choice[:len(q)], n

    def SearchForInitialization(self, kps1, desc1, F2, vbPrevMatched, windowSize=10):
        """ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (ORBmatcher.cc:706-821).
        kps1 / desc1 = F1.mvKeysUn / F1.mDescriptors, F2 = projection.GridFrame of the second frame, vbPrevMatched
        (n1, 2) float32.  Returns (nmatches, vnMatches12, updated vbPrevMatched)."""
        from ._lib import KEYPOINT_DTYPE
        k1 = np.ascontiguousarray(kps1, KEYPOINT_DTYPE)
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        prev = np.array(vbPrevMatched, np.float32).reshape(-1, 2).copy()
        if len(d1) != len(k1) or len(prev) != len(k1):
            raise ValueError("SearchForInitialization: kps1, desc1 and vbPrevMatched must have one row per F1 feature")
        g = GridFrameC()
        g.kps, g.desc, g.n, g.nleft = _p(F2.kps), _p(F2.desc), F2.n, -1
        g.min_x, g.min_y, g.grid_w_inv, g.grid_h_inv = float(F2.min_x), float(F2.min_y), float(F2.grid_w_inv), float(F2.grid_h_inv)
        m12 = np.full(max(len(k1), 1), -1, np.int32)
        n = self._lib.orbfe_search_for_initialization(self._m, _p(k1), _p(d1), len(k1), C.byref(g), _p(prev), int(windowSize), _p(m12))
        if n < 0:
            raise OrbfeError(n, "orbfe_search_for_initialization")
        return n, m12[:len(k1)], prev

    def SearchForTriangulationKB8(self, desc1, fv1, has_mp1, kps1, desc2, fv2, has_mp2, kps2, cam_a, cam_b, nleft_a, nleft_b, R12,
                                  t12, epipole, sigma2_a, sigma2_b, scale_b, bOnlyStereo=False, bCoarse=False):
        """ORBmatcher::SearchForTriangulation_ (ORBmatcher.cc:1208-1432) for Kannala-Brandt keyframes
        (KannalaBrandt8::epipolarConstrain_).  cam_*: (2, 8) = mpCamera, mpCamera2; R12 (4, 9) / t12 (4, 3) for the
        pairings ll, lr, rl, rr.  Returns (vMatchedPairs (n, 2), nmatches, vMatches12, points (n1, 3))."""
        A, B = _Side(desc1, fv1, has_mp1, kps1), _Side(desc2, fv2, has_mp2, kps2)
        sa, sb = A.c_struct(), B.c_struct()
        f = lambda a: np.ascontiguousarray(a, np.float32)
        sigma2_a, sigma2_b, scale_b = f(sigma2_a), f(sigma2_b), f(scale_b)
        geo = Kb8Epipolar()
        for dst, src, n in ((geo.cam_a, cam_a, 16), (geo.cam_b, cam_b, 16), (geo.R12, R12, 36), (geo.t12, t12, 12)):
            src = f(src).reshape(-1)
            if src.size != n:
                raise ValueError("cam_* must hold 2 x 8, R12 4 x 9 and t12 4 x 3 floats")
            for i in range(n):
                dst[i] = float(src[i])
        geo.nleft_a, geo.nleft_b = int(nleft_a), int(nleft_b)
        geo.ex, geo.ey = float(epipole[0]), float(epipole[1])
        geo.sigma2_a, geo.sigma2_b, geo.scale_b, geo.nlevels = _p(sigma2_a), _p(sigma2_b), _p(scale_b), len(sigma2_b)
        m12 = np.full(max(len(A.desc), 1), -1, np.int32)
        pts = np.zeros((max(len(A.desc), 1), 3), np.float32)
        n = self._lib.orbfe_search_triangulation_kb8(self._m, C.byref(sa), C.byref(sb), C.byref(geo), int(bool(bOnlyStereo)),
                                                     int(bool(bCoarse)), _p(m12), _p(pts))
        if n < 0:
            raise OrbfeError(n, "orbfe_search_triangulation_kb8")
        m12, pts = m12[:len(A.desc)], pts[:len(A.desc)]
        i1 = np.nonzero(m12 >= 0)[0]
        return np.stack([i1, m12[i1]], axis=1).astype(np.int64), n, m12, pts
