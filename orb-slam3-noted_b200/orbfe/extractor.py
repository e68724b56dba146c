"""ORBextractor — Python mirror of ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109) over the C ABI.

Same constructor arguments and the same call: `mono_index, keypoints, descriptors =
extractor(image, None, [lap0, lap1])`, where keypoints is a structured array with cv::KeyPoint's
layout and descriptors an (n, 32) uint8 matrix.  `mvImagePyramid` is available through
`image_pyramid(level)` (bordered level, like the reference's cv::Mat parent buffers).
"""
import ctypes as C

import numpy as np

from ._lib import KEYPOINT_DTYPE, OrbfeError, Params, lib, ptr


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0, max_width=1920,
                 max_height=1080, max_batch=1):
        self._h = C.c_void_p()
        self._lib = lib()
        p = Params(int(nfeatures), float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST), int(device),
                   int(max_width), int(max_height), int(max_batch))
        st = self._lib.orbfe_create(C.byref(p), C.byref(self._h))
        if st != 0:
            self._h = C.c_void_p()
            raise OrbfeError(st, "orbfe_create")
        self.nfeatures, self.nlevels, self.max_batch = int(nfeatures), int(nlevels), int(max_batch)
        self.cap = self._lib.orbfe_max_keypoints(self._h)
        nl = self.nlevels
        self._scale = np.zeros(nl, np.float32)
        self._inv_scale = np.zeros(nl, np.float32)
        self._sigma2 = np.zeros(nl, np.float32)
        self._inv_sigma2 = np.zeros(nl, np.float32)
        self._per_level = np.zeros(nl, np.int32)
        self._lib.orbfe_get_tables(self._h, ptr(self._scale), ptr(self._inv_scale), ptr(self._sigma2),
                                   ptr(self._inv_sigma2), ptr(self._per_level))
        self._scale_factor = float(np.float32(scaleFactor))

    # ---- getters of ORBextractor.h:62-81
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self._scale_factor

    def GetScaleFactors(self):
        return self._scale.copy()

    def GetInverseScaleFactors(self):
        return self._inv_scale.copy()

    def GetScaleSigmaSquares(self):
        return self._sigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self._inv_sigma2.copy()

    def features_per_level(self):
        return self._per_level.copy()

    @property
    def handle(self):
        return self._h

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.orbfe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    FORMATS = {"gray": 0, "rgb": 1, "bgr": 2, "rgba": 3, "bgra": 4}

    def set_input_format(self, fmt):
        """Colour input: the cv::cvtColor(..2GRAY) of Tracking::GrabImage* (Tracking.cc:1233-1258) runs inside the
        pyramid's level-0 kernel.  fmt: "gray", "rgb", "bgr", "rgba", "bgra"."""
        st = self._lib.orbfe_set_input_format(self._h, self.FORMATS[fmt])
        if st < 0:
            raise OrbfeError(st, "orbfe_set_input_format")
        self._bpp = 1 if fmt == "gray" else 3 if fmt in ("rgb", "bgr") else 4

    # ---- operator()
    def __call__(self, image, mask=None, vLappingArea=(0, 0)):
        """Returns (monoIndex, keypoints, descriptors); monoIndex == -1 for an empty image, like the reference."""
        if image is None or image.size == 0:
            return -1, np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        bpp = getattr(self, "_bpp", 1)
        if bpp == 1:
            assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 image expected (ORBextractor.cc:1076)"
            assert image.strides[1] == 1
        else:
            assert image.dtype == np.uint8 and image.ndim == 3 and image.shape[2] == bpp and image.strides[2] == 1 and \
                image.strides[1] == bpp
        h, w = image.shape[:2]
        kps = np.zeros(self.cap, KEYPOINT_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        st = self._lib.orbfe_extract(self._h, C.c_void_p(image.ctypes.data), w, h, image.strides[0],
                                     int(vLappingArea[0]), int(vLappingArea[1]), ptr(kps), ptr(desc), self.cap, C.byref(n))
        if st < -1:
            raise OrbfeError(st, "orbfe_extract")
        if st == -1:
            return -1, kps[:0], desc[:0]
        return st, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, vLappingArea=(0, 0)):
        """images: (B, H, W) uint8 host array.  Returns a list of (monoIndex, keypoints, descriptors)."""
        images = np.ascontiguousarray(images)
        assert images.dtype == np.uint8 and images.ndim == 3
        B, h, w = images.shape
        kps = np.zeros((B, self.cap), KEYPOINT_DTYPE)
        desc = np.zeros((B, self.cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        mono = np.zeros(B, np.int32)
        st = self._lib.orbfe_extract_batch(self._h, ptr(images), B, w, h, w, w * h, int(vLappingArea[0]),
                                           int(vLappingArea[1]), ptr(kps), ptr(desc), self.cap, ptr(n), ptr(mono))
        if st < 0:
            raise OrbfeError(st, "orbfe_extract_batch")
        return [(int(mono[b]), kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def extract_batch_device(self, d_ptr, batch, width, height, stride=None, frame_stride=None, vLappingArea=(0, 0)):
        """Enqueue extraction of frames already in device memory (asynchronous on the handle's stream)."""
        stride = width if stride is None else stride
        frame_stride = stride * height if frame_stride is None else frame_stride
        st = self._lib.orbfe_extract_batch_device(self._h, C.c_void_p(d_ptr), batch, width, height, stride, frame_stride,
                                                  int(vLappingArea[0]), int(vLappingArea[1]))
        if st < 0:
            raise OrbfeError(st, "orbfe_extract_batch_device")

    def fetch_results(self, batch):
        kps = np.zeros((batch, self.cap), KEYPOINT_DTYPE)
        desc = np.zeros((batch, self.cap, 32), np.uint8)
        n = np.zeros(batch, np.int32)
        mono = np.zeros(batch, np.int32)
        st = self._lib.orbfe_fetch_results(self._h, batch, ptr(kps), ptr(desc), self.cap, ptr(n), ptr(mono))
        if st < 0:
            raise OrbfeError(st, "orbfe_fetch_results")
        return [(int(mono[b]), kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(batch)]

    def device_results(self):
        """(d_kps, d_desc, d_n, d_mono, cap) raw device pointers of the last extract."""
        a, b, c, d = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        cap = C.c_int()
        st = self._lib.orbfe_device_results(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d), C.byref(cap))
        if st < 0:
            raise OrbfeError(st, "orbfe_device_results")
        return a.value, b.value, c.value, d.value, cap.value

    def image_pyramid(self, level, frame=0):
        """mvImagePyramid[level] with its 19-px border: returns (bordered array, roi view)."""
        w, h = C.c_int(), C.c_int()
        st = self._lib.orbfe_get_pyramid_level(self._h, frame, level, None, 0, C.byref(w), C.byref(h))
        if st < 0:
            raise OrbfeError(st, "orbfe_get_pyramid_level")
        buf = np.zeros((h.value + 38, w.value + 38), np.uint8)
        st = self._lib.orbfe_get_pyramid_level(self._h, frame, level, ptr(buf), buf.strides[0], C.byref(w), C.byref(h))
        if st < 0:
            raise OrbfeError(st, "orbfe_get_pyramid_level")
        return buf, buf[19:19 + h.value, 19:19 + w.value]

    def candidates(self, level, frame=0):
        """FAST candidates of a level, (n, 3) int32 (x, y, response) in the reference's push order."""
        n = self._lib.orbfe_get_candidates(self._h, frame, level, None, 0)
        if n < 0:
            raise OrbfeError(n, "orbfe_get_candidates")
        out = np.zeros((max(n, 1), 3), np.int32)
        n = self._lib.orbfe_get_candidates(self._h, frame, level, ptr(out), out.shape[0])
        return out[:n]

    def synchronize(self):
        st = self._lib.orbfe_synchronize(self._h)
        if st < 0:
            raise OrbfeError(st, "orbfe_synchronize")

    def stream(self):
        return self._lib.orbfe_stream(self._h)

    def launch_count(self):
        return int(self._lib.orbfe_launch_count(self._h))

    def graph_launch_count(self):
        return int(self._lib.orbfe_graph_launch_count(self._h))

    def order_consumer(self, stream):
        """Work enqueued on `stream` (a matcher's or vocabulary's cudaStream_t) from now on waits for this extractor."""
        st = self._lib.orbfe_stream_wait_extractor(self._h, C.c_void_p(stream))
        if st < 0:
            raise OrbfeError(st, "orbfe_stream_wait_extractor")

    def wait_consumer(self, stream):
        """The next extraction waits for everything already enqueued on `stream` (it overwrites what that work reads)."""
        st = self._lib.orbfe_extractor_wait_stream(self._h, C.c_void_p(stream))
        if st < 0:
            raise OrbfeError(st, "orbfe_extractor_wait_stream")

    BLUR_ALU, BLUR_TENSOR = 0, 1

    def set_blur_backend(self, backend):
        """k_blur (integer ALU) or k_blur_tc (tcgen05 int8 GEMMs); same bytes either way."""
        st = self._lib.orbfe_set_blur_backend(self._h, int(backend))
        if st < 0:
            raise OrbfeError(st, "orbfe_set_blur_backend")

    def blurred_level(self, level, frame=0):
        """The blurred level the descriptors were sampled from (the reference's workingMat after GaussianBlur), ROI only."""
        w, h = C.c_int(), C.c_int()
        st = self._lib.orbfe_get_pyramid_level(self._h, frame, level, None, 0, C.byref(w), C.byref(h))
        if st < 0:
            raise OrbfeError(st, "orbfe_get_pyramid_level")
        buf = np.zeros((h.value, w.value), np.uint8)
        st = self._lib.orbfe_get_blurred_level(self._h, frame, level, ptr(buf), buf.strides[0])
        if st < 0:
            raise OrbfeError(st, "orbfe_get_blurred_level")
        return buf

    def set_pipeline(self, chunk_frames=0, compute_streams=0, use_graphs=-1):
        st = self._lib.orbfe_set_pipeline(self._h, int(chunk_frames), int(compute_streams), int(use_graphs))
        if st < 0:
            raise OrbfeError(st, "orbfe_set_pipeline")

    STAGES = ("pyramid", "fast", "octree", "assemble", "describe")

    def set_profiling(self, enabled=True):
        st = self._lib.orbfe_set_profiling(self._h, int(bool(enabled)))
        if st < 0:
            raise OrbfeError(st, "orbfe_set_profiling")

    def stage_times(self):
        """({stage: total ms since the last read}, extract calls covered); synchronises the stream."""
        ms = np.zeros(len(self.STAGES), np.float64)
        calls = C.c_longlong(0)
        st = self._lib.orbfe_get_stage_times(self._h, ptr(ms), C.byref(calls))
        if st < 0:
            raise OrbfeError(st, "orbfe_get_stage_times")
        return dict(zip(self.STAGES, ms.tolist())), int(calls.value)
