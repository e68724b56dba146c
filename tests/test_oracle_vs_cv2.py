"""CPU: known-answer tests of the oracle's OpenCV-free primitives (oracle/cpu_ref/cvprim.cc) against
the real OpenCV (cv2 4.13.0, the pinned version), and of the whole C++
extractor against the cv2-backed restatement (oracle/py/orb_cv2.py).  Skipped where cv2 is absent;
the golden-vector test covers that case."""
import os
import sys

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

import orbfe  # noqa: E402
import oracle_binding as ob  # noqa: E402
from oracle_binding import OrbRef, _p  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "py"))
from orb_cv2 import ORBextractorCv2  # noqa: E402


def rnd_img(seed, w, h):
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)


@pytest.mark.parametrize("sw,sh,dw,dh", [(752, 480, 627, 400), (627, 400, 522, 333), (252, 161, 210, 134),
                                         (1920, 1080, 1600, 900), (512, 512, 427, 427), (97, 61, 80, 51)])
def test_resize_linear(sw, sh, dw, dh):
    src = rnd_img(sw * 7 + dh, sw, sh)
    want = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
    got = np.zeros((dh, dw), np.uint8)
    ob.lib().orbref_resize(_p(src), sw, sh, sw, _p(got), dw, dh, dw)
    assert np.array_equal(got, want)


def test_border_reflect101():
    src = rnd_img(3, 70, 50)
    want = cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    buf = np.zeros_like(want)
    buf[19:-19, 19:-19] = src
    ob.lib().orbref_border101(_p(buf), 70, 50, 19, buf.strides[0])
    assert np.array_equal(buf, want)


@pytest.mark.parametrize("t", [7, 20])
def test_fast_9_16_nms(t):
    img = orbfe.synth.frame(5, 160, 120)
    want = cv2.FastFeatureDetector_create(t, True).detect(img)
    out = np.zeros((20000, 3), np.int32)
    n = ob.lib().orbref_fast(_p(img), 160, 120, 160, t, _p(out), len(out))
    assert n == len(want) and n > 50
    w = np.array([[k.pt[0], k.pt[1], k.response] for k in want]).astype(np.int32)
    assert np.array_equal(out[:n], w)
    # a 42x44 cell-sized ROI behaves the same (ORBextractor.cc:808 passes views of this size)
    cell = np.ascontiguousarray(img[30:74, 50:92])
    want = cv2.FastFeatureDetector_create(t, True).detect(cell)
    n = ob.lib().orbref_fast(_p(cell), 42, 44, 42, t, _p(out), len(out))
    assert n == len(want)
    if n:
        assert np.array_equal(out[:n], np.array([[k.pt[0], k.pt[1], k.response] for k in want]).astype(np.int32))


def test_gaussian_blur_7x7_fixed_point():
    for (w, h, seed) in [(160, 120, 1), (37, 41, 2), (752, 480, 3)]:
        src = rnd_img(seed, w, h)
        want = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        got = np.zeros_like(src)
        ob.lib().orbref_blur(_p(src), w, h, w, _p(got), w)
        assert np.array_equal(got, want)


def test_fast_atan2():
    rng = np.random.default_rng(11)
    ys = rng.integers(-200000, 200000, 20000).astype(np.float32)
    xs = rng.integers(-200000, 200000, 20000).astype(np.float32)
    ys[:4], xs[:4] = [0, 0, 5, -5], [0, 7, 0, 0]
    for y, x in zip(ys, xs):
        assert ob.lib().orbref_fast_atan2(float(y), float(x)) == np.float32(cv2.fastAtan2(float(y), float(x)))


def test_knn2_matches_bfmatcher_tie_rule():
    q = orbfe.synth.descriptors(1, 300)
    t = orbfe.synth.descriptors(2, 5000)
    t[100] = t[7]           # crafted ties: the lower train index must win (Appendix A.7)
    t[4000] = t[7]
    q[0] = t[7]
    idx, dist = ob.knn2(q, t)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
    want_i = np.array([[a.trainIdx, b.trainIdx] for a, b in m])
    want_d = np.array([[a.distance, b.distance] for a, b in m]).astype(np.int32)
    assert np.array_equal(idx, want_i) and np.array_equal(dist, want_d)
    assert idx[0].tolist() == [7, 100] and dist[0].tolist() == [0, 0]


@pytest.mark.parametrize("w,h,nf,lap,seed", [(752, 480, 1000, (0, 1000), 1003), (752, 480, 1200, (0, 0), 1104),
                                             (512, 512, 1500, (128, 511), 1205), (1241, 376, 2000, (0, 0), 1306),
                                             (640, 480, 500, (100, 300), 9)])
def test_cpp_oracle_equals_cv2_restatement(w, h, nf, lap, seed):
    img = orbfe.synth.frame(seed, w, h)
    py = ORBextractorCv2(nf, 1.2, 8, 20, 7)
    kp, desc, mono = py(img, lap)
    ref = OrbRef(nf, 1.2, 8, 20, 7)
    rmono, rk, rd = ref(img, lap)
    assert rmono == mono and len(rk) == len(kp)
    for l in range(8):
        assert np.array_equal(ref.level(l), py.pyramid[l])
        assert np.array_equal(ref.candidates(l), np.array(py.candidates[l], np.float64).reshape(-1, 3).astype(np.int32))
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(rk[f], kp[:, i].astype(np.float32)), f
    assert np.array_equal(rk["octave"], kp[:, 5].astype(np.int32))
    assert np.array_equal(rd, desc)


def test_grey_conversion_formula_is_opencvs():
    """The fixed-point luminance the CUDA level-0 kernel applies to colour input (csrc/k_pyramid.cu) == cv2.cvtColor."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    im = rng.integers(0, 256, (97, 131, 4), dtype=np.uint8)
    f = lambda r, g, b: ((r.astype(np.int64) * 9798 + g.astype(np.int64) * 19235 + b.astype(np.int64) * 3735 + (1 << 14)) >> 15).astype(np.uint8)
    assert np.array_equal(cv2.cvtColor(im[..., :3].copy(), cv2.COLOR_RGB2GRAY), f(im[..., 0], im[..., 1], im[..., 2]))
    assert np.array_equal(cv2.cvtColor(im[..., :3].copy(), cv2.COLOR_BGR2GRAY), f(im[..., 2], im[..., 1], im[..., 0]))
    assert np.array_equal(cv2.cvtColor(im, cv2.COLOR_RGBA2GRAY), f(im[..., 0], im[..., 1], im[..., 2]))
    assert np.array_equal(cv2.cvtColor(im, cv2.COLOR_BGRA2GRAY), f(im[..., 2], im[..., 1], im[..., 0]))


# EuRoC cam0 (Examples/Monocular/EuRoC.yaml:9-17) and a 5-coefficient set
UND_K = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
UND_CASES = [np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32),
             np.array([-0.2917, 0.08228, 5.333e-05, -1.578e-04, 0.01], np.float32),
             np.array([0.0, 0.1, 0.001, 0.001], np.float32)]


def und_points(n=3000):
    rng = np.random.default_rng(9)
    k = np.zeros(n, orbfe.KEYPOINT_DTYPE)
    k["x"], k["y"] = rng.uniform(0, 752, n).astype(np.float32), rng.uniform(0, 480, n).astype(np.float32)
    k["octave"], k["angle"], k["size"] = rng.integers(0, 8, n), rng.uniform(0, 360, n), 31
    return k


@pytest.mark.parametrize("ci", range(len(UND_CASES)))
def test_undistort_keypoints_equals_cv2(ci):
    """Frame::UndistortKeyPoints = cv::undistortPoints(pts, K, D, I, K): the oracle's double-precision restatement
    is bit-identical to the library on the float32 outputs."""
    D = UND_CASES[ci]
    k = und_points()
    out = ob.undistort_keypoints(k, UND_K, D)
    if D[0] == 0:
        assert out.tobytes() == k.tobytes()        # Frame.cc:732-736
        return
    Kmat = np.array([[UND_K[0], 0, UND_K[2]], [0, UND_K[1], UND_K[3]], [0, 0, 1]], np.float32)
    pts = np.stack([k["x"], k["y"]], 1).reshape(-1, 1, 2).copy()
    want = cv2.undistortPoints(pts, Kmat, D, None, Kmat).reshape(-1, 2)
    assert np.array_equal(out["x"], want[:, 0]) and np.array_equal(out["y"], want[:, 1])
    assert np.array_equal(out["octave"], k["octave"]) and np.array_equal(out["angle"], k["angle"])
    assert np.abs(out["x"] - k["x"]).max() > 1.0


def test_bench_opencv_crosscheck_runs_the_same_calls_both_ways():
    """bench.py's cpu_baseline.opencv_crosscheck: the primitive calls of one reference frame through cv2 and
    through the scalar restatement.  The two FAST cell loops must find the same number of candidates (they run the same calls)."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("orbfe_bench_for_test", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    frames = orbfe.synth.frames(1000, 3, bench.W, bench.H)
    threads = cv2.getNumThreads()
    r = bench.cpu_opencv_crosscheck(frames, bench.ORB_PARAMS, budget_s=0.5)
    assert cv2.getNumThreads() == threads                      # restored
    assert r["calls_per_frame"] == {"resize": 7, "copyMakeBorder": 8, "FAST": 700, "GaussianBlur": 8}
    assert r["fast_candidates_first_frame"]["cv2"] == r["fast_candidates_first_frame"]["scalar"] > 2000
    assert r["cv2_ms_per_frame"] > 0 and r["scalar_restatement_ms_per_frame"] > 0
