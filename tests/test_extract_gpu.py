"""GPU parity: the CUDA extractor (through the C ABI) against the CPU oracle, stage by stage.

Bars (BASELINE.json north_star / SURVEY.md Appendix B): pyramid bytes, FAST candidates, selected
keypoint set + order + response/octave/size/pt: bit-exact; angles within 1e-4 rad; >= 99.99 % of
descriptors bit-identical.
"""
import numpy as np
import pytest

import orbfe
from common import check_against_golden, golden_cases, golden_input
from oracle_binding import OrbRef

pytestmark = pytest.mark.gpu

ANGLE_TOL_DEG = 1e-4 * 180.0 / np.pi   # 1e-4 rad

# (name, w, h, nfeatures, lapping) — the five BASELINE.json shapes
SHAPES = [
    ("euroc_mono", 752, 480, 1000, (0, 1000)),
    ("euroc_stereo", 752, 480, 1200, (0, 0)),
    ("tumvi_fisheye", 512, 512, 1500, (128, 511)),
    ("kitti", 1241, 376, 2000, (0, 0)),
    ("hd", 1920, 1080, 4000, (0, 1000)),
]


def check_frame(ex, ref, img, lap, frame=0, got=None, stages=True):
    rmono, rkps, rdesc = ref(img, lap)
    if got is None:
        got = ex(img, None, lap)
    mono, kps, desc = got
    if stages:
        for l in range(ex.nlevels):
            buf, _ = ex.image_pyramid(l, frame)
            assert np.array_equal(buf, ref.level(l)), f"pyramid level {l}"
        for l in range(ex.nlevels):
            assert np.array_equal(ex.candidates(l, frame), ref.candidates(l)), f"FAST candidates level {l}"
    assert len(kps) == len(rkps)
    assert mono == rmono
    for f in ("octave", "x", "y", "response", "size", "class_id"):
        assert np.array_equal(kps[f], rkps[f]), f
    assert np.max(np.abs(kps["angle"] - rkps["angle"]), initial=0) <= ANGLE_TOL_DEG
    same = np.all(desc == rdesc, axis=1)
    return int(same.sum()), len(same)


@pytest.mark.parametrize("name,w,h,nf,lap", SHAPES, ids=[s[0] for s in SHAPES])
def test_extract_parity(name, w, h, nf, lap):
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ref = OrbRef(nf, 1.2, 8, 20, 7)
    ok = tot = 0
    for seed in range(3):
        img = orbfe.synth.frame(1000 + 100 * SHAPES.index((name, w, h, nf, lap)) + seed, w, h)
        a, b = check_frame(ex, ref, img, lap)
        ok += a
        tot += b
    assert tot > 0.9 * 3 * nf
    assert ok >= 0.9999 * tot, (ok, tot)


@pytest.mark.parametrize("name,w,h,nf,lap", SHAPES, ids=[s[0] for s in SHAPES])
def test_cuda_equals_reference_source(name, w, h, nf, lap):
    """The CUDA path against the reference's own src/ORBextractor.cc (oracle/_ref, prebuilt; see
    tests/test_reference_orbextractor.py for what that library is)."""
    import oracle_binding as ob
    if not ob.ref_available():
        pytest.skip("oracle/_ref not built")
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ref = ob.OrbSlam3Extractor(nf, 1.2, 8, 20, 7)
    ok = tot = 0
    for seed in range(2):
        img = orbfe.synth.frame(2000 + 100 * SHAPES.index((name, w, h, nf, lap)) + seed, w, h)
        a, b = check_frame(ex, ref, img, lap, stages=False)
        for l in range(ex.nlevels):
            assert np.array_equal(ex.image_pyramid(l, 0)[0], ref.level_bordered(l)), f"pyramid level {l}"
        ok += a
        tot += b
    assert tot > 0.9 * 2 * nf and ok >= 0.9999 * tot, (ok, tot)


def test_batch_equals_single():
    w, h, nf = 752, 480, 1000
    imgs = orbfe.synth.frames(2000, 6, w, h)
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=6)
    ref = OrbRef(nf, 1.2, 8, 20, 7)
    res = ex.extract_batch(imgs, (0, 1000))
    ok = tot = 0
    for b in range(6):
        a, t = check_frame(ex, ref, imgs[b], (0, 1000), frame=b, got=res[b])
        ok += a
        tot += t
    assert ok >= 0.9999 * tot


def test_low_texture_retry_and_strided_input():
    """Flat regions force the minThFAST retry (ORBextractor.cc:825-828); input with a row stride > width."""
    w, h = 752, 480
    img = orbfe.synth.frame(77, w, h).copy()
    img[:, : w // 2] = (img[:, : w // 2].astype(np.int32) // 8 + 112).astype(np.uint8)   # contrast / 8
    img[: h // 3, : w // 4] = 90                                                        # fully flat block
    big = np.zeros((h, w + 40), np.uint8)
    big[:, :w] = img
    view = big[:, :w]
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ref = OrbRef(1000, 1.2, 8, 20, 7)
    ok, tot = check_frame(ex, ref, img, (0, 0), got=ex(view, None, (0, 0)))
    assert tot > 300 and ok >= 0.9999 * tot


def test_shape_change_and_errors():
    ex = orbfe.ORBextractor(500, 1.2, 8, 20, 7, max_width=800, max_height=600)
    ref = OrbRef(500, 1.2, 8, 20, 7)
    for (w, h) in [(640, 480), (800, 600), (376, 240)]:
        img = orbfe.synth.frame(5, w, h)
        ok, tot = check_frame(ex, ref, img, (0, 0))
        assert ok >= 0.9999 * tot
    mono, kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(kps) == 0
    with pytest.raises(orbfe.OrbfeError):
        ex(np.zeros((700, 900), np.uint8))


def test_blank_image_gives_no_keypoints():
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=752, max_height=480)
    mono, kps, desc = ex(np.full((480, 752), 128, np.uint8), None, (0, 0))
    assert mono == 0 and len(kps) == 0 and desc.shape == (0, 32)


GOLD = golden_cases("extract_")


@pytest.mark.parametrize("path", GOLD, ids=[p.split("extract_")[1][:-4] for p in GOLD])
def test_extract_matches_golden(path):
    """The CUDA path against the committed cv2-generated fixtures (no oracle in the loop)."""
    g = np.load(path)
    img = golden_input(g)
    sf, nl, ini, mn = g["orb"]
    ex = orbfe.ORBextractor(int(g["nfeatures"]), float(sf), int(nl), int(ini), int(mn), max_width=int(g["w"]),
                            max_height=int(g["h"]))
    mono, kps, desc = ex(img, None, tuple(int(v) for v in g["lapping"]))
    ok, tot = check_against_golden(g, mono, kps, desc, pyramid_level=lambda l: ex.image_pyramid(l)[0],
                                   candidates=ex.candidates)
    assert ok >= 0.9999 * tot, (ok, tot)


@pytest.mark.parametrize("fmt", ["rgb", "bgr", "rgba", "bgra"])
def test_colour_input_equals_cvtcolor_then_extract(fmt):
    """Tracking::GrabImageMonocular converts colour frames with cv::cvtColor before extraction (Tracking.cc:1233-1258);
    with orbfe_set_input_format the conversion is fused into the pyramid.  Level 0 must equal OpenCV's grey image
    (fixed point (R*9798 + G*19235 + B*3735 + 16384) >> 15, verified against cv2 4.13 in the oracle tests) and the
    features must equal those of the grey path bit for bit."""
    w, h, nf = 752, 480, 1000
    rng = np.random.default_rng(11)
    grey = orbfe.synth.frame(1000, w, h, 0, 0)
    nch = 3 if fmt in ("rgb", "bgr") else 4
    col = np.zeros((h, w, nch), np.uint8)
    # a colourful image whose luminance keeps the synthetic frame's structure
    col[..., 0] = np.clip(grey.astype(np.int32) + rng.integers(-30, 31, (h, w)), 0, 255)
    col[..., 1] = grey
    col[..., 2] = np.clip(grey.astype(np.int32) + rng.integers(-50, 51, (h, w)), 0, 255)
    if nch == 4:
        col[..., 3] = rng.integers(0, 256, (h, w))
    r_idx, b_idx = (0, 2) if fmt in ("rgb", "rgba") else (2, 0)
    R, G, B = (col[..., i].astype(np.int64) for i in (r_idx, 1, b_idx))
    want_grey = ((R * 9798 + G * 19235 + B * 3735 + (1 << 14)) >> 15).astype(np.uint8)
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    mono0, k0, d0 = ex(want_grey, None, (0, 0))
    ex.set_input_format(fmt)
    mono1, k1, d1 = ex(col, None, (0, 0))
    assert np.array_equal(ex.image_pyramid(0)[1], want_grey)
    assert mono0 == mono1 and k0.tobytes() == k1.tobytes() and np.array_equal(d0, d1) and len(k0) > 900
    ex.set_input_format("gray")
    mono2, k2, d2 = ex(want_grey, None, (0, 0))
    assert k2.tobytes() == k0.tobytes()


def _random_configs():
    from test_reference_orbextractor import random_configs
    return random_configs(20261106, 8)


@pytest.mark.parametrize("cfg", _random_configs(), ids=lambda c: f"{c[0]}x{c[1]}_n{c[2]}_s{c[3]}_l{c[4]}_t{c[5]}-{c[6]}")
def test_cuda_equals_reference_source_random_configs(cfg):
    """Random image sizes, level counts, scale factors, budgets and thresholds: CUDA vs the reference's own ORBextractor.cc."""
    import oracle_binding as ob
    if not ob.ref_available():
        pytest.skip("oracle/_ref not built")
    w, h, nf, sf, nl, ini, mn = cfg
    ex = orbfe.ORBextractor(nf, sf, nl, ini, mn, max_width=w, max_height=h)
    ref = ob.OrbSlam3Extractor(nf, sf, nl, ini, mn)
    img = orbfe.synth.frame(8000 + w + h, w, h)
    lap = (w // 3, 2 * w // 3)
    ok, tot = check_frame(ex, ref, img, lap, stages=False)
    for l in range(nl):
        assert np.array_equal(ex.image_pyramid(l, 0)[0], ref.level_bordered(l)), f"pyramid level {l}"
    assert ok >= 0.9999 * tot and tot > 0.5 * nf, (ok, tot)


@pytest.mark.parametrize("ini,mn", [(150, 130), (200, 127), (128, 5), (127, 126), (254, 253)])
def test_cuda_equals_reference_source_high_thresholds(ini, mn):
    """FAST thresholds on both sides of 128 (the packed byte compare of k_fast_cells switches form there) on a black / white
    checker pattern with noise: CUDA vs the reference's own ORBextractor.cc."""
    import oracle_binding as ob
    if not ob.ref_available():
        pytest.skip("oracle/_ref not built")
    w, h, nf = 640, 480, 800
    rng = np.random.default_rng(ini * 1000 + mn)
    yy, xx = np.mgrid[0:h, 0:w]
    img = (((xx // 23 + yy // 17) % 2) * 235 + 10 + rng.integers(-10, 11, (h, w))).astype(np.uint8)
    img[rng.random((h, w)) < 0.002] = 255                      # isolated bright dots: corners at any threshold below 235
    ex = orbfe.ORBextractor(nf, 1.2, 8, ini, mn, max_width=w, max_height=h)
    ref = ob.OrbSlam3Extractor(nf, 1.2, 8, ini, mn)
    ok, tot = check_frame(ex, ref, img, (0, 0), stages=False)
    assert ok >= 0.9999 * tot
    assert tot > 50 or ini > 235, tot


@pytest.mark.parametrize("params", [(1000, 1.2, 8, 20, 7), (1200, 1.2, 8, 20, 7), (2000, 1.25, 5, 26, 6), (777, 1.5, 4, 18, 9),
                                    (1500, 2.0, 3, 20, 7), (4000, 1.2, 8, 20, 7), (500, 1.1, 12, 20, 7)])
def test_tables_equal_reference_constructor(params):
    """orbfe_get_tables (row a1) straight against the tables the reference's own constructor builds (ORBextractor.cc:413-444,
    read out of oracle/_ref's ORB_SLAM3::ORBextractor object) and the oracle's per-level quotas: float bit patterns."""
    import oracle_binding as ob
    w, h = 752, 480
    ex = orbfe.ORBextractor(*params, max_width=w, max_height=h)
    sc, isc, s2, is2, per, umax = OrbRef(*params).tables()
    got = (ex.GetScaleFactors(), ex.GetInverseScaleFactors(), ex.GetScaleSigmaSquares(), ex.GetInverseScaleSigmaSquares())
    for a, b in zip(got, (sc, isc, s2, is2)):
        assert a.dtype == np.float32 and a.tobytes() == b.tobytes()
    assert np.array_equal(ex.features_per_level(), per) and per.sum() >= min(params[0], per.sum())
    if ob.ref_available():
        for a, b in zip(got, ob.OrbSlam3Extractor(*params).tables()):
            assert a.tobytes() == b.tobytes()


def test_fast_corner_list_overflow_path():
    """k_fast_warp keeps a bounded corner list per cell and walks the score map when a cell has more corners than that.
    ORBFE_FAST_L2CAP (read once per process) shrinks the bound to 8, so almost every cell of an ordinary frame overflows;
    the candidates and key points must not change.  Runs in a child process because of the read-once knob."""
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = (
        "import sys, numpy as np\n"
        f"sys.path[:0] = [{here!r}, {os.path.join(os.path.dirname(here), 'orb-slam3-noted_b200')!r}]\n"
        "import orbfe\n"
        "from oracle_binding import OrbRef\n"
        "ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=752, max_height=480)\n"
        "ref = OrbRef(1000, 1.2, 8, 20, 7)\n"
        "for seed in (4100, 4101):\n"
        "    img = orbfe.synth.frame(seed, 752, 480)\n"
        "    rmono, rk, rd = ref(img, (0, 1000)); mono, k, d = ex(img, None, (0, 1000))\n"
        "    for l in range(8): assert np.array_equal(ex.candidates(l, 0), ref.candidates(l)), l\n"
        "    assert mono == rmono and all(np.array_equal(k[f], rk[f]) for f in ('octave', 'x', 'y', 'response'))\n"
        "    assert np.array_equal(d, rd)\n"
        "print('overflow path ok')\n")
    env = dict(os.environ, ORBFE_FAST_L2CAP="8")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "overflow path ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("env", [{"ORBFE_PYR_TMA": "0"}, {"ORBFE_BLUR_OVERLAP": "1", "ORBFE_BLUR_PERSIST": "2"},
                                 {"ORBFE_BLUR": "tensor"}, {"ORBFE_GRAPHS": "0", "ORBFE_DEV_CHUNK": "2", "ORBFE_DEV_STREAMS": "3"},
                                 {"ORBFE_FAST_SIDE": "0", "ORBFE_TEST_FRAMES": "8"}, {"ORBFE_FAST_SIDE": "1", "ORBFE_TEST_FRAMES": "8"},
                                 {"ORBFE_FAST_SIDE": "2", "ORBFE_TEST_FRAMES": "8"}],
                         ids=["resize_through_L1", "blur_on_side_stream", "tensor_blur", "chunked_no_graphs", "fast_in_line",
                              "fast_tall_cells_on_side_stream", "fast_tall_cells_and_their_levels_on_side_stream"])
def test_selectable_paths_give_the_same_result(env):
    """The kernels and schedules that are not the default (read-once environment knobs) against the oracle: the L1 form of the
    resize, the blur on a side stream with resident CTAs, the tensor-core blur, eager chunked launches on three streams, FAST's
    two launches in line / forked (eight frames: more cells than fit the GPU at once, so that there ARE two launches)."""
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = (
        "import os, sys, numpy as np\n"
        f"sys.path[:0] = [{here!r}, {os.path.join(os.path.dirname(here), 'orb-slam3-noted_b200')!r}]\n"
        "import orbfe\n"
        "from oracle_binding import OrbRef\n"
        "w, h, nfr = 640, 400, int(os.environ.get('ORBFE_TEST_FRAMES', '5'))\n"
        "ex = orbfe.ORBextractor(800, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nfr)\n"
        "ref = OrbRef(800, 1.2, 8, 20, 7)\n"
        "frames = np.stack([orbfe.synth.frame(4200 + i, w, h) for i in range(nfr)])\n"
        "for rep in range(3):\n"                       # eager, captured, replayed
        "    res = ex.extract_batch(frames, (0, 0))\n"
        "for f in range(nfr):\n"
        "    rmono, rk, rd = ref(frames[f], (0, 0)); mono, k, d = res[f]\n"
        "    assert mono == rmono and all(np.array_equal(k[x], rk[x]) for x in ('octave', 'x', 'y', 'response')), f\n"
        "    assert np.array_equal(d, rd), f\n"
        "    if f == nfr - 1:\n"
        "        for l in range(8): assert np.array_equal(ex.image_pyramid(l, f)[0], ref.level(l)), l\n"
        "print('selectable path ok')\n")
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "selectable path ok" in r.stdout, r.stdout + r.stderr


def test_largest_supported_image_and_refusal_beyond():
    """Level-0 ROI up to 2063 px per side (11-bit packed coordinates, INTEGRATION.md "Limits"): the widest frame the
    library takes is extracted exactly; a wider handle or image is refused with ORBFE_ERR_TOO_LARGE (not a crash, not a wrong
    result), and the handle stays usable."""
    w, h, nf = 2063, 600, 3000
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ref = OrbRef(nf, 1.2, 8, 20, 7)
    img = orbfe.synth.frame(6100, w, h)
    ok, tot = check_frame(ex, ref, img, (0, 0), stages=True)
    assert tot > 0.9 * nf and ok >= 0.9999 * tot
    with pytest.raises(orbfe.OrbfeError) as e:
        ex(orbfe.synth.frame(6101, w + 1, h), None, (0, 0))
    assert e.value.status == -3, e.value            # ORBFE_ERR_TOO_LARGE
    with pytest.raises(orbfe.OrbfeError) as e:
        orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w + 1, max_height=h)
    assert e.value.status == -3, e.value
    ok2, tot2 = check_frame(ex, ref, img, (0, 0), stages=False)
    assert (ok2, tot2) == (ok, tot)
