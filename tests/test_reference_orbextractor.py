"""CPU: the reference's OWN src/ORBextractor.cc, compiled unmodified into oracle/_ref (oracle/Makefile `ref`: OpenCV
stand-in whose primitives are oracle/cpu_ref/cvprim.cc, monotonic heap so that DistributeOctTree's pointer tie-break
is "later node first"), against
  * the OpenCV-free C++ oracle (oracle/cpu_ref/orb_ref.cc): every byte of the keypoints, descriptors and pyramid, and
  * the golden vectors of the cv2-backed Python restatement (real OpenCV 4.13 primitives, restated control flow).
Together: reference control flow x restated primitives == restated control flow x real primitives, so both halves of the
oracle are pinned.  The GPU parity tests then hold the CUDA path to the same oracle (and tests/test_extract_gpu.py
to this library directly)."""
import numpy as np
import pytest

import oracle_binding as ob
import orbfe
from common import check_against_golden, golden_cases, golden_input, low_texture

pytestmark = pytest.mark.skipif(not ob.ref_available(), reason="oracle/_ref not built (needs /root/reference; make -C oracle ref)")

SHAPES = [("c1", 752, 480, 1000, (0, 1000)), ("c2", 752, 480, 1200, (0, 0)), ("c3", 512, 512, 1500, (128, 511)),
          ("c4", 1241, 376, 2000, (0, 0)), ("c5", 1920, 1080, 4000, (0, 1000))]
ORB = (1.2, 8, 20, 7)


def both(nf, orb=ORB):
    return ob.OrbSlam3Extractor(nf, *orb), ob.OrbRef(nf, *orb)


def same_result(a, b):
    (m1, k1, d1), (m2, k2, d2) = a, b
    assert m1 == m2 and len(k1) == len(k2)
    assert k1.tobytes() == k2.tobytes(), "keypoints (x, y, size, angle, response, octave, class_id) differ"
    assert np.array_equal(d1, d2), "descriptors differ"


@pytest.mark.parametrize("name,w,h,nf,lap", SHAPES, ids=[s[0] for s in SHAPES])
def test_oracle_equals_reference_source(name, w, h, nf, lap):
    ref, orc = both(nf)
    seeds = range(1000, 1003) if w * h > 10 ** 6 else range(1000, 1008)
    for seed in seeds:
        img = orbfe.synth.frame(seed, w, h)
        same_result(ref(img, lap), orc(img, lap))
        if seed == 1000:
            for l in range(8):
                assert np.array_equal(ref.level_bordered(l), orc.level(l)), f"pyramid level {l}"


def test_tables_equal():
    for nf, sf, nl in ((1000, 1.2, 8), (1200, 1.2, 8), (2000, 1.5, 5), (500, 1.1, 12)):
        ref, orc = ob.OrbSlam3Extractor(nf, sf, nl, 20, 7), ob.OrbRef(nf, sf, nl, 20, 7)
        for a, b in zip(ref.tables(), orc.tables()[:4]):
            assert np.array_equal(a, b)


def test_retry_threshold_blank_and_lapping_variants():
    ref, orc = both(1000)
    img = orbfe.synth.frame(1003, 752, 480)
    same_result(ref(low_texture(img), (0, 1000)), orc(low_texture(img), (0, 1000)))      # minThFAST retry (:825-828)
    for lap in ((0, 0), (100, 400), (376, 751), (0, 1000)):
        same_result(ref(img, lap), orc(img, lap))
    blank = np.full((480, 752), 77, np.uint8)
    m, k, d = ref(blank, (0, 0))
    assert m == 0 and len(k) == 0
    same_result((m, k, d), orc(blank, (0, 0)))
    # other extractor settings: thresholds, scale factor, level count
    for nf, orb in ((700, (1.2, 8, 40, 12)), (1500, (1.5, 5, 20, 7)), (900, (1.1, 10, 15, 5))):
        r2, o2 = both(nf, orb)
        same_result(r2(img, (0, 0)), o2(img, (0, 0)))


@pytest.mark.parametrize("path", golden_cases("extract_"), ids=[p.split("extract_")[1][:-4] for p in golden_cases("extract_")])
def test_reference_source_matches_cv2_golden(path):
    g = np.load(path)
    img = golden_input(g)
    nf, lap = int(g["nfeatures"]), tuple(int(v) for v in g["lapping"])
    sf, nl, ini, mn = g["orb"]
    ref = ob.OrbSlam3Extractor(nf, float(sf), int(nl), int(ini), int(mn))
    mono, kps, desc = ref(img, lap)
    ok, tot = check_against_golden(g, mono, kps, desc, pyramid_level=ref.level_bordered)
    assert ok == tot


def test_timing_build_agrees_in_count():
    """The process-allocator build (bench.py's CPU baseline) runs the same code; only pointer ties may order differently,
    so keypoint totals stay within a fraction of a percent."""
    frames = orbfe.synth.frames(1000, 4, 752, 480)
    total = ob.ref_extract_many((1000,) + ORB, frames, (0, 1000), 2)
    ref = ob.OrbSlam3Extractor(1000, *ORB)
    want = sum(len(ref(f, (0, 1000))[1]) for f in frames)
    assert abs(total - want) <= 0.005 * want


def random_configs(seed, count):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(count):
        nl = int(rng.integers(3, 9))
        sf = float(rng.choice([1.15, 1.2, 1.25, 1.3, 1.41]))
        lo = int(np.ceil(120 * sf ** (nl - 1)))                      # the smallest level must still hold a FAST cell row
        w, h = int(rng.integers(max(lo, 200), max(lo, 200) + 700)), int(rng.integers(max(lo, 160), max(lo, 160) + 500))
        if w < h:                 # portrait frames narrower than ~1:2 make DistributeOctTree's nIni = round(width / height) zero
            w, h = h, w           # (a division by zero in the reference, ORBextractor.cc:541-543; orbfe_create refuses them)
        nf = int(rng.integers(200, 2500))
        ini = int(rng.integers(10, 41))
        out.append((w, h, nf, sf, nl, ini, int(rng.integers(3, ini))))
    return out


@pytest.mark.parametrize("cfg", random_configs(20261105, 12), ids=lambda c: f"{c[0]}x{c[1]}_n{c[2]}_s{c[3]}_l{c[4]}_t{c[5]}-{c[6]}")
def test_oracle_equals_reference_source_random_configs(cfg):
    """Odd image sizes, level counts, scale factors, feature budgets and FAST thresholds drawn at random."""
    w, h, nf, sf, nl, ini, mn = cfg
    ref, orc = ob.OrbSlam3Extractor(nf, sf, nl, ini, mn), ob.OrbRef(nf, sf, nl, ini, mn)
    for seed in (7000 + w, 7001 + h):
        img = orbfe.synth.frame(seed, w, h)
        lap = (0, 0) if seed & 1 else (w // 4, 3 * w // 4)
        same_result(ref(img, lap), orc(img, lap))
