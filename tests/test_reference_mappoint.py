"""The reference's own src/MapPoint.cc (compiled unmodified into oracle/_ref/liborbslam3_mappoint.so, recipe in
oracle/Makefile) against the C++ oracle, the golden fixtures and the host-side restatements the product uses:
MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:349-423) and MapPoint::PredictScale (:534-566)."""
import os

import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
from common import GOLDEN
from orbfe import projection as pj

pytestmark = pytest.mark.skipif(not ob.ref_mappoint_available(), reason="oracle/_ref/liborbslam3_mappoint.so not built")


def gold(name):
    return np.load(os.path.join(GOLDEN, f"match_{name}.npz"))


@pytest.mark.parametrize("seed,npoints", [(5200, 400), (5201, 3000), (5202, 1500)])
@pytest.mark.parametrize("stereo", [False, True], ids=["mono", "fisheye_pairs"])
def test_distinctive_descriptors_oracle_equals_reference_source(seed, npoints, stereo):
    d, off = mc.distinctive_inputs(seed, npoints)
    ref = ob.ref_distinctive_descriptors(d, off, stereo)
    assert np.array_equal(ref, ob.distinctive_descriptors(d, off))
    if seed == 5200:
        assert np.array_equal(ref, gold("distinctive")["best"])
    n = np.diff(off)
    assert (ref[n == 0] == -1).all() and (ref[n > 0] >= 0).all() and (ref < np.maximum(n, 1)).all()


def test_observation_counts():
    L = ob.ref_mappoint()
    for n in (1, 2, 7, 64):
        assert L.orbslam3_mappoint_nobs(n, 0) == n      # monocular keyframes: one per observation (MapPoint.cc:182-185)
        assert L.orbslam3_mappoint_nobs(n, 1) == n      # second camera present: left and right count one each


def predict_scale_inputs(seed, n, sf=1.2, nlev=8):
    rng = np.random.default_rng(seed)
    md = rng.uniform(0.2, 40.0, n).astype(np.float32)
    cd = (md / np.float32(sf) ** rng.uniform(-2.5, nlev + 1.5, n)).astype(np.float32)
    k = n // 4                                                     # ratios on (and one ulp around) exact powers of the scale factor
    lev = rng.integers(0, nlev, k)
    chain = np.ones(nlev, np.float32)
    for i in range(1, nlev):
        chain[i] = chain[i - 1] * np.float32(sf)
    cd[:k] = md[:k] / chain[lev]
    cd[k // 3:2 * k // 3] = np.nextafter(cd[k // 3:2 * k // 3], np.float32(0))
    cd[2 * k // 3:k] = np.nextafter(cd[2 * k // 3:k], np.float32(1e9))
    return md, cd


@pytest.mark.parametrize("sf,nlev", [(1.2, 8), (1.1, 12), (1.5, 5), (2.0, 4)])
@pytest.mark.parametrize("via_frame", [False, True], ids=["keyframe", "frame"])
def test_predict_scale_equals_reference_source(sf, nlev, via_frame):
    md, cd = predict_scale_inputs(77 + nlev, 20000, sf, nlev)
    logsf = np.float32(np.log(np.float32(sf)))          # ORBextractor / Frame: mfLogScaleFactor = log(mfScaleFactor) (Frame.cc:109)
    ref = ob.ref_predict_scale(md, cd, logsf, nlev, via_frame)
    got = np.array([pj.predict_scale(a, b, logsf, nlev) for a, b in zip(md, cd)], np.int32)
    assert np.array_equal(got, ref), np.flatnonzero(got != ref)[:10]
    assert ref.min() == 0 and ref.max() == nlev - 1
