"""CPU: the OpenCV-free C++ oracle (oracle/cpu_ref) against the committed golden vectors, which were
produced by the cv2-backed restatement (oracle/py/make_golden.py, real OpenCV 4.13 primitives).
This is the pin that lets the GPU box (no cv2 needed, no /root/reference) trust the C++ oracle."""
import numpy as np
import pytest

from common import check_against_golden, golden_cases, golden_input
from oracle_binding import OrbRef

CASES = golden_cases("extract_")


def test_fixtures_present():
    assert len(CASES) >= 7


@pytest.mark.parametrize("path", CASES, ids=[p.split("extract_")[1][:-4] for p in CASES])
def test_cpu_oracle_matches_golden(path):
    g = np.load(path)
    img = golden_input(g)
    nf, lap = int(g["nfeatures"]), tuple(int(v) for v in g["lapping"])
    sf, nl, ini, mn = g["orb"]
    ref = OrbRef(nf, float(sf), int(nl), int(ini), int(mn))
    mono, kps, desc = ref(img, lap)
    ok, tot = check_against_golden(g, mono, kps, desc, pyramid_level=ref.level, candidates=ref.candidates)
    assert ok == tot, "the C++ oracle and the cv2 restatement share libm and rounding: descriptors must be identical"


def test_tables_match_reference_ctor():
    """ORBextractor.cc:413-444 for the EuRoC settings; quotas from SURVEY.md §8."""
    ref = OrbRef(1000, 1.2, 8, 20, 7)
    sc, isc, s2, is2, per, umax = ref.tables()
    assert per.tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert umax.tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert sc[0] == 1.0 and abs(sc[7] - 1.2 ** 7) < 1e-5
    assert np.array_equal(s2, (sc * sc).astype(np.float32))
    assert np.array_equal(isc, (np.float32(1) / sc).astype(np.float32))
    ref = OrbRef(1200, 1.2, 8, 20, 7)
    assert ref.tables()[4].tolist() == [261, 217, 181, 151, 126, 105, 87, 72]
