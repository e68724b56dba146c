"""CPU: the flattened window-search engine of the C++ oracle (oracle/cpu_ref/proj_ref.cc) fed by the host-side
flattening (orbfe/projection.py) against the golden results of the OBJECT-level Python restatement of
SearchByProjection (oracle/py/proj_py.py), plus GetFeaturesInArea known answers."""
import os
import zlib

import numpy as np
import pytest

import oracle_binding as ob
import proj_cases as pc
from common import GOLDEN
from orbfe import projection as pj


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c & 0xFFFFFFFF


def owners(assign, owner):
    out = np.full(len(assign), -1, np.int32)
    ok = assign >= 0
    out[ok] = owner[assign[ok]]
    return out


def run_mp(c, search):
    q, owner = pj.queries_from_mappoints(c["F"], c["mp"], c["th"], c["far"], c["th_far"])
    assign, choice, n = search(c["F"], q, 1, 100, c["nnratio"], True, False)
    return owners(assign, owner), n, q, choice


def run_lf(c, search):
    q, owner = pj.queries_from_lastframe(c["Cur"], c["last"], c["Tcw"], c["th"], c["fb"], c["project"], c["project_right"])
    assign, choice, n = search(c["Cur"], q, 0, 100, 0.9, False, c["ori"])
    return owners(assign, owner), n, q, choice


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_local_points_matches_golden(name):
    c = pc.mp_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["F"].kps, c["F"].desc, c["mp"]["desc"], c["mp"]["proj_x"]) == int(g["in_crc"])
    res, n, q, choice = run_mp(c, ob.search_by_projection)
    assert n == int(g["n"]) and np.array_equal(res, g["mappoint_of_feature"])
    assert n > 300 and (q["right"] > 0).any() == (c["F"].nleft != -1)
    # MapPoints with observations hold their feature exclusively; those without (obs == 0) can be overwritten
    held = choice[(choice >= 0) & (q["obs"] > 0)]
    assert len(set(held.tolist())) == len(held)
    assert len(set(choice[choice >= 0].tolist())) < (choice >= 0).sum()


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_track_last_frame_matches_golden(name):
    c = pc.lf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    assert crc(c["Cur"].kps, c["Cur"].desc, c["last"]["desc"], c["last"]["world_pos"]) == int(g["in_crc"])
    res, n, q, choice = run_lf(c, ob.search_by_projection)
    assert n == int(g["n"]) and np.array_equal(res, g["lastframe_feature_of_feature"])
    assert n > 200


def test_features_in_area_known_answers():
    import orbfe
    k = np.zeros(6, orbfe.KEYPOINT_DTYPE)
    k["x"] = [100, 104, 96, 100, 400, 751.9]
    k["y"] = [100, 100, 100, 109.9, 300, 479.9]
    k["octave"] = [0, 1, 2, 0, 0, 0]
    F = pj.GridFrame(k, np.zeros((6, 32), np.uint8), (0, 0, 752, 480), pc.SCALE)
    assert ob.features_in_area(F, 100, 100, 5, -1, -1, False).tolist() == [2, 0, 1]       # cell order: x cell first
    assert ob.features_in_area(F, 100, 100, 4, -1, -1, False).tolist() == [0]             # |dx| < r is strict
    assert ob.features_in_area(F, 100, 100, 10, 0, 0, False).tolist() == [0, 3]
    assert ob.features_in_area(F, 100, 100, 10, 1, -1, False).tolist() == [2, 1]          # minLevel only
    assert ob.features_in_area(F, 100, 100, 10, -1, 1, False).tolist() == [0, 1, 3]       # cell (9,10) before (9,11)
    assert ob.features_in_area(F, 2000, 100, 10, -1, -1, False).tolist() == []
    # the last feature rounds to grid cell (64, 48): PosInGrid drops it (Frame.cc:714)
    assert ob.features_in_area(F, 751, 479, 5, -1, -1, False).tolist() == []
