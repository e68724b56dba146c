"""GPU parity of the projection-guided window searches (through the C ABI) against the C++ oracle engine on the same
flattened windows (assign / choice / count bit-exact) and against the golden results of the object-level Python
restatement of ORBmatcher::SearchByProjection."""
import os

import numpy as np
import pytest

import oracle_binding as ob
import orbfe
import proj_cases as pc
from common import GOLDEN
from orbfe import projection as pj
from test_proj_oracle import owners, run_lf, run_mp

pytestmark = pytest.mark.gpu


def gpu_search(ratio, check_ori=True):
    m = orbfe.ORBmatcher(ratio, check_ori)

    def f(F, q, mode, th, nnratio, link, rot):
        assert abs(nnratio - ratio) < 1e-9 or mode == 0
        return m.search_by_projection(F, q, mode, th, link, rot)
    return f


@pytest.mark.parametrize("name", list(pc.MP_CASES))
def test_search_local_points(name):
    c = pc.mp_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    res, n, q, choice = run_mp(c, gpu_search(c["nnratio"]))
    rres, rn, _, rchoice = run_mp(c, ob.search_by_projection)
    assert n == rn and np.array_equal(choice, rchoice) and np.array_equal(res, rres)
    assert n == int(g["n"]) and np.array_equal(res, g["mappoint_of_feature"])


@pytest.mark.parametrize("name", list(pc.LF_CASES))
def test_track_last_frame(name):
    c = pc.lf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    res, n, q, choice = run_lf(c, gpu_search(0.9, c["ori"]))
    rres, rn, _, rchoice = run_lf(c, ob.search_by_projection)
    assert n == rn and np.array_equal(choice, rchoice) and np.array_equal(res, rres)
    assert n == int(g["n"]) and np.array_equal(res, g["lastframe_feature_of_feature"])


def test_long_claim_chain_converges():
    """300 identical MapPoints over a tight cluster of 40 near-identical features: window i can only settle after
    windows < i have (each takes the best feature still free), the worst case for the parallel sweeps."""
    rng = np.random.default_rng(5)
    n = 40
    k = np.zeros(n, orbfe.KEYPOINT_DTYPE)
    k["x"], k["y"] = pc.snap(300 + rng.uniform(-3, 3, n)), pc.snap(200 + rng.uniform(-3, 3, n))
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    d = np.stack([pc.noisy_desc(rng, base, 30) for _ in range(n)])
    F = pj.GridFrame(k, d, (0, 0, pc.W, pc.H), pc.SCALE)
    q = np.zeros(300, pj.PROJ_QUERY_DTYPE)
    q["u"], q["v"], q["r"], q["min_level"], q["max_level"] = 300, 200, 8, -1, -1
    q["obs"], q["valid"], q["desc"] = 1, 1, base
    m = orbfe.ORBmatcher(0.9, False)
    for mode in (0, 1):
        assign, choice, cnt = m.search_by_projection(F, q, mode, 100)
        rassign, rchoice, rcnt = ob.search_by_projection(F, q, mode, 100, 0.9, False, False)
        assert cnt == rcnt and np.array_equal(choice, rchoice) and np.array_equal(assign, rassign)
        if mode == 0:       # every feature is taken exactly once, by the first 40 windows
            assert cnt == n and sorted(choice[:n].tolist()) == list(range(n)) and (choice[n:] == -1).all()


def test_empty_inputs():
    m = orbfe.ORBmatcher(0.9, True)
    c = pc.mp_inputs("proj_local_mono")
    q, _ = pj.queries_from_mappoints(c["F"], c["mp"], 1.0)
    assign, choice, n = m.search_by_projection(c["F"], q[:0], 1)
    assert n == 0 and (assign == -1).all() and len(choice) == 0
    q2 = q.copy()
    q2["valid"] = 0
    assign, choice, n = m.search_by_projection(c["F"], q2, 1)
    assert n == 0 and (choice == -1).all()


def test_relocalisation_overload():
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist): same engine, mode 0 with th = ORBdist,
    every assignment blocks; GPU == oracle engine on the windows built by projection.queries_from_keyframe."""
    c = pc.lf_inputs("proj_last_mono")
    rng = np.random.default_rng(77)
    last, m = c["last"], len(c["last"]["has_mp"])
    dist = np.sqrt((last["world_pos"].astype(np.float64) ** 2).sum(1)).astype(np.float32)
    kf = dict(has_mp=last["has_mp"], bad=rng.random(m) < 0.03, already_found=rng.random(m) < 0.2, world_pos=last["world_pos"],
              min_dist=(dist * rng.uniform(0.3, 1.05, m)).astype(np.float32), max_dist=(dist * rng.uniform(0.95, 4.0, m)).astype(np.float32),
              angle=last["angle"], desc=last["desc"])
    q, owner = pj.queries_from_keyframe(c["Cur"], kf, c["Tcw"], 10.0, c["project"], np.float32(np.log(1.2)))
    assert len(q) > 400 and len(set(q["min_level"].tolist())) > 3
    for orbdist in (100, 64):
        assign, choice, n = orbfe.ORBmatcher(0.9, True).search_by_projection(c["Cur"], q, 0, orbdist, False, True)
        rassign, rchoice, rn = ob.search_by_projection(c["Cur"], q, 0, orbdist, 0.9, False, True)
        assert n == rn and np.array_equal(assign, rassign) and np.array_equal(choice, rchoice) and n > 60


@pytest.mark.parametrize("name", list(pc.KF_CASES))
def test_keyframe_side_searches(name):
    """SearchByProjection(KeyFrame*, Scw, ...) :473, Fuse :1629 (chi-square gate, bRight) and :1823: GPU == oracle engine
    on the same windows == object-level golden."""
    from test_proj_oracle import run_kf
    c = pc.kf_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    res, n, q, choice = run_kf(c, gpu_search(0.9, False))
    rres, rn, _, rchoice = run_kf(c, ob.search_by_projection)
    assert n == rn and np.array_equal(choice, rchoice) and np.array_equal(res, rres)
    key = "point_of_feature" if c["kind"] == "sim3" else "feature_of_point"
    assert n == int(g["n"]) and np.array_equal(res, g[key])


@pytest.mark.parametrize("name", list(pc.SIM3_CASES))
def test_search_by_sim3(name):
    from test_proj_oracle import run_sim3
    c = pc.sim3_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    vn1, vn2, pairs, n = run_sim3(c, gpu_search(0.9, False))
    assert np.array_equal(vn1, g["vnMatch1"]) and np.array_equal(vn2, g["vnMatch2"]) and n == int(g["n"])


def test_chi2_needs_sigma_table():
    c = pc.kf_inputs("proj_kf_fuse")
    from test_proj_oracle import kf_windows
    q, owner, th = kf_windows(c)
    c["KF"].inv_level_sigma2 = None
    with pytest.raises(orbfe.OrbfeError):
        orbfe.ORBmatcher(0.9, False).search_by_projection(c["KF"], q, 0, th)


@pytest.mark.parametrize("name", list(pc.INIT_CASES))
def test_search_for_initialization(name):
    """ORBmatcher::SearchForInitialization :706: GPU == C++ oracle == object-level golden (vnMatches12, vbPrevMatched, count)."""
    c = pc.init_inputs(name)
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    m = orbfe.ORBmatcher(c["nnratio"], c["ori"])
    n, m12, prev = m.SearchForInitialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"])
    rn, rm12, rprev = ob.search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], c["win"], c["nnratio"], c["ori"])
    assert n == rn and np.array_equal(m12, rm12) and np.array_equal(prev, rprev)
    assert n == int(g["n"]) and np.array_equal(m12, g["vnMatches12"]) and np.array_equal(prev, g["prev_matched"])


@pytest.mark.parametrize("seed,n1,n2,win,ratio,ori", [(4510, 3000, 900, 100, 0.9, True), (4511, 5000, 5000, 100, 0.9, True),
                                                       (4512, 800, 3000, 15, 0.7, False), (4513, 2500, 400, 200, 1.0, True)])
def test_search_for_initialization_vs_oracle(seed, n1, n2, win, ratio, ori):
    """More shapes incl. the initialiser's 5000-feature frames and heavy competition (many F1 features per F2 source)."""
    c = pc.init_inputs(spec=(seed, n1, n2, win, ratio, ori))
    m = orbfe.ORBmatcher(ratio, ori)
    n, m12, prev = m.SearchForInitialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], win)
    rn, rm12, rprev = ob.search_for_initialization(c["F1"].kps, c["F1"].desc, c["F2"], c["prev"], win, ratio, ori)
    assert n == rn and np.array_equal(m12, rm12) and np.array_equal(prev, rprev) and n > 20


def test_search_for_initialization_steal_chain():
    """400 F1 features all derived from 3 F2 features with steadily decreasing distance: every window steals from the
    one before it, so the sweeps need one round per window (worst case), and only the last claimants survive."""
    rng = np.random.default_rng(9)
    k2 = np.zeros(3, orbfe.KEYPOINT_DTYPE)
    k2["x"], k2["y"] = (300, 304, 296), (200, 203, 198)
    d2 = rng.integers(0, 256, (3, 32), dtype=np.uint8)
    F2 = pj.GridFrame(k2, d2, (0, 0, pc.W, pc.H), pc.SCALE)
    n1 = 400
    k1 = np.zeros(n1, orbfe.KEYPOINT_DTYPE)
    k1["x"], k1["y"] = 300, 200
    d1 = np.zeros((n1, 32), np.uint8)
    for i in range(n1):
        d = d2[i % 3].copy()
        for b in rng.choice(256, max(0, 45 - i // 9), replace=False):
            d[b >> 3] ^= 1 << (b & 7)
        d1[i] = d
    prev = np.tile(np.float32([300, 200]), (n1, 1))
    m = orbfe.ORBmatcher(0.9, False)
    n, m12, _ = m.SearchForInitialization(k1, d1, F2, prev, 20)
    rn, rm12, _ = ob.search_for_initialization(k1, d1, F2, prev, 20, 0.9, False)
    assert n == rn == 3 and np.array_equal(m12, rm12)


def test_search_for_initialization_edge_cases():
    c = pc.init_inputs("proj_init_w30_noori")
    m = orbfe.ORBmatcher(0.8, False)
    n, m12, prev = m.SearchForInitialization(c["F1"].kps[:0], c["F1"].desc[:0], c["F2"], c["prev"][:0], 30)
    assert n == 0 and len(m12) == 0
    k = c["F1"].kps.copy()
    k["octave"] = 1                                   # nothing on level 0: every F1 feature is skipped (:724)
    n, m12, prev = m.SearchForInitialization(k, c["F1"].desc, c["F2"], c["prev"], 30)
    assert n == 0 and (m12 == -1).all() and np.array_equal(prev, c["prev"])
    k2 = c["F2"].kps.copy()
    k2["octave"] = 2                                  # no level-0 candidates in F2
    F2 = pj.GridFrame(k2, c["F2"].desc, (0, 0, pc.W, pc.H), pc.SCALE)
    n, m12, _ = m.SearchForInitialization(c["F1"].kps, c["F1"].desc, F2, c["prev"], 30)
    assert n == 0 and (m12 == -1).all()
