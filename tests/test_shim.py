"""The C++ shim classes (include/orbfe_shim/: ORB_SLAM3::ORBextractor / ORBmatcher with the reference's
signatures) compiled against the OpenCV stand-in and driven like Frame.cc drives the reference.
CPU: it compiles, links against liborbfe.so and reports "no device" instead of computing anything.
GPU: its outputs equal the Python-ABI path (which the parity tests hold to the oracle)."""
import os
import subprocess

import numpy as np
import pytest

import match_cases as mc
import orbfe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")
EXE = os.path.join(BUILD, "shim_test")


def build_shim():
    os.makedirs(BUILD, exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", "shim_test.cc")
    deps = [src] + [os.path.join(ROOT, "include", "orbfe_shim", f) for f in os.listdir(os.path.join(ROOT, "include", "orbfe_shim"))]
    deps.append(os.path.join(ROOT, "include", "orbfe.h"))
    if os.path.exists(EXE) and all(os.path.getmtime(d) <= os.path.getmtime(EXE) for d in deps):
        return EXE
    libdir = os.path.join(ROOT, "orb-slam3-noted_b200", "csrc")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include", "orbfe_shim"),
           "-I" + os.path.join(ROOT, "include"), src, "-o", EXE, "-L" + libdir, "-lorbfe", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return EXE


def write_voc(tmp_path, name="voc_k10_L4"):
    """A synthetic vocabulary in the ORBvoc text format, written by the Python restatement."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import bow_cases as bc
    import bow_py
    c = bc.voc_inputs(name)
    v = bow_py.Vocabulary(c["k"], c["L"], c["weighting"], c["norm"])
    for i in range(1, len(c["parent"])):
        v.add_node(int(c["parent"][i]), bool(c["leaf"][i]), c["desc"][i], c["weight"][i])
    p = tmp_path / "voc.txt"
    v.save_text(str(p))
    return p, c


def run_shim(tmp_path, seed=1100, w=752, h=480, nf=1200, bf=47.9064, b=0.110078, voc=None):
    exe = build_shim()
    l, r = orbfe.synth.stereo_pair(seed, w, h)
    l.tofile(tmp_path / "l.raw")
    r.tofile(tmp_path / "r.raw")
    out = tmp_path / "o.bin"
    p = subprocess.run([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(nf), repr(bf), repr(b), str(out)] +
                       ([str(voc)] if voc else []),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return p, out, (l, r)


def test_shim_compiles_and_refuses_without_gpu(tmp_path):
    import torch
    p, out, _ = run_shim(tmp_path, voc=write_voc(tmp_path)[0])
    if torch.cuda.is_available():
        assert p.returncode == 0, p.stdout
    else:
        assert p.returncode == 3 and "no CUDA device" in p.stdout, p.stdout


@pytest.mark.gpu
def test_shim_equals_python_abi_path(tmp_path):
    w, h, nf = 752, 480, 1200
    bf = float(np.float32(47.9064))
    b = float(np.float32(np.float32(47.9064) / np.float32(435.2047)))
    vocp, vc = write_voc(tmp_path)
    p, out, (l, r) = run_shim(tmp_path, 1100, w, h, nf, bf, b, voc=vocp)
    assert p.returncode == 0, p.stdout
    buf = open(out, "rb").read()
    nL, monoL, nR = np.frombuffer(buf, np.int32, 3)
    o = 12
    kps = np.frombuffer(buf, orbfe.KEYPOINT_DTYPE, nL, o); o += 28 * nL
    desc = np.frombuffer(buf, np.uint8, 32 * nL, o).reshape(nL, 32); o += 32 * nL
    u = np.frombuffer(buf, np.float32, nL, o); o += 4 * nL
    d = np.frombuffer(buf, np.float32, nL, o); o += 4 * nL
    nb = int(np.frombuffer(buf, np.int32, 1, o)[0]); o += 4
    bow = np.frombuffer(buf, np.int32, nR, o); o += 4 * nR
    sf = np.frombuffer(buf, np.float32, 8, o); o += 32
    pw, ph = np.frombuffer(buf, np.int32, 2, o); o += 8
    lvl1 = np.frombuffer(buf, np.uint8, pw * ph, o).reshape(ph, pw); o += pw * ph
    und = np.frombuffer(buf, np.float32, 2 * nL, o).reshape(nL, 2); o += 8 * nL
    nproj = int(np.frombuffer(buf, np.int32, 1, o)[0]); o += 4
    passign = np.frombuffer(buf, np.int32, nR, o); o += 4 * nR
    pbest = np.frombuffer(buf, np.int32, 3, o); o += 12
    ninit = int(np.frombuffer(buf, np.int32, 1, o)[0]); o += 4
    init12 = np.frombuffer(buf, np.int32, nL, o); o += 4 * nL
    initprev = np.frombuffer(buf, np.float32, 2 * nL, o).reshape(nL, 2); o += 8 * nL
    nw, nn = np.frombuffer(buf, np.int32, 2, o); o += 8
    bv = np.frombuffer(buf, np.dtype([("w", "<u4"), ("v", "<f8")]), nw, o); o += 12 * nw
    fv = {}
    for _ in range(nn):
        nid, cnt = np.frombuffer(buf, np.int32, 2, o); o += 8
        fv[int(nid)] = np.frombuffer(buf, np.uint32, cnt, o).tolist(); o += 4 * cnt
    assert o == len(buf)

    exL = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h)
    exR = orbfe.ORBextractor(nf, *mc.ORB, max_width=w, max_height=h)
    mono, k, dd = exL(l, None, (0, 0))
    _, kR, dR = exR(r, None, (0, 0))
    assert (nL, monoL, nR) == (len(k), mono, len(kR))
    assert kps.tobytes() == k.tobytes() and np.array_equal(desc, dd)
    m = orbfe.ORBmatcher(0.7, True)
    pu, pd, n = m.stereo_match(exL, exR, bf, b)
    assert np.array_equal(u, pu[:nL]) and np.array_equal(d, pd[:nL])
    b2a, n = m.SearchByBoW_KF_F(dd, mc.fv_of(dd), np.ones(nL, np.uint8), k, dR, mc.fv_of(dR), kR)
    assert n == nb and np.array_equal(bow, b2a)
    assert np.array_equal(sf, exL.GetScaleFactors())
    assert np.array_equal(lvl1, exL.image_pyramid(1)[1])
    # ORBVocabulary::transform through the shim == the oracle on the same descriptors
    from oracle_binding import VocRef
    rw, rv, rn, ro, ri = VocRef(vc["k"], vc["L"], vc["weighting"], vc["norm"], vc["parent"], vc["leaf"], vc["desc"],
                                vc["weight"]).transform(dd, 4)
    assert np.array_equal(bv["w"], rw) and np.array_equal(bv["v"], rv)
    assert sorted(fv) == rn.tolist() and all(fv[int(n)] == ri[ro[i]:ro[i + 1]].tolist() for i, n in enumerate(rn))
    # undistortion, projection windows and distinctive descriptors through the shim == the oracle
    import oracle_binding as ob
    from orbfe import projection as pj
    ku = ob.undistort_keypoints(k, [458.654, 457.296, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05])
    assert np.array_equal(und[:, 0], ku["x"]) and np.array_equal(und[:, 1], ku["y"])
    G = pj.GridFrame(kR, dR, (0, 0, w, h), exL.GetScaleFactors())
    win = np.zeros(nL, pj.PROJ_QUERY_DTYPE)
    win["u"], win["v"], win["r"] = k["x"] - np.float32(12), k["y"], 30
    win["min_level"], win["max_level"], win["angle"], win["obs"], win["valid"], win["desc"] = k["octave"] - 1, k["octave"] + 1, k["angle"], 1, 1, dd
    rassign, _, rn = ob.search_by_projection(G, win, 0, 100, 0.7, False, True)
    assert nproj == rn and np.array_equal(passign, rassign) and rn > 200
    offs = np.array([0, min(5, nL), min(5, nL), min(40, nL)], np.int32)
    assert np.array_equal(pbest, ob.distinctive_descriptors(dd, offs))
    rni, r12, rprev = ob.search_for_initialization(k, dd, G, np.stack([k["x"], k["y"]], 1), 40, 0.7, True)
    assert ninit == rni and np.array_equal(init12, r12) and np.array_equal(initprev, rprev) and rni > 30
