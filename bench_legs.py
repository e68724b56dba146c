"""bench_legs.py — the legs of bench.py for BASELINE.json's other named shapes (configs 1-4) and the drop-in latency.

Every leg returns one object of the bench line: metric / value (device-resident, CUDA events) / e2e (through the host-buffer
C-ABI calls, copies inside the timed region) / roofline / cpu_baseline (kind "reference": the reference's own sources from
oracle/_ref timed on the host cores; "port" only where no single reference build runs the path).  Legs run on rank 0 of a
multi-GPU job as well as at N = 1 (they are shapes of the same frame-sharded path; the scaling run is the headline's).
"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding
    return oracle_binding


def _cases():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import match_cases
    return match_cases


def _percentiles(samples_s):
    a = np.sort(np.asarray(samples_s)) * 1e3
    return {"p50_ms": float(a[len(a) // 2]), "p99_ms": float(a[min(len(a) - 1, int(0.99 * len(a)))]), "mean_ms": float(a.mean()),
            "min_ms": float(a[0]), "calls": len(a)}


def _level_sizes(ex):
    out = []
    for l in range(ex.nlevels):
        _, roi = ex.image_pyramid(l, 0)
        out.append((roi.shape[1], roi.shape[0]))
    return out


def _b_alg(w, h, n_kp, sizes):
    return w * h + sum((a + 38) * (b + 38) for a, b in sizes) + n_kp * 60


def _pinned(torch, shape, dtype=None):
    return torch.empty(shape, dtype=dtype or torch.uint8).pin_memory()


def _extract_throughput(torch, orbfe, ex, frames, W, H, lap, K, Wm, peak):
    """Device-resident frames/s + stage times + e2e through orbfe_extract_batch for one shape.  frames: (B, H, W) uint8."""
    B = len(frames)
    host = _pinned(torch, (B, H, W))
    host.numpy()[:] = frames
    dev = host.cuda()
    stream = torch.cuda.ExternalStream(ex.stream())
    for _ in range(Wm):
        ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=lap)
    ex.synchronize()
    res = ex.fetch_results(B)
    n_kp = float(np.mean([len(r[1]) for r in res]))
    l0 = ex.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(K):
        ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=lap)
    e1.record(stream)
    ex.synchronize()
    ms = e0.elapsed_time(e1)
    launches = ex.launch_count() - l0
    ex.set_profiling(True)
    ex.stage_times()
    for _ in range(K):
        ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=lap)
    stage_ms, calls = ex.stage_times()
    ex.set_profiling(False)
    # end to end
    kps = _pinned(torch, (B, ex.cap, 28)).numpy().view(orbfe.KEYPOINT_DTYPE).reshape(B, ex.cap)
    desc = _pinned(torch, (B, ex.cap, 32)).numpy()
    n_out, mono = np.zeros(B, np.int32), np.zeros(B, np.int32)
    L = orbfe.lib()
    hnp = host.numpy()

    def e2e_step():
        st = L.orbfe_extract_batch(ex.handle, hnp.ctypes.data, B, W, H, W, W * H, lap[0], lap[1], kps.ctypes.data, desc.ctypes.data,
                                   ex.cap, n_out.ctypes.data, mono.ctypes.data)
        if st < 0:
            raise orbfe.OrbfeError(st, "orbfe_extract_batch")
    for _ in range(Wm):
        e2e_step()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    e2e_s = time.perf_counter() - t0
    sizes = _level_sizes(ex)
    b_alg = _b_alg(W, H, n_kp, sizes)
    calls = max(calls, 1)
    top = max(stage_ms, key=stage_ms.get)
    frames_per_launch = B * K / calls
    top_ms = stage_ms[top] / calls
    achieved = b_alg * frames_per_launch / (top_ms / 1e3) / 1e9
    whole = b_alg * B * K / (ms / 1e3) / 1e9
    return {
        "value": K * B / (ms / 1e3), "ms_per_step": ms / K, "gpu_launches": int(launches), "mean_keypoints_per_frame": n_kp,
        "e2e": {"value": K * B / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(B * W * H), "d2h_bytes_per_step": int(B * ex.cap * 60 + 8 * B),
                "ms_per_step": 1e3 * e2e_s / K, "api": "orbfe_extract_batch (pinned host buffers)"},
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "algorithmic_bytes_per_frame": b_alg, "frames_per_launch": frames_per_launch,
                     "kernel_ms_per_launch": top_ms, "whole_path_achieved": whole, "whole_path_frac": whole / peak,
                     "stage_ms_per_step": {k: v / K for k, v in stage_ms.items()}},
    }, res, dev, host


# ---- config 1: EuRoC-shape stereo pair 752x480, 1200 features per eye + Frame::ComputeStereoMatches ------------------------
def run_stereo(args, torch, orbfe, local, peak):
    mc = _cases()
    cam = mc.EUROC
    W, H, NF = cam["w"], cam["h"], 1200
    P, K, Wm = args.stereo_pairs, max(3, args.steps // 3), args.warmup
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    nd = min(P, 16)
    pairs = [orbfe.synth.stereo_pair(1100 + i, W, H) for i in range(nd)]
    frames = np.stack([pairs[p % nd][eye] for p in range(P) for eye in (0, 1)])
    ex = orbfe.ORBextractor(NF, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=2 * P)
    m = orbfe.ORBmatcher(0.6, True, device=local)
    base, res, dev, host = _extract_throughput(torch, orbfe, ex, frames, W, H, (0, 0), K, Wm, peak)
    cap = ex.cap
    ur = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    dp = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    s_ex, s_m = torch.cuda.ExternalStream(ex.stream()), torch.cuda.ExternalStream(m.stream())

    def step():
        ex.extract_batch_device(dev.data_ptr(), 2 * P, W, H)
        m.stereo_match_batch_device(ex, P, bf, b, ur.data_ptr(), dp.data_ptr())     # waits for the extraction, releases the handle
    for _ in range(Wm):
        step()
    m.synchronize()
    l0 = ex.launch_count() + m.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s_ex)
    for _ in range(K):
        step()
    e1.record(s_m)
    m.synchronize()
    ex.synchronize()
    ms = e0.elapsed_time(e1)
    launches = ex.launch_count() + m.launch_count() - l0
    # the stereo kernels alone (k_stereo_search + k_stereo_filter)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(s_m)
    for _ in range(K):
        m.stereo_match_batch_device(ex, P, bf, b, ur.data_ptr(), dp.data_ptr())
    e3.record(s_m)
    m.synchronize()
    ms_stereo = e2.elapsed_time(e3) / K
    depth = dp.cpu().numpy()
    n_depth = float(np.mean([(depth[p, :len(res[2 * p][1])] > 0).sum() for p in range(P)]))
    # parity spot check outside the timed regions: pair 0 of the batch against the reference's own stereo Frame constructor
    ob = _oracle()
    parity = None
    if ob.ref_frame_available():
        K4 = np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32)
        F = ob.ref_stereo_frame(pairs[0][0], pairs[0][1], (NF,) + mc.ORB, K4, np.zeros(4, np.float32), bf)
        n0 = len(F["kps"])
        parity = bool(F["kps"].tobytes() == res[0][1].tobytes() and np.array_equal(F["desc"], res[0][2]) and
                      np.array_equal(F["u_right"], ur[0, :n0].cpu().numpy()) and np.array_equal(F["depth"], depth[0, :n0]))
    # end to end: host frames in, mvuRight / mvDepth of every pair out
    kps = _pinned(torch, (2 * P, cap, 28)).numpy().view(orbfe.KEYPOINT_DTYPE).reshape(2 * P, cap)
    desc = _pinned(torch, (2 * P, cap, 32)).numpy()
    n_out, mono = np.zeros(2 * P, np.int32), np.zeros(2 * P, np.int32)
    hur, hdp = _pinned(torch, (P, cap), torch.float32), _pinned(torch, (P, cap), torch.float32)
    L, hnp = orbfe.lib(), host.numpy()

    def e2e_step():
        st = L.orbfe_extract_batch(ex.handle, hnp.ctypes.data, 2 * P, W, H, W, W * H, 0, 0, kps.ctypes.data, desc.ctypes.data, cap,
                                   n_out.ctypes.data, mono.ctypes.data)
        if st < 0:
            raise orbfe.OrbfeError(st, "orbfe_extract_batch")
        m.stereo_match_batch_device(ex, P, bf, b, ur.data_ptr(), dp.data_ptr())
        with torch.cuda.stream(s_m):
            hur.copy_(ur, non_blocking=True)
            hdp.copy_(dp, non_blocking=True)
        m.synchronize()
    for _ in range(Wm):
        e2e_step()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    e2e_s = time.perf_counter() - t0
    b_alg_pair = 2 * base["roofline"]["algorithmic_bytes_per_frame"] + 8 * base["mean_keypoints_per_frame"]
    stage = dict(base["roofline"]["stage_ms_per_step"])
    stage["stereo_match"] = ms_stereo
    top = max(stage, key=stage.get)
    achieved = b_alg_pair * P / (stage[top] / 1e3) / 1e9
    whole = b_alg_pair * P * K / (ms / 1e3) / 1e9
    out = {
        "metric": "stereo pairs/s: ORB extract+describe of both eyes (752x480, 1200 features per eye) + Frame::ComputeStereoMatches",
        "value": K * P / (ms / 1e3), "unit": "pairs/s", "ms_per_step": ms / K, "gpu_launches": int(launches),
        "config": {"workload": "BASELINE config 1: EuRoC-shape stereo pair 752x480, nFeatures=1200 per eye, ORBmatcher::SearchStereo "
                               "(= Frame::ComputeStereoMatches) row-band matching; batched: one extraction launch set over all eyes, "
                               "one stereo launch set over all pairs", "pairs_per_step": P, "distinct_pairs": nd,
                   "mean_keypoints_per_eye": base["mean_keypoints_per_frame"], "mean_depths_per_pair": n_depth,
                   "parity_pair0_equals_reference_frame_constructor": parity},
        "e2e": {"value": K * P / e2e_s, "unit": "pairs/s", "h2d_bytes_per_step": int(2 * P * W * H),
                "d2h_bytes_per_step": int(2 * P * cap * 60 + 16 * P + 8 * P * cap), "ms_per_step": 1e3 * e2e_s / K,
                "api": "orbfe_extract_batch (pinned host frames) + orbfe_stereo_match_batch_device + D2H of mvuRight / mvDepth"},
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "algorithmic_bytes_per_pair": b_alg_pair, "pairs_per_launch": P, "kernel_ms_per_launch": stage[top],
                     "whole_path_achieved": whole, "whole_path_frac": whole / peak, "stage_ms_per_step": stage,
                     "extraction_only_frames_per_s": base["value"]},
    }
    if not args.no_cpu and ob.ref_frame_available():
        K4 = np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32)
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < max(args.cpu_seconds / 4, 2.0):
            ob.ref_stereo_frame(pairs[n % nd][0], pairs[n % nd][1], (NF,) + mc.ORB, K4, np.zeros(4, np.float32), bf)
            n += 1
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dt, "unit": "pairs/s", "cores": 2, "kind": "reference", "ms_per_pair": 1e3 * dt / n,
                               "sample": f"{n} stereo Frame constructors (src/Frame.cc:98-200 of the reference, oracle/_ref/liborbslam3_frame.so: "
                                         "its ORBextractor on two threads as the reference does, UndistortKeyPoints, ComputeStereoMatches), one after "
                                         "the other; OpenCV primitives are the scalar restatement"}
    return out


# ---- config 4 (0-based): 1920x1080, 4000 features per frame -----------------------------------------------------------------
def run_hd(args, torch, orbfe, local, peak):
    mc = _cases()
    W, H, NF = 1920, 1080, 4000
    B, K, Wm = args.hd_frames, max(3, args.steps // 3), args.warmup
    nd = min(B, 8)
    base = orbfe.synth.frames(1400, nd, W, H)
    frames = np.stack([base[i % nd] for i in range(B)])
    ex = orbfe.ORBextractor(NF, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=B)
    r, res, dev, host = _extract_throughput(torch, orbfe, ex, frames, W, H, (0, 0), K, Wm, peak)
    ob = _oracle()
    parity = None
    if ob.ref_available():
        smono, skps, sdesc = ob.OrbSlam3Extractor(NF, *mc.ORB)(base[0], (0, 0))
        parity = bool(len(skps) == len(res[0][1]) and
                      all(np.array_equal(skps[f], res[0][1][f]) for f in ("x", "y", "size", "response", "octave")) and
                      np.all(sdesc == res[0][2], axis=1).mean() >= 0.9999)
    out = {"metric": "ORB extract+describe frames/s (1920x1080, 4000 features)", "value": r["value"], "unit": "frames/s",
           "ms_per_step": r["ms_per_step"], "gpu_launches": r["gpu_launches"],
           "config": {"workload": "BASELINE config 4: batched 1920x1080 synthetic frames, nFeatures=4000, 8 levels, scale 1.2", "frames_per_step": B,
                      "distinct_frames": nd, "mean_keypoints_per_frame": r["mean_keypoints_per_frame"],
                      "parity_frame0_equals_reference_ORBextractor": parity},
           "e2e": r["e2e"], "roofline": r["roofline"]}
    if not args.no_cpu:
        import bench
        cores = os.cpu_count() or 1
        kind, extract_many, what = bench.cpu_extractor()
        cf = np.stack([base[i % nd] for i in range(cores)])
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < max(args.cpu_seconds / 4, 2.0):
            extract_many((NF,) + mc.ORB, cf, (0, 0), cores)
            n += len(cf)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dt, "unit": "frames/s", "cores": cores, "kind": kind,
                               "sample": f"{n} frames of 1920x1080 / 4000 features, frame-parallel over {cores} host threads; {what}"}
    return out


# ---- config 3: KITTI-shape 1241x376, 2000 features + SearchByBoW against a 50-keyframe window -----------------------------
def run_kitti_bow(args, torch, orbfe, local, peak):
    mc = _cases()
    cam = mc.KITTI
    W, H, NF, NKF = cam["w"], cam["h"], 2000, 50
    K, Wm = max(3, args.steps // 3), args.warmup
    # frame 0 = the query frame, frames 1..50 = the covisibility window: three scenes under growing camera shifts (the generator allows 0..24 px)
    frames = np.stack([orbfe.synth.frame(1300 + s // 25, W, H, 0, s % 25) for s in range(NKF + 1)])
    ex = orbfe.ORBextractor(NF, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=NKF + 1)
    r, res, dev, host = _extract_throughput(torch, orbfe, ex, frames, W, H, (0, 0), K, Wm, peak)
    m = orbfe.ORBmatcher(0.7, True, device=local)
    _, fk, fd = res[0]
    ffv = mc.fv_of(fd)
    kfs = []
    for i in range(1, NKF + 1):
        _, k, d = res[i]
        kfs.append((d, mc.fv_of(d), orbfe.synth.mask(1300 * 3 + i, len(d), 60), k))

    def node_pairs(fa, ma, fb):     # descriptor pairs the search evaluates: per shared node, (features with a MapPoint) x (frame features)
        ia = {int(n): k for k, n in enumerate(fa.node_ids)}
        tot = 0
        for kb, n in enumerate(fb.node_ids):
            ka = ia.get(int(n))
            if ka is None:
                continue
            a = fa.indices[fa.offsets[ka]:fa.offsets[ka + 1]]
            tot += int(ma[a].sum()) * int(fb.offsets[kb + 1] - fb.offsets[kb])
        return tot
    pairs_per_call = sum(node_pairs(kf[1], kf[2], ffv) for kf in kfs)
    for _ in range(Wm):
        b2a, cnt = m.SearchByBoW_KFs_F(kfs, fd, ffv, fk)
    l0 = m.launch_count()
    reps = max(20, 4 * K)
    t0 = time.perf_counter()
    for _ in range(reps):
        b2a, cnt = m.SearchByBoW_KFs_F(kfs, fd, ffv, fk)
    dt = time.perf_counter() - t0
    launches = m.launch_count() - l0
    popc = orbfe.measure_popc_peak(local)
    ob = _oracle()
    out = {
        "metric": "SearchByBoW: one frame against a 50-keyframe window, calls/s (1241x376, 2000 features; host buffers in and out)",
        "value": reps / dt, "unit": "calls/s", "ms_per_call": 1e3 * dt / reps, "gpu_launches": int(launches),
        "config": {"workload": "BASELINE config 3: KITTI-shape 1241x376, nFeatures=2000; ORBmatcher::SearchByBoW(KeyFrame*, Frame&) of the frame "
                               "against 50 synthetic keyframes in ONE orbfe_search_bow call (nprob = 50, b_shared = 1), FeatureVector = stand-in "
                               "100-node partition, MapPoint mask 60 %", "keyframes": NKF, "mean_features": float(np.mean([len(x[1]) for x in res])),
                   "descriptor_pairs_per_call": pairs_per_call, "mean_matches_per_keyframe": float(np.mean(cnt))},
        "e2e": {"value": reps / dt, "unit": "calls/s", "h2d_bytes_per_step": int(sum(len(k[0]) for k in kfs) * 53 + len(fd) * 48),
                "d2h_bytes_per_step": int(4 * (sum(len(k[0]) for k in kfs) + NKF * len(fd))),
                "api": "orbfe_search_bow (host buffers): the timed call IS the host call, value == e2e"},
        "roofline": {"bound": "int-popc", "kernel": "k_bow_nodes", "achieved": pairs_per_call * 8 * reps / dt / 1e12, "peak": popc / 1e12,
                     "unit": "Tpopc/s", "frac": pairs_per_call * 8 * reps / dt / popc, "traffic": None,
                     "note": "a 50-keyframe window is 1.1 M descriptor pairs = 0.02 ms of POPC work: the call is bound by launch + copy latency"},
        "extraction": {"value": r["value"], "unit": "frames/s", "e2e": r["e2e"], "roofline": r["roofline"],
                       "frames_per_step": NKF + 1},
    }
    if not args.no_cpu and ob.ref_matcher_available():
        t0, n = time.perf_counter(), 0
        rn = []
        while time.perf_counter() - t0 < max(args.cpu_seconds / 6, 1.0):
            rn = [ob.ref_search_bow(0, kf[0], kf[1], kf[2], kf[3]["angle"], fd, ffv, None, fk["angle"], 0.7, True)[2] for kf in kfs]
            n += 1
        dtc = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dtc, "unit": "calls/s", "cores": 1, "kind": "reference", "ms_per_call": 1e3 * dtc / n,
                               "sample": f"{n} x 50 ORBmatcher::SearchByBoW(KeyFrame*, Frame&) calls of the reference's own src/ORBmatcher.cc "
                                         "(oracle/_ref/liborbslam3_orbmatcher.so), sequential on one thread as Tracking::Relocalization runs them"}
        out["config"]["parity_match_counts_equal_reference"] = bool(list(rn) == list(cnt))
    return out


# ---- config 2: TUM-VI-shape fisheye stereo 512x512, 1500 features, lapping split + SearchForTriangulation ----------------------
def run_fisheye(args, torch, orbfe, local, peak):
    mc = _cases()
    ob = _oracle()
    W = H = 512
    NF, T = 1500, 8
    lapL, lapR = (128, 511), (0, 383)
    # one scene seen from a rig that moves a few pixels per frame: consecutive keyframes share most of their content
    ims = [(orbfe.synth.frame(1200, W, H, 0, 2 * t), orbfe.synth.frame(1200, W, H, 1, 2 * t)) for t in range(T)]
    ex = orbfe.ORBextractor(NF, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=1)
    m = orbfe.ORBmatcher(0.6, True, device=local)
    s2 = ex.GetScaleSigmaSquares()
    sc = ex.GetScaleFactors()
    R, t = mc.TLR[:, :3].copy(), mc.TLR[:, 3].copy()

    def frame(i):
        monoL, kL, dL = ex(ims[i][0], None, lapL)
        monoR, kR, dR = ex(ims[i][1], None, lapR)
        l2r, r2l, depth, p3d, n = m.ComputeStereoFishEyeMatches(kL, dL, monoL, kR, dR, monoR, mc.KB8_LEFT, mc.KB8_RIGHT, R, t, s2)
        return kL, dL, kR, dR, n
    kfs = [frame(i) for i in range(T)]          # also the warm-up
    l0 = ex.launch_count() + m.launch_count()
    reps = max(3, args.steps // 6)
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in range(T):
            frame(i)
    dt = time.perf_counter() - t0
    launches = ex.launch_count() + m.launch_count() - l0
    # SearchForTriangulation_ between consecutive keyframes of the rig (5 cm forward-left motion, stand-in FeatureVectors)
    Tlr = np.eye(4)
    Tlr[:3] = mc.TLR.astype(np.float64)
    T2 = np.eye(4)
    T2[:3, :3] = mc._rot(0.01, -0.015, 0.008)
    T2[:3, 3] = (0.05, 0.005, 0.01)
    R12, t12 = np.zeros((4, 9), np.float32), np.zeros((4, 3), np.float32)
    for c in range(4):
        Ta = Tlr if c >> 1 else np.eye(4)
        Tb = T2 @ Tlr if c & 1 else T2
        T12 = np.linalg.inv(Ta) @ Tb
        R12[c], t12[c] = T12[:3, :3].reshape(9), T12[:3, 3]
    ep = mc._kb8_project64(mc.KB8_LEFT, np.linalg.inv(T2)[:3, 3][None, :])[0]
    ep = (float(np.float32(ep[0])), float(np.float32(ep[1])))
    cam = np.stack([mc.KB8_LEFT, mc.KB8_RIGHT]).astype(np.float32)
    sides = []
    for kL, dL, kR, dR, _ in kfs:
        k, d = np.concatenate([kL, kR]), np.concatenate([dL, dR])
        sides.append((k, d, mc.fv_of(d), orbfe.synth.mask(77 + len(sides), len(d), 40), len(kL)))

    def tri(a, b):
        return m.SearchForTriangulationKB8(a[1], a[2], a[3], a[0], b[1], b[2], b[3], b[0], cam, cam, a[4], b[4], R12, t12, ep, s2, s2, sc,
                                           False, False)
    ntri = [tri(sides[i], sides[i + 1])[1] for i in range(T - 1)]
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in range(T - 1):
            tri(sides[i], sides[i + 1])
    dtt = time.perf_counter() - t0
    out = {
        "metric": "stereo-fisheye frames/s: ORB extract+describe of both eyes with their lapping areas (512x512, 1500 features) + "
                  "Frame::ComputeStereoFishEyeMatches, through the host-buffer calls",
        "value": reps * T / dt, "unit": "frames/s", "ms_per_frame": 1e3 * dt / (reps * T), "gpu_launches": int(launches),
        "config": {"workload": "BASELINE config 2: TUM-VI-shape fisheye stereo 512x512, nFeatures=1500, vLappingArea left [128, 511] / right [0, 383], "
                               "KB8 intrinsics + Tlr of Examples/Stereo/TUM_512.yaml; then SearchForTriangulation_ between consecutive keyframes",
                   "frames": T, "mean_stereo_matches": float(np.mean([k[4] for k in kfs])),
                   "mean_features_per_rig": float(np.mean([len(s[0]) for s in sides]))},
        "e2e": {"value": reps * T / dt, "unit": "frames/s", "h2d_bytes_per_step": int(2 * W * H), "d2h_bytes_per_step": int(2 * ex.cap * 60),
                "api": "2 x orbfe_extract + orbfe_fisheye_stereo_match per frame (host buffers): the timed calls ARE the host calls, value == e2e"},
        "roofline": {"bound": "latency", "kernel": "per-frame call chain (2 extraction graphs + k_knn2 + k_fisheye_triangulate)", "achieved": None,
                     "peak": None, "unit": None, "frac": None, "traffic": None,
                     "note": "one 512x512 frame pair per call: ~30 short launches, bound by launch latency and the 4 host<->device copies, not by a pipe"},
        "search_for_triangulation": {"metric": "SearchForTriangulation_ (Kannala-Brandt rig, 4 camera pairings) calls/s", "value": reps * (T - 1) / dtt,
                                     "unit": "calls/s", "ms_per_call": 1e3 * dtt / (reps * (T - 1)), "mean_pairs_found": float(np.mean(ntri)),
                                     "api": "orbfe_search_triangulation_kb8 (host buffers)"},
    }
    if not args.no_cpu and ob.ref_frame_available():
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < max(args.cpu_seconds / 6, 1.5):
            ob.ref_fisheye_frame(ims[n % T][0], ims[n % T][1], (NF,) + mc.ORB, mc.KB8_LEFT, mc.KB8_RIGHT, lapL, lapR, mc.TLR)
            n += 1
        dtc = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dtc, "unit": "frames/s", "cores": 2, "kind": "reference", "ms_per_frame": 1e3 * dtc / n,
                               "sample": f"{n} stereo-fisheye Frame constructors (src/Frame.cc:1017-1114 of the reference, oracle/_ref: two extraction "
                                         "threads, ComputeStereoFishEyeMatches with cv::BFMatcher + KannalaBrandt8::TriangulateMatches), one after the other"}
        a, b = sides[0], sides[1]
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < 1.0:
            ob.search_triangulation_kb8(a[0], a[1], a[2], a[3], s2, b[0], b[1], b[2], b[3], s2, sc, cam, cam, a[4], b[4], R12, t12, ep, True, False, False)
            n += 1
        out["search_for_triangulation"]["cpu_baseline"] = {"value": n / (time.perf_counter() - t0), "unit": "calls/s", "cores": 1, "kind": "port",
                                                           "sample": f"{n} calls of oracle/cpu_ref (no single reference build runs SearchForTriangulation_ "
                                                                     "together with the KannalaBrandt8 camera, DESIGN.md section 2)"}
    return out


# ---- the drop-in call: one image (or one stereo pair) per call, as Frame::ExtractORB makes it ----------------------------------
def run_latency(args, torch, orbfe, local):
    mc = _cases()
    ob = _oracle()
    W, H = 752, 480
    L = orbfe.lib()
    calls = 400
    out = {"metric": "latency of the drop-in calls (host image in, host keypoints + descriptors out)", "unit": "ms",
           "config": {"workload": "orbfe_extract on one 752x480 image (nFeatures 1000), and orbfe_extract_batch(2) + orbfe_stereo_match on one "
                                  "EuRoC stereo pair (nFeatures 1200): what Frame::ExtractORB / the stereo Frame constructor cost per frame",
                      "calls": calls, "images": "pageable numpy buffers (as a cv::Mat is)", "pipeline": "one H2D, one CUDA-graph launch, one D2H of the "
                      "result arena, one synchronise"}}
    # mono
    ex = orbfe.ORBextractor(1000, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=1)
    imgs = [orbfe.synth.frame(1000 + i, W, H) for i in range(8)]
    kps, desc, n = np.zeros(ex.cap, orbfe.KEYPOINT_DTYPE), np.zeros((ex.cap, 32), np.uint8), C.c_int(0)

    def mono(i):
        im = imgs[i % 8]
        return L.orbfe_extract(ex.handle, im.ctypes.data, W, H, W, 0, 1000, kps.ctypes.data, desc.ctypes.data, ex.cap, C.byref(n))
    for i in range(20):
        assert mono(i) >= 0
    g0, l0 = ex.graph_launch_count(), ex.launch_count()
    ts = []
    for i in range(calls):
        t0 = time.perf_counter()
        mono(i)
        ts.append(time.perf_counter() - t0)
    out["extract_1"] = dict(_percentiles(ts), api="orbfe_extract", graph_launches_per_call=(ex.graph_launch_count() - g0) / calls,
                            kernels_per_call=(ex.launch_count() - l0) / calls)
    ex.set_pipeline(0, 0, 0)           # the same call without the CUDA graph: 13 separate launches
    for i in range(20):
        mono(i)
    ts = []
    for i in range(calls):
        t0 = time.perf_counter()
        mono(i)
        ts.append(time.perf_counter() - t0)
    out["extract_1_no_graph"] = _percentiles(ts)
    # stereo pair through one handle (batch of 2) + the stereo matcher
    cam = mc.EUROC
    bf = float(np.float32(cam["bf"]))
    b = float(np.float32(np.float32(cam["bf"]) / np.float32(cam["fx"])))
    ex2 = orbfe.ORBextractor(1200, *mc.ORB, device=local, max_width=W, max_height=H, max_batch=2)
    m = orbfe.ORBmatcher(0.6, True, device=local)
    pairs = [np.stack(orbfe.synth.stereo_pair(1100 + i, W, H)) for i in range(8)]
    k2, d2 = np.zeros((2, ex2.cap), orbfe.KEYPOINT_DTYPE), np.zeros((2, ex2.cap, 32), np.uint8)
    n2, m2 = np.zeros(2, np.int32), np.zeros(2, np.int32)
    ur, dp = np.zeros(ex2.cap, np.float32), np.zeros(ex2.cap, np.float32)

    def stereo(i):
        im = pairs[i % 8]
        st = L.orbfe_extract_batch(ex2.handle, im.ctypes.data, 2, W, H, W, W * H, 0, 0, k2.ctypes.data, d2.ctypes.data, ex2.cap, n2.ctypes.data,
                                   m2.ctypes.data)
        assert st >= 0
        return L.orbfe_stereo_match(m.handle, ex2.handle, 0, ex2.handle, 1, bf, b, ur.ctypes.data, dp.ctypes.data, ex2.cap)
    for i in range(20):
        assert stereo(i) >= 0
    ts = []
    for i in range(calls):
        t0 = time.perf_counter()
        stereo(i)
        ts.append(time.perf_counter() - t0)
    out["stereo_pair"] = dict(_percentiles(ts), api="orbfe_extract_batch(2) + orbfe_stereo_match")
    out["value"] = out["extract_1"]["p50_ms"]
    if not args.no_cpu:
        import bench
        kind, extract_many, what = bench.cpu_extractor()
        ts = []
        for i in range(24):
            t0 = time.perf_counter()
            extract_many((1000,) + mc.ORB, imgs[i % 8][None], (0, 1000), 1)
            ts.append(time.perf_counter() - t0)
        out["cpu_baseline"] = {"extract_1": _percentiles(ts[4:]), "cores": 1, "kind": kind, "unit": "ms",
                               "sample": "one image per call on one thread (what the reference does per frame); " + what}
        if ob.ref_frame_available():
            K4 = np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32)
            ts = []
            for i in range(16):
                t0 = time.perf_counter()
                ob.ref_stereo_frame(pairs[i % 8][0], pairs[i % 8][1], (1200,) + mc.ORB, K4, np.zeros(4, np.float32), bf)
                ts.append(time.perf_counter() - t0)
            out["cpu_baseline"]["stereo_pair"] = dict(_percentiles(ts[2:]), cores=2,
                                                      what="the reference's stereo Frame constructor (two extraction threads + ComputeStereoMatches)")
    return out
