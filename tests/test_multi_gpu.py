"""1-vs-N-GPU byte identity on hardware (test plan item 4) and several devices in ONE process.

Skipped on a 1-GPU box; run with `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`.
  * one frame list sharded with orbfe.shard.chunk over every visible device (one host thread + one handle per device,
    as the reference forks one thread per eye, Frame.cc:119-122) equals the unsharded single-device run, byte for byte;
  * the > 48 KB dynamic-shared-memory kernels (quadtree at 1920x1080 / 4000 features, ~59 KB) launch on EVERY device of the
    process: cudaFuncSetAttribute is per device (the process-wide `static size_t configured` skipped device 1);
  * matchers on a second device; a NCCL process-group variant of the sharded run (torch.multiprocessing, one rank per GPU)."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import orbfe
from orbfe import shard

pytestmark = pytest.mark.gpu


def ndev():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


need2 = pytest.mark.skipif(ndev() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")


def _run_sharded(frames, params, W, H, lap, devices):
    n = len(frames)
    out = [None] * n
    errs = []

    def work(rank):
        try:
            b, e = shard.chunk(n, rank, len(devices))
            if e == b:
                return
            ex = orbfe.ORBextractor(*params, device=devices[rank], max_width=W, max_height=H, max_batch=e - b)
            for i, r in enumerate(ex.extract_batch(frames[b:e], lap)):
                out[b + i] = r
        except Exception as ex_:      # surfaced by the main thread
            errs.append((rank, repr(ex_)))
    ts = [threading.Thread(target=work, args=(r,)) for r in range(len(devices))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    return out


@need2
@pytest.mark.parametrize("W,H,nf,n", [(752, 480, 1000, 13), (1920, 1080, 4000, 5)])
def test_sharded_over_all_devices_equals_single_device(W, H, nf, n):
    params = (nf, 1.2, 8, 20, 7)
    frames = orbfe.synth.frames(7000 + W, n, W, H)
    lap = (0, 1000)
    single = _run_sharded(frames, params, W, H, lap, [0])
    multi = _run_sharded(frames, params, W, H, lap, list(range(ndev())))
    last_only = _run_sharded(frames, params, W, H, lap, [ndev() - 1])        # the last device alone, first use of it in a new handle
    for a, b, c in zip(single, multi, last_only):
        assert a[0] == b[0] == c[0]
        assert a[1].tobytes() == b[1].tobytes() == c[1].tobytes()
        assert a[2].tobytes() == b[2].tobytes() == c[2].tobytes()
    assert len(single[0][1]) >= nf


@need2
def test_matchers_on_the_second_device():
    import oracle_binding as ob
    d = ndev() - 1
    q, t = orbfe.synth.descriptors(31, 3000), orbfe.synth.descriptors(32, 5000)
    m = orbfe.ORBmatcher(0.7, True, device=d)
    idx, dist = m.knn2(q, t)
    ridx, rdist = ob.knn2(q, t)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    # the stereo matcher uses > 48 KB of dynamic shared memory for large right-keypoint tables
    W, H = 1920, 1080
    imL, imR = orbfe.synth.stereo_pair(1400, W, H)
    res = []
    for dev in (0, d):
        ex = orbfe.ORBextractor(6000, 1.2, 8, 20, 7, device=dev, max_width=W, max_height=H, max_batch=2)
        ex.extract_batch(np.stack([imL, imR]), (0, 0))
        res.append(orbfe.ORBmatcher(0.6, True, device=dev).stereo_match(ex, ex, 400.0, 0.5, 0, 1))
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]) and res[0][2] == res[1][2] > 100


@need2
def test_two_devices_two_threads_cpp():
    """tests/cpp/two_device_test.cc: two std::threads, one orbfe handle each on devices 0 and 1, 1920x1080 / 4000 features."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    build = os.path.join(root, "tests", "cpp", "_build")
    os.makedirs(build, exist_ok=True)
    exe = os.path.join(build, "two_device_test")
    libdir = os.path.join(root, "orb-slam3-noted_b200", "csrc")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-pthread", "-I" + os.path.join(root, "include"),
                        os.path.join(root, "tests", "cpp", "two_device_test.cc"), "-o", exe, "-L" + libdir, "-lorbfe", "-Wl,-rpath," + libdir,
                        "-L" + os.path.join(libdir), "-l:liborbfe_synth.so"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    p = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert p.returncode == 0, p.stdout
    assert "identical on 2 devices" in p.stdout, p.stdout


def _nccl_worker(rank, world, port, n, W, H, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    frames = orbfe.synth.frames(7100, n, W, H)
    b, e = shard.chunk(n, rank, world)
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, device=rank, max_width=W, max_height=H, max_batch=max(e - b, 1))
    local = [(b + i, r[0], r[1].tobytes(), r[2].tobytes()) for i, r in enumerate(ex.extract_batch(frames[b:e], (0, 0)))] if e > b else []
    got = shard.gather(local)
    if rank == 0:
        q.put(got)
    dist.barrier()
    dist.destroy_process_group()


@need2
def test_one_rank_per_gpu_nccl_equals_unsharded():
    import torch.multiprocessing as mp
    world, n, W, H = ndev(), 11, 752, 480
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 2000
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, n, W, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    frames = orbfe.synth.frames(7100, n, W, H)
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=n)
    ref = ex.extract_batch(frames, (0, 0))
    assert [g[0] for g in got] == list(range(n))
    for g, r in zip(got, ref):
        assert g[1] == r[0] and g[2] == r[1].tobytes() and g[3] == r[2].tobytes()
