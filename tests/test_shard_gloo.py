"""CPU, world_size 2 over gloo: the host-side sharding logic (orbfe/shard.py).  Each rank processes its
chunk with a stand-in worker (the CPU oracle, test infrastructure) and the gathered result must equal the
unsharded run; sequence sharding must cover every consecutive pair exactly once."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

import orbfe
from orbfe import shard


def test_chunks_partition():
    for n in (0, 1, 2, 7, 8, 9, 512):
        for world in (1, 2, 3, 8):
            spans = [shard.chunk(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1


def test_sequence_chunks_cover_every_pair_once():
    for n in (1, 2, 5, 16, 17):
        for world in (1, 2, 4, 8):
            pairs = []
            for r in range(world):
                b, e, npairs = shard.sequence_chunk(n, r, world)
                assert npairs == max(e - b - 1, 0)
                pairs += [(t, t + 1) for t in range(b, b + npairs)]
            assert pairs == [(t, t + 1) for t in range(n - 1)]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_binding as ob
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 7
    frames = orbfe.synth.frames(500, n, 376, 240)
    b, e = shard.chunk(n, rank, world)
    ref = ob.OrbRef(300, 1.2, 8, 20, 7)
    local = []
    for i in range(b, e):
        mono, k, d = ref(frames[i], (0, 0))
        local.append((i, mono, k.tobytes(), d.tobytes()))
    # consecutive-frame matching, sharded by pair
    sb, se, npairs = shard.sequence_chunk(n, rank, world)
    descs = [ref(frames[i], (0, 0))[2] for i in range(sb, se)]
    local_pairs = [(sb + t, ob.knn2(descs[t], descs[t + 1])[0].tobytes()) for t in range(npairs)]
    got = shard.gather(local)
    got_pairs = shard.gather(local_pairs)
    if rank == 0:
        q.put((got, got_pairs))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_equal_unsharded():
    import oracle_binding as ob
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got, got_pairs = q.get(timeout=240)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    n = 7
    frames = orbfe.synth.frames(500, n, 376, 240)
    ref = ob.OrbRef(300, 1.2, 8, 20, 7)
    descs = []
    assert [g[0] for g in got] == list(range(n))
    for i in range(n):
        mono, k, d = ref(frames[i], (0, 0))
        descs.append(d)
        assert got[i][1] == mono and got[i][2] == k.tobytes() and got[i][3] == d.tobytes()
    assert [g[0] for g in got_pairs] == list(range(n - 1))
    for t in range(n - 1):
        assert got_pairs[t][1] == ob.knn2(descs[t], descs[t + 1])[0].tobytes()
