"""Frame sharding across the GPUs of one box: frames, stereo pairs and sequences are
independent units, so the path shards with NO data-path collective.  One process per GPU; each rank
takes a contiguous chunk (consecutive-frame matching t <-> t+1 stays local; the first frame of the next
chunk is extracted twice rather than shipped between GPUs: 0.36 MB of pixels vs a peer copy + sync).
`gather` only collects the small per-frame results on rank 0 for callers that want one list.
"""
import torch.distributed as dist


def chunk(n_units, rank, world):
    """[begin, end) of rank's contiguous share of n_units (sizes differ by at most one)."""
    base, rem = divmod(n_units, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def sequence_chunk(n_frames, rank, world):
    """Frames a rank must EXTRACT to match every consecutive pair (t, t+1) whose t lies in its chunk:
    its own frames plus the first frame of the next chunk.  Returns (begin, end, n_pairs)."""
    b, e = chunk(n_frames - 1, rank, world) if n_frames > 1 else (0, 0)   # shard the PAIRS
    if e == b:
        return b, b, 0
    return b, e + 1, e - b


def gather(local_results, dst=0):
    """Collects per-rank result lists on `dst` in rank order (host objects: keypoints, descriptors,
    match arrays).  Returns the concatenated list on dst, None elsewhere.  Works on any backend."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(local_results)
    world, rank = dist.get_world_size(), dist.get_rank()
    out = [None] * world if rank == dst else None
    dist.gather_object(list(local_results), out, dst=dst)
    if rank != dst:
        return None
    return [r for part in out for r in part]
