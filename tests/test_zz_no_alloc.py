"""orbfe_hamming_knn2_noalloc (include/orbfe.h): the allocation-free 2-NN on device buffers must give exactly what the
existing host-buffer entry point orbfe_hamming_knn2 gives (bit-exact: integer work), on two calls in a row that share
the static scratch array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(L, q, t, stream=None):
    import torch
    dq = torch.from_numpy(np.ascontiguousarray(q)).cuda()
    dt = torch.from_numpy(np.ascontiguousarray(t)).cuda() if len(t) else torch.zeros((1, 32), dtype=torch.uint8, device="cuda")
    idx = torch.full((len(q), 2), -9, dtype=torch.int32, device="cuda")
    dist = torch.full((len(q), 2), -9, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    st = L.orbfe_hamming_knn2_noalloc(0, C.c_void_p(stream), C.c_void_p(dq.data_ptr()), len(q), C.c_void_p(dt.data_ptr()), len(t),
                                      C.c_void_p(idx.data_ptr()), C.c_void_p(dist.data_ptr()))
    assert st == 0, (st, L.orbfe_last_cuda_error())
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return idx.cpu().numpy(), dist.cpu().numpy(), free0, free1


def test_noalloc_knn2_equals_host_entry_point_twice():
    import orbfe
    L = orbfe.lib()
    m = orbfe.ORBmatcher(0.7, True)
    # call 1: large problem (many splits -> the scratch is used); call 2: a smaller, different one right after it
    # (stale partial results of call 1 stay in the array); call 3: the maximum query count
    cases = [(4000, 4100, 11), (1500, 777, 12), (8192, 3000, 13), (5, 1, 14), (300, 0, 15)]
    for k, (nq, nt, seed) in enumerate(cases):
        q = orbfe.synth.descriptors(seed, nq)
        t = orbfe.synth.descriptors(seed + 100, max(nt, 1))[:nt]
        if nt > 40:       # duplicates and near-duplicates: distance ties must resolve to the lowest train index
            t[nt // 2] = t[3]
            q[7] = t[3]
        want_idx, want_dist = m.knn2(q, t)
        idx, dist, free0, free1 = _noalloc(L, q, t)
        assert np.array_equal(idx, want_idx), k
        assert np.array_equal(dist, want_dist), k
        if k > 0:          # the first call may load the module lazily; later calls must not take device memory
            assert free1 >= free0, (free0, free1)


def test_noalloc_knn2_on_a_caller_stream():
    import torch
    import orbfe
    L = orbfe.lib()
    m = orbfe.ORBmatcher(0.7, True)
    s = torch.cuda.Stream()
    q, t = orbfe.synth.descriptors(21, 2500), orbfe.synth.descriptors(22, 2600)
    want = m.knn2(q, t)
    for _ in range(2):
        idx, dist, _, _ = _noalloc(L, q, t, s.cuda_stream)
        assert np.array_equal(idx, want[0]) and np.array_equal(dist, want[1])


def test_noalloc_knn2_rejects_sizes_beyond_the_static_scratch():
    import torch
    import orbfe
    L = orbfe.lib()
    d = torch.zeros((16, 32), dtype=torch.uint8, device="cuda")
    o = torch.zeros((16, 2), dtype=torch.int32, device="cuda")
    p = lambda x: C.c_void_p(x.data_ptr())
    assert L.orbfe_hamming_knn2_noalloc(0, None, p(d), 8193, p(d), 16, p(o), p(o)) == -3      # ORBFE_ERR_TOO_LARGE
    assert b"ORBFE_NOALLOC_MAX_QUERIES" in L.orbfe_last_cuda_error()
    assert L.orbfe_hamming_knn2_noalloc(0, None, p(d), 16, p(d), 65537, p(o), p(o)) == -3
    assert b"ORBFE_NOALLOC_MAX_TRAIN" in L.orbfe_last_cuda_error()
    assert L.orbfe_hamming_knn2_noalloc(0, None, None, 16, p(d), 16, p(o), p(o)) == -2        # ORBFE_ERR_BAD_ARGUMENT
    assert L.orbfe_hamming_knn2_noalloc(0, None, p(d), 0, p(d), 16, p(o), p(o)) == 0
