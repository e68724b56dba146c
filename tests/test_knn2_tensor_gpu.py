"""The tensor-core 2-NN (csrc/k_knn2_tc.cu: tcgen05.mma kind::i8, accumulator = 256 * hamming + index) against the POPC kernel
and the oracle: indices and distances bit-identical (integer arithmetic), ties to the lowest train index, ragged sizes."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe

pytestmark = pytest.mark.gpu


def both_backends(q, t):
    m = orbfe.ORBmatcher(0.7, True)
    m.set_knn2_backend(m.KNN2_POPC)
    a = m.knn2(q, t)
    m.set_knn2_backend(m.KNN2_TENSOR)
    b = m.knn2(q, t)
    return a, b


@pytest.mark.parametrize("nq,nt", [(700, 3000), (4000, 4000), (128, 256), (129, 257), (5, 1), (1, 300), (300, 2), (2047, 513), (4100, 9000)])
def test_tensor_knn2_equals_popc_and_oracle(nq, nt):
    q, t = orbfe.synth.descriptors(100 + nq, nq), orbfe.synth.descriptors(200 + nt, nt)
    if nt > 300:
        t[100] = t[7]
        t[nt - 1] = t[7]
        q[0] = t[7]                 # three-way tie at distance 0: lowest index first, then the next
        t[255:259] = q[min(3, nq - 1)]      # a tie across a 256-row tile boundary
    (pi, pd), (ti, td) = both_backends(q, t)
    assert np.array_equal(ti, pi) and np.array_equal(td, pd)
    ri, rd = ob.knn2(q, t)
    assert np.array_equal(ti, ri) and np.array_equal(td, rd)


def test_tensor_knn2_extreme_distances():
    """distance 0 and 256 (all bits differ) and everything between on structured descriptors."""
    nq = 257
    q = np.zeros((nq, 32), np.uint8)
    t = np.zeros((nq, 32), np.uint8)
    for i in range(nq):             # train row i has its first i bits set
        bits = np.zeros(256, np.uint8)
        bits[:min(i, 256)] = 1
        t[i] = np.packbits(bits, bitorder="little")
    q[:] = t
    q[0] = 0xFF                     # distance 256 to train row 0
    (pi, pd), (ti, td) = both_backends(q, t)
    assert np.array_equal(ti, pi) and np.array_equal(td, pd)
    assert td[5, 0] == 0 and ti[5, 0] == 5 and td[0].tolist() == [0, 1] and ti[0].tolist() == [256, 255]


def test_tensor_knn2_batch_device_ragged():
    import torch
    B, W, H = 5, 752, 480
    frames = orbfe.synth.frames(2300, B, W, H)
    ex = orbfe.ORBextractor(1000, *mc.ORB, max_width=W, max_height=H, max_batch=B)
    res = ex.extract_batch(frames, (0, 0))
    d_kps, d_desc, d_n, d_mono, cap = ex.device_results()
    out = []
    for backend in (0, 1):
        m = orbfe.ORBmatcher(0.7, True)
        m.set_knn2_backend(backend)
        idx = torch.full((B - 1, cap, 2), -5, dtype=torch.int32, device="cuda")
        dist = torch.full((B - 1, cap, 2), -5, dtype=torch.int32, device="cuda")
        ex.order_consumer(m.stream())
        m.knn2_batch_device(d_desc, cap * 32, d_n, cap, d_desc + cap * 32, cap * 32, d_n + 4, cap, B - 1, idx.data_ptr(), dist.data_ptr())
        m.synchronize()
        out.append((idx.cpu().numpy(), dist.cpu().numpy()))
    for b in range(B - 1):
        n = len(res[b][1])
        assert np.array_equal(out[0][0][b, :n], out[1][0][b, :n]) and np.array_equal(out[0][1][b, :n], out[1][1][b, :n])
        ri, rd = ob.knn2(res[b][2], res[b + 1][2])
        assert np.array_equal(out[1][0][b, :n], ri) and np.array_equal(out[1][1][b, :n], rd)


def test_tensor_knn2_empty_train():
    q = orbfe.synth.descriptors(1, 300)
    (pi, pd), (ti, td) = both_backends(q, q[:0])
    assert np.array_equal(ti, pi) and np.array_equal(td, pd) and (ti == -1).all()
