"""Host pipeline of the extractor (csrc/orbfe_api.cu): chunked multi-stream device path, CUDA graphs, the one-copy
small-batch path and the stream-ordering calls.  Every configuration must give byte-identical results."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 752, 480


def _same(a, b):
    assert len(a) == len(b)
    for (ma, ka, da), (mb, kb, db) in zip(a, b):
        assert ma == mb
        assert ka.tobytes() == kb.tobytes()
        assert da.tobytes() == db.tobytes()


def _device_frames(frames):
    import torch
    return torch.from_numpy(frames).cuda()


@pytest.mark.parametrize("chunk,streams,graphs", [(1000, 1, 0), (7, 1, 0), (7, 3, 0), (7, 3, 1), (5, 4, 1), (32, 2, 1)])
def test_chunking_streams_graphs_identical(chunk, streams, graphs):
    import orbfe
    B = 23
    frames = orbfe.synth.frames(4100, B, W, H)
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    ex.set_pipeline(1000, 1, 0)
    dev = _device_frames(frames)
    ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=(0, 1000))
    ref = ex.fetch_results(B)
    ex.set_pipeline(chunk, streams, graphs)
    g0 = ex.graph_launch_count()
    for rep in range(4):   # rep 0 eager, rep 1 captures + launches, reps 2-3 replay
        ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=(0, 1000))
        _same(ex.fetch_results(B), ref)
    assert (ex.graph_launch_count() - g0 == 3) if graphs else (ex.graph_launch_count() == g0)
    # a different lapping range must not replay the old graph
    ex.extract_batch_device(dev.data_ptr(), B, W, H, vLappingArea=(0, 0))
    other = ex.fetch_results(B)
    assert other[0][0] != ref[0][0] or len(ref[0][1]) == 0
    # the host-buffer path of a large-batch handle (chunked copy pipeline) agrees as well
    _same(ex.extract_batch(frames, (0, 1000)), ref)


def test_small_batch_path_equals_large_handle():
    """max_batch <= 8 handles take the one-copy path (pinned mirror of the result arena)."""
    import orbfe
    frames = orbfe.synth.frames(4200, 2, W, H)
    big = orbfe.ORBextractor(1200, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=16)
    ref = big.extract_batch(frames, (0, 0))
    for mb in (1, 2, 8):
        ex = orbfe.ORBextractor(1200, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=mb)
        for rep in range(3):
            if mb == 1:
                got = [ex(frames[i], None, (0, 0)) for i in range(2)]
            else:
                got = ex.extract_batch(frames, (0, 0))
            _same(got, ref)
        # strided rows (a cv::Mat ROI): same result
        wide = np.zeros((H, W + 40), np.uint8)
        wide[:, :W] = frames[0]
        _same([ex(wide[:, :W], None, (0, 0))], ref[:1])


def test_shape_change_drops_graphs():
    import orbfe
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    a = orbfe.synth.frame(4300, W, H)
    b = orbfe.synth.frame(4301, 640, 400)
    ra = [ex(a, None, (0, 0)) for _ in range(3)]
    rb = [ex(b, None, (0, 0)) for _ in range(3)]
    ra2 = [ex(a, None, (0, 0)) for _ in range(3)]
    _same(ra[:1] * 3, ra)
    _same(rb[:1] * 3, rb)
    _same(ra, ra2)
    assert len(ra[0][1]) != len(rb[0][1]) or ra[0][1].tobytes() != rb[0][1].tobytes()


def test_chained_device_consumers_without_host_sync():
    """extract_batch_device -> knn2 / BoW / stereo on the device results -> next extract, no host synchronise in
    between: the ordering calls (orbfe_stream_wait_extractor / orbfe_extractor_wait_stream) make it equal to the
    synchronised sequence (consumers ran on their own streams with nothing ordering them)."""
    import torch
    import orbfe
    B = 8
    fa = orbfe.synth.frames(4400, B, W, H)
    fb = orbfe.synth.frames(4500, B, W, H)
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    m = orbfe.ORBmatcher(0.7, True)
    da, db = _device_frames(fa), _device_frames(fb)
    cap = ex.cap

    def knn(sync):
        idx = torch.full((B - 1, cap, 2), -7, dtype=torch.int32, device="cuda")
        dst = torch.full((B - 1, cap, 2), -7, dtype=torch.int32, device="cuda")
        ur = torch.zeros((B // 2, cap), dtype=torch.float32, device="cuda")
        dp = torch.zeros((B // 2, cap), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        ex.extract_batch_device(da.data_ptr(), B, W, H)
        if sync:
            ex.synchronize()
        else:
            ex.order_consumer(m.stream())
        d_kps, d_desc, d_n, d_mono, c = ex.device_results()
        m.knn2_batch_device(d_desc, c * 32, d_n, c, d_desc + c * 32, c * 32, d_n + 4, c, B - 1, idx.data_ptr(), dst.data_ptr())
        m.stereo_match_batch_device(ex, B // 2, 47.9, 0.11, ur.data_ptr(), dp.data_ptr())
        if sync:
            m.synchronize()
        else:
            ex.wait_consumer(m.stream())
        ex.extract_batch_device(db.data_ptr(), B, W, H)     # overwrites what the matcher reads
        ex.synchronize()
        m.synchronize()
        n = torch.empty(B, dtype=torch.int32)
        return idx.cpu().numpy(), dst.cpu().numpy(), ur.cpu().numpy(), dp.cpu().numpy()

    want = knn(True)
    for _ in range(5):
        got = knn(False)
        for a, b in zip(want, got):
            assert np.array_equal(a, b)
