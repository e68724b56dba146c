"""k_blur_tc (the 7x7 Gaussian as banded int8 GEMMs on tcgen05.mma) against k_blur and against the oracle's restatement of
OpenCV's fixed-point GaussianBlur (held to cv2 4.13 by tests/test_oracle_vs_cv2.py): the blurred levels are the same bytes,
and so are the descriptors sampled from them."""
import ctypes as C

import numpy as np
import pytest

import orbfe
import oracle_binding as ob

pytestmark = pytest.mark.gpu

SHAPES = [(752, 480, 1000), (512, 512, 1500), (1241, 376, 2000), (1920, 1080, 4000), (262, 251, 300), (385, 257, 400)]


def oracle_blur(img):
    out = np.zeros_like(img)
    ob.lib().orbref_blur(img.ctypes.data_as(C.c_void_p), img.shape[1], img.shape[0], img.strides[0],
                         out.ctypes.data_as(C.c_void_p), out.strides[0])
    return out


@pytest.mark.parametrize("w,h,nf", SHAPES, ids=[f"{s[0]}x{s[1]}" for s in SHAPES])
def test_blurred_levels_equal_oracle_and_alu_kernel(w, h, nf):
    ex = orbfe.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=3)
    frames = np.stack([orbfe.synth.frame(5200 + i, w, h) for i in range(3)])
    out = {}
    for backend in (ex.BLUR_TENSOR, ex.BLUR_ALU):      # tensor first: the blurred arena is zero until a kernel writes it
        ex.set_blur_backend(backend)
        res = ex.extract_batch(frames, (0, 0))
        lv = [[ex.blurred_level(l, f) for l in range(ex.nlevels)] for f in range(3)]
        out[backend] = (res, lv)
    for f in range(3):
        for l in range(ex.nlevels):
            a, t = out[ex.BLUR_ALU][1][f][l], out[ex.BLUR_TENSOR][1][f][l]
            _, roi = ex.image_pyramid(l, f)
            ref = oracle_blur(np.ascontiguousarray(roi))
            assert np.array_equal(t, ref), f"tensor blur, frame {f} level {l}: {np.argwhere(t != ref)[:5]}"
            assert np.array_equal(a, ref), f"ALU blur, frame {f} level {l}"
    ra, rt = out[ex.BLUR_ALU][0], out[ex.BLUR_TENSOR][0]
    for f in range(3):
        assert np.array_equal(ra[f][2], rt[f][2])          # descriptors
        assert np.array_equal(ra[f][1]["angle"], rt[f][1]["angle"])


def test_tensor_blur_large_batch_matches_alu():
    """More tiles than SMs (persistent CTAs walk several tiles each, both pipeline stages wrap)."""
    w, h = 752, 480
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=24)
    frames = np.stack([orbfe.synth.frame(5300 + i % 5, w, h) for i in range(24)])
    ex.set_blur_backend(ex.BLUR_TENSOR)                # first, on the still-zero arena
    rt = ex.extract_batch(frames, (0, 1000))
    lt = [ex.blurred_level(l, 23) for l in range(8)]
    assert all(x.any() for x in lt)
    ex.set_blur_backend(ex.BLUR_ALU)
    ra = ex.extract_batch(frames, (0, 1000))
    la = [ex.blurred_level(l, 23) for l in range(8)]
    for l in range(8):
        assert np.array_equal(la[l], lt[l]), l
    for f in range(24):
        assert np.array_equal(ra[f][2], rt[f][2]), f
