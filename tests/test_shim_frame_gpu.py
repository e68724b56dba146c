"""The reference's own src/Frame.cc, compiled UNMODIFIED against the orbfe shim headers (include/orbfe_shim/ORBextractor.h,
ORBmatcher.h) and linked to liborbfe.so (tests/cpp/_build/liborbfe_shim_frame.so, `make -C oracle shim`), against the same
Frame.cc built with the reference's own ORBextractor.cc (oracle/_ref/liborbslam3_frame.so).  The monocular, rectified-stereo
and stereo-fisheye Frame constructors run from images; the stereo one calls the unmodified Frame::ComputeStereoMatches, which
reads mvImagePyramid of the shim extractors on the host.  mvKeys, mDescriptors, mvKeysUn, mvuRight, mvDepth must be identical."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
import orbfe

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (ob.ref_frame_available() and ob.shim_frame_available()),
                                                  reason="oracle/_ref or the shim frame library not built")]

DIST = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)


def both(fn, *a, **k):
    want = fn(*a, **k)
    with ob.shim_frame_library():
        got = fn(*a, **k)
    return want, got


def identical(want, got, keys):
    for key in keys:
        w, g = want[key], got[key]
        if isinstance(w, np.ndarray):
            assert w.dtype == g.dtype and w.shape == g.shape and w.tobytes() == g.tobytes(), key
        else:
            assert w == g, key


@pytest.mark.parametrize("name", list(mc.STEREO_CASES))
def test_stereo_frame_constructor_on_the_shim(name):
    cam, nf, seed = mc.STEREO_CASES[name]
    w, h = cam["w"], cam["h"]
    imL, imR = orbfe.synth.stereo_pair(seed, w, h)
    bf = float(np.float32(cam["bf"]))
    K4 = np.array([cam["fx"], cam.get("fy", cam["fx"]), cam.get("cx", w / 2), cam.get("cy", h / 2)], np.float32)
    want, got = both(ob.ref_stereo_frame, imL, imR, (nf,) + mc.ORB, K4, DIST, bf)
    identical(want, got, ("kps", "kps_r", "desc", "desc_r", "kps_un", "u_right", "depth", "bounds"))
    assert (want["depth"] > 0).sum() > 0.4 * len(want["kps"])


def test_fisheye_frame_constructor_on_the_shim():
    from test_reference_frame import fisheye_frame_case
    imL, imR, nf, lapL, lapR = fisheye_frame_case()
    want, got = both(ob.ref_fisheye_frame, imL, imR, (nf,) + mc.ORB, mc.KB8_LEFT, mc.KB8_RIGHT, lapL, lapR, mc.TLR)
    identical(want, got, ("kps", "desc", "nleft", "nright", "mono_left", "mono_right", "l2r", "r2l", "depth", "p3d"))
    assert (want["l2r"] >= 0).sum() > 50


def test_mono_frame_constructor_on_the_shim(tmp_path):
    from test_reference_dbow2 import write_text
    img = orbfe.synth.frame(1000, 752, 480)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    parent, leaf, desc, weight = orbfe.synth.vocabulary(4242, 10, 5, 1)
    write_text(dict(k=10, L=5, weighting=0, norm=1, parent=parent, leaf=leaf, desc=desc, weight=weight), tmp_path / "voc.txt")
    want, got = both(ob.ref_mono_frame, img, (1000,) + mc.ORB, K4, DIST, tmp_path / "voc.txt")
    identical(want, got, ("kps", "desc", "kps_un"))
    for a, b in zip(want["bow"], got["bow"]):
        assert np.array_equal(a, b)
    assert len(want["kps"]) >= 1000


def test_stereo_frame_random_configs_on_the_shim():
    from test_reference_frame import _stereo_random_configs
    for w, h, nf, sf, nl, ini, mn in _stereo_random_configs(20261109, 4):
        imL, imR = orbfe.synth.stereo_pair(9300 + w + h, w, h)
        K4 = np.array([0.6 * w, 0.6 * w, w / 2 - 3.25, h / 2 + 1.5], np.float32)
        D = np.array([-0.21, 0.05, 0.0004, -0.0003], np.float32)
        bf = float(np.float32(0.11 * K4[0]))
        want, got = both(ob.ref_stereo_frame, imL, imR, (nf, sf, nl, ini, mn), K4, D, bf)
        identical(want, got, ("kps", "kps_r", "kps_un", "bounds"))
        n = len(want["desc"]) + len(want["desc_r"])
        same = (want["desc"] == got["desc"]).all(1).sum() + (want["desc_r"] == got["desc_r"]).all(1).sum()
        assert same >= 0.9999 * n
        if same == n:
            identical(want, got, ("u_right", "depth"))
