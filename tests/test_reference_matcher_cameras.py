"""CPU: the reference's OWN src/ORBmatcher.cc running on its OWN camera models (src/CameraModels/Pinhole.cpp, KannalaBrandt8.cpp),
all unmodified, in one library (oracle/_ref/liborbslam3_orbmatcher_cam.so): SearchForTriangulation_ starts from keyframe POSES,
forms the epipole, the relative poses of the four camera pairings and (pinhole) F12 itself, and asks the real
GeometricCamera::epipolarConstrain_ about every candidate pair.  The oracle (and the C ABI) are handed R12 / t12 / F12 / the epipole
by the caller; this file checks that the two compositions agree."""
import numpy as np
import pytest

import match_cases as mc
import oracle_binding as ob
from oracle_binding import OrbRef

pytestmark = pytest.mark.skipif(not ob.ref_matcher_cam_available(),
                                reason="oracle/_ref not built (needs /root/reference; make -C oracle ref)")


def _sides(c, s2, sc, cam1, cam2, ur1=None, ur2=None):
    A = dict(kps=c["k1"], desc=c["d1"], fv=c["fv1"], mp=c["mp1"], nleft=c.get("nleft_a", -1), u_right=ur1, sigma2=s2, scale=sc,
             cam=cam1, Tcw=c["Tcw1"], Trl=c["Trl"])
    B = dict(kps=c["k2"], desc=c["d2"], fv=c["fv2"], mp=c["mp2"], nleft=c.get("nleft_b", -1), u_right=ur2, sigma2=s2, scale=sc,
             cam=cam2, Tcw=c["Tcw2"], Trl=c["Trl"])
    return A, B


@pytest.mark.parametrize("name", list(mc.TRI_KB8_CASES) + list(mc.TRI_KB8_EXTRA))
def test_triangulation_search_on_the_real_kannala_brandt_camera(name):
    """Rows whose candidate pairs all sit > 1e-3 (relative) away from every threshold of the triangulation test must be identical:
    the reference multiplies the poses in float32 and ends in cv::SVD, the oracle receives float32 R12 / t12 computed in double."""
    sc, isc, s2 = OrbRef(1000, *mc.ORB).tables()[:3]
    c = mc.tri_kb8_inputs(name)
    A, B = _sides(c, s2, sc, c["cam_a"], c["cam_b"])
    m12, n = ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], False, c["coarse"])
    om12, on, _, mg = ob.search_triangulation_kb8(c["k1"], c["d1"], c["fv1"], c["mp1"], s2, c["k2"], c["d2"], c["fv2"], c["mp2"], s2, sc,
                                                  c["cam_a"], c["cam_b"], c["nleft_a"], c["nleft_b"], c["R12"], c["t12"], c["ep"],
                                                  c["check_ori"], False, c["coarse"])
    assert n == (m12 >= 0).sum() and n > 60
    safe, differs = mg > 1e-3, m12 != om12
    if not (c["check_ori"] and (differs & ~safe).any()):
        assert not (differs & safe).any(), np.flatnonzero(differs & safe)
        assert n == on or (differs & ~safe).any()
    else:                                  # an unsafe row flipped and moved a rotation-histogram bin across the 10 % rule
        assert differs.mean() < 0.01
    assert safe.mean() > 0.97
    # fisheye keyframes carry no mvuRight: bOnlyStereo matches nothing (:1280-1283)
    assert ob.ref_search_for_triangulation_cam(1, A, B, 0.6, c["check_ori"], True, c["coarse"])[1] == 0


@pytest.mark.parametrize("name", [k for k, v in mc.TRI_CASES.items() if v[7] is None])
def test_triangulation_search_on_the_real_pinhole_camera(name):
    """Pinhole keyframes one pure x translation apart (x1 = x2 + (0.1, 0, 0)): the reference derives R12 = I, t12 and, inside
    Pinhole::epipolarConstrain_ (Pinhole.cpp:159-181), F12 = K1^-T [t12]x R12 K2^-1 per candidate pair; the oracle is handed the
    F12 of orbfe.synth.fundamental_x_shift.  Camera 1's centre lies in camera 2's focal plane, so the epipole is at infinity
    (the cases that place it inside the image stay with liborbslam3_orbmatcher.so, which takes the epipole as an input)."""
    from test_match_oracle import extract, tables
    sc, isc, s2, is2, per, umax = tables()
    c = mc.tri_inputs(name, extract, sc, s2)
    cam = mc.TRI_CASES[name][0]
    K = np.zeros((2, 8), np.float32)
    K[:, :4] = (cam["fx"], cam["fy"], cam["cx"], cam["cy"])
    c["Tcw1"] = np.eye(4, dtype=np.float32)
    c["Tcw2"] = np.eye(4, dtype=np.float32)
    c["Tcw2"][0, 3] = -0.1
    c["Trl"] = np.eye(4, dtype=np.float32)
    A, B = _sides(c, s2, sc, K, K, c["ur1"], c["ur2"])
    m12, n = ob.ref_search_for_triangulation_cam(0, A, B, 0.6, c["check_ori"], c["only_stereo"], c["coarse"])
    om12, on = ob.search_triangulation(c["k1"], c["d1"], c["fv1"], c["mp1"], c["ur1"], c["k2"], c["d2"], c["fv2"], c["mp2"], c["ur2"],
                                       s2, sc, c["F12"], c["ep"], c["check_ori"], c["only_stereo"], c["coarse"])
    assert n == on and np.array_equal(m12, om12) and n > 20
