"""ctypes structures and prototypes of the matcher entry points of include/orbfe.h."""
import ctypes as C

_P, _I, _F, _PD = C.c_void_p, C.c_int, C.c_float, C.c_ssize_t


class FeatureVector(C.Structure):
    """orbfe_feature_vector"""
    _fields_ = [("nnodes", _I), ("node_ids", _P), ("offsets", _P), ("indices", _P)]


class BowSide(C.Structure):
    """orbfe_bow_side"""
    _fields_ = [("desc", _P), ("n", _I), ("fv", FeatureVector), ("mask", _P), ("kps", _P)]


class Epipolar(C.Structure):
    """orbfe_epipolar"""
    _fields_ = [("F12", _F * 9), ("ex", _F), ("ey", _F), ("sigma2_b", _P), ("scale_b", _P), ("nlevels", _I)]


class Kb8Epipolar(C.Structure):
    """orbfe_kb8_epipolar"""
    _fields_ = [("cam_a", _F * 16), ("cam_b", _F * 16), ("nleft_a", _I), ("nleft_b", _I), ("R12", _F * 36), ("t12", _F * 12),
                ("ex", _F), ("ey", _F), ("sigma2_a", _P), ("sigma2_b", _P), ("scale_b", _P), ("nlevels", _I)]


class GridFrameC(C.Structure):
    """orbfe_grid_frame"""
    _fields_ = [("kps", _P), ("desc", _P), ("n", _I), ("nleft", _I), ("blocked", _P), ("u_right", _P), ("left_to_right", _P),
                ("right_to_left", _P), ("min_x", _F), ("min_y", _F), ("grid_w_inv", _F), ("grid_h_inv", _F),
                ("inv_level_sigma2", _P), ("n_levels", _I)]


PROTO = {
    "orbfe_undistort_keypoints": (_I, [_P, _P, _I, _P, _P, _I, _P]),
    "orbfe_distinctive_descriptors": (_I, [_P, _P, _P, _I, _P]),
    "orbfe_search_by_projection": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "orbfe_search_for_initialization": (_I, [_P, _P, _P, _I, _P, _P, _I, _P]),
    "orbfe_vocabulary_create": (_I, [_I, _I, _I, _I, _I, _P, _P, _P, _P, _I, C.POINTER(_P)]),
    "orbfe_vocabulary_load_text": (_I, [C.c_char_p, _I, _I, C.POINTER(_P)]),
    "orbfe_vocabulary_destroy": (None, [_P]),
    "orbfe_vocabulary_words": (_I, [_P]),
    "orbfe_vocabulary_nodes": (_I, [_P]),
    "orbfe_vocabulary_stream": (_P, [_P]),
    "orbfe_vocabulary_synchronize": (_I, [_P]),
    "orbfe_vocabulary_launch_count": (C.c_longlong, [_P]),
    "orbfe_bow_transform": (_I, [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P]),
    "orbfe_bow_transform_batch_device": (_I, [_P, _P, _PD, _P, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P]),
    "orbfe_search_triangulation_kb8": (_I, [_P, _P, _P, _P, _I, _I, _P, _P]),
    "orbfe_matcher_create": (_I, [_F, _I, _I, C.POINTER(_P)]),
    "orbfe_matcher_destroy": (None, [_P]),
    "orbfe_matcher_stream": (_P, [_P]),
    "orbfe_matcher_synchronize": (_I, [_P]),
    "orbfe_matcher_launch_count": (C.c_longlong, [_P]),
    "orbfe_matcher_set_knn2_backend": (_I, [_P, _I]),
    "orbfe_hamming_knn2_noalloc": (_I, [_I, _P, _P, _I, _P, _I, _P, _P]),
    "orbfe_hamming_knn2": (_I, [_P, _P, _I, _P, _I, _P, _P]),
    "orbfe_hamming_knn2_batch_device": (_I, [_P, _P, _PD, _P, _I, _P, _PD, _P, _I, _I, _P, _P]),
    "orbfe_match_lapping": (_I, [_P, _P, _I, _P, _I, C.c_double, _P]),
    "orbfe_fisheye_stereo_match": (_I, [_P, _P, _P, _I, _I, _P, _P, _I, _I, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P]),
    "orbfe_stereo_match": (_I, [_P, _P, _I, _P, _I, _F, _F, _P, _P, _I]),
    "orbfe_stereo_match_batch_device": (_I, [_P, _P, _I, _F, _F, _P, _P]),
    "orbfe_search_bow": (_I, [_P, _I, C.POINTER(BowSide), C.POINTER(BowSide), _I, _I, _I, C.POINTER(_P), C.POINTER(_P), _P]),
    "orbfe_measure_popc_peak": (_I, [_I, C.POINTER(C.c_double)]),
    "orbfe_search_triangulation": (_I, [_P, C.POINTER(BowSide), _P, C.POINTER(BowSide), _P, C.POINTER(Epipolar), _I, _I, _P]),
}
