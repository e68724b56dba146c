"""Loads csrc/liborbfe.so and declares the prototypes of include/orbfe.h."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), "csrc", "liborbfe.so")

# Bit-compatible with cv::KeyPoint / orbfe_keypoint (28 bytes).
KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                           ("octave", "<i4"), ("class_id", "<i4")])
assert KEYPOINT_DTYPE.itemsize == 28


class OrbfeError(RuntimeError):
    def __init__(self, status, where=""):
        self.status = status
        msg = strerror(status)
        if status == -5:
            msg += ": " + lib().orbfe_last_cuda_error().decode()
        super().__init__(f"{where}: {msg} ({status})" if where else f"{msg} ({status})")


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int), ("ini_th_fast", C.c_int),
                ("min_th_fast", C.c_int), ("device", C.c_int), ("max_width", C.c_int), ("max_height", C.c_int),
                ("max_batch", C.c_int)]


_lib = None

_P = C.c_void_p
_I = C.c_int
_F = C.c_float
_PD = C.c_ssize_t


def lib_path():
    return _LIB_PATH


def lib():
    """The product library.  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ImportError(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a).  orbfe has no CPU fallback.")
    L = C.CDLL(_LIB_PATH)
    proto = {
        "orbfe_create": (_I, [C.POINTER(Params), C.POINTER(_P)]),
        "orbfe_destroy": (None, [_P]),
        "orbfe_get_tables": (_I, [_P, _P, _P, _P, _P, _P]),
        "orbfe_max_keypoints": (_I, [_P]),
        "orbfe_extract": (_I, [_P, _P, _I, _I, _PD, _I, _I, _P, _P, _I, C.POINTER(_I)]),
        "orbfe_extract_batch": (_I, [_P, _P, _I, _I, _I, _PD, _PD, _I, _I, _P, _P, _I, _P, _P]),
        "orbfe_extract_batch_device": (_I, [_P, _P, _I, _I, _I, _PD, _PD, _I, _I]),
        "orbfe_device_results": (_I, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_I)]),
        "orbfe_fetch_results": (_I, [_P, _I, _P, _P, _I, _P, _P]),
        "orbfe_get_pyramid_level": (_I, [_P, _I, _I, _P, _PD, C.POINTER(_I), C.POINTER(_I)]),
        "orbfe_get_candidates": (_I, [_P, _I, _I, _P, _I]),
        "orbfe_get_blurred_level": (_I, [_P, _I, _I, _P, _PD]),
        "orbfe_set_blur_backend": (_I, [_P, _I]),
        "orbfe_get_pyramid": (_I, [_P, _I, _P, _P]),
        "orbfe_get_level_size": (_I, [_P, _I, C.POINTER(_I), C.POINTER(_I)]),
        "orbfe_stream": (_P, [_P]),
        "orbfe_synchronize": (_I, [_P]),
        "orbfe_launch_count": (C.c_longlong, [_P]),
        "orbfe_graph_launch_count": (C.c_longlong, [_P]),
        "orbfe_stream_wait_extractor": (_I, [_P, _P]),
        "orbfe_extractor_wait_stream": (_I, [_P, _P]),
        "orbfe_set_pipeline": (_I, [_P, _I, _I, _I]),
        "orbfe_set_profiling": (_I, [_P, _I]),
        "orbfe_set_input_format": (_I, [_P, _I]),
        "orbfe_get_stage_times": (_I, [_P, _P, C.POINTER(C.c_longlong)]),
        "orbfe_stage_name": (C.c_char_p, [_I]),
        "orbfe_strerror": (C.c_char_p, [_I]),
        "orbfe_last_cuda_error": (C.c_char_p, []),
        "orbfe_version": (C.c_char_p, []),
    }
    from . import _matcher_proto
    proto.update(_matcher_proto.PROTO)
    for name, (res, args) in proto.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def exported_symbols():
    """Every entry point include/orbfe.h declares (used by the CPU-side ABI test)."""
    from . import _matcher_proto
    return ["orbfe_create", "orbfe_destroy", "orbfe_get_tables", "orbfe_max_keypoints", "orbfe_extract",
            "orbfe_extract_batch", "orbfe_extract_batch_device", "orbfe_device_results", "orbfe_fetch_results",
            "orbfe_get_pyramid_level", "orbfe_get_pyramid", "orbfe_get_level_size", "orbfe_get_candidates", "orbfe_get_blurred_level", "orbfe_set_blur_backend", "orbfe_stream", "orbfe_synchronize",
            "orbfe_launch_count", "orbfe_graph_launch_count", "orbfe_stream_wait_extractor", "orbfe_extractor_wait_stream",
            "orbfe_set_pipeline", "orbfe_set_profiling", "orbfe_set_input_format", "orbfe_get_stage_times", "orbfe_stage_name", "orbfe_strerror", "orbfe_last_cuda_error", "orbfe_version"] + list(_matcher_proto.PROTO)


def strerror(status):
    try:
        return lib().orbfe_strerror(int(status)).decode()
    except Exception:  # library missing: still give a readable message
        return f"orbfe status {status}"


def version():
    return lib().orbfe_version().decode()


def ptr(a):
    """Host pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_P)
