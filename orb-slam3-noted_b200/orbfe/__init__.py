"""orbfe — ctypes view of the B200-native ORB front end's C ABI (include/orbfe.h).

The product is the CUDA library `csrc/liborbfe.so`; this module is the thin harness tests and
bench.py use to reach it from Python, mirroring the reference's operator interface
(`ORBextractor::operator()`, include/ORBextractor.h:57-59 in the ORB-SLAM3 tree; the matchers of
include/ORBmatcher.h:39-91).  There is no CPU fallback: if the library is missing, or no CUDA
device is present, calls raise.
"""
from ._lib import (KEYPOINT_DTYPE, OrbfeError, lib, lib_path, strerror, version)  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
from .matcher import ORBmatcher, FeatureVectorCSR, measure_popc_peak  # noqa: F401
from .vocabulary import ORBVocabulary  # noqa: F401
from . import synth  # noqa: F401
from . import shard  # noqa: F401
