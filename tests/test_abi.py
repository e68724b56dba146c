"""CPU: the product library loads, exports every symbol include/orbfe.h declares, and refuses to
compute without a CUDA device (there is no CPU fallback to fall into)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import orbfe
from orbfe import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "orbfe.h")


def declared_functions():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    txt = re.sub(r"static inline[^{;]*\{.*?\n\}", "", txt, flags=re.S)   # host-side inline helpers are not exports
    return sorted(set(re.findall(r"\b(orbfe_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_all_exported():
    names = declared_functions()
    assert len(names) >= 20
    L = orbfe.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/orbfe.h but not exported by liborbfe.so"


def test_binding_covers_header():
    assert sorted(_lib.exported_symbols()) == declared_functions()


def test_no_oracle_symbols_or_deps_in_product():
    """The product must not link or embed the checker."""
    out = subprocess.run(["nm", "-D", "--defined-only", orbfe.lib_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "orbref" not in out
    ldd = subprocess.run(["ldd", orbfe.lib_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "orbref" not in ldd and "opencv" not in ldd


def test_keypoint_layout_is_cv_keypoint():
    assert orbfe.KEYPOINT_DTYPE.itemsize == 28
    assert [orbfe.KEYPOINT_DTYPE.fields[f][1] for f in ("x", "y", "size", "angle", "response", "octave", "class_id")] == \
        [0, 4, 8, 12, 16, 20, 24]


def test_strerror_and_version():
    assert orbfe.strerror(0) == "ok"
    assert "empty" in orbfe.strerror(-1)
    assert orbfe.version().startswith("orbfe ") and "sm_100a" in orbfe.version()


def test_bad_arguments_rejected_before_any_device_work():
    L = orbfe.lib()
    h = C.c_void_p()
    p = _lib.Params(0, 1.2, 8, 20, 7, 0, 752, 480, 1)            # nfeatures = 0
    assert L.orbfe_create(C.byref(p), C.byref(h)) == -2
    p = _lib.Params(1000, 1.0, 8, 20, 7, 0, 752, 480, 1)          # scale factor must be > 1
    assert L.orbfe_create(C.byref(p), C.byref(h)) == -2
    assert L.orbfe_create(None, C.byref(h)) == -2
    assert L.orbfe_max_keypoints(None) == -2
    assert L.orbfe_synchronize(None) == -2


def test_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(orbfe.OrbfeError) as e:
        orbfe.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.status == -6   # ORBFE_ERR_NO_DEVICE
    with pytest.raises(orbfe.OrbfeError) as e:
        orbfe.ORBmatcher()
    assert e.value.status == -6
