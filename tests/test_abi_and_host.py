"""CPU-side checks of the product boundary: the shared library loads and exports every symbol that
include/ygzorb.h declares, compute entry points fail loudly without a GPU (no CPU fallback), and the
host-side C++ classes compile and link."""
import ctypes
import os
import re
import subprocess

import pytest

from ygz import Y, PKG, ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "ygzorb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ygzorb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 30
    lib = Y.lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_structs_match_header_layout():
    assert ctypes.sizeof(Y.Params) == 20 and Y.KP_DTYPE.itemsize == 28
    assert ctypes.sizeof(Y.FrameFeatures) == 8 + 5 * 8 + 16 + 8 + 8        # ints, pointers, bounds, nlevels(+pad), pointer
    assert ctypes.sizeof(Y.MapPointProj) == 8 + 9 * 8
    assert ctypes.sizeof(Y.KeyFrameFeatures) == 16 + 9 * 8


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(Y.YgzError) as e:
        Y.Extractor()
    assert "CUDA" in str(e.value)
    with pytest.raises(Y.YgzError):
        Y.Matcher()
    import numpy as np
    with pytest.raises(Y.YgzError):          # the pyramid object of the direct-tracking entry points lives on the device too
        Y.Pyramid([np.zeros((48, 64), np.uint8)])
    assert Y.lib().ygzorb_version().startswith(b"ygzorb-b200")


def test_direct_tracking_entry_points_reject_null_handles():
    """Row f4 entry points: argument errors come back as codes with a message, never as a crash or a silent success."""
    L = Y.lib()
    i64, f7 = ctypes.c_int64(), (ctypes.c_float * 7)()
    L.ygzorb_sparse_align_run.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    assert L.ygzorb_sparse_align_run(None, 7, 1, f7, ctypes.byref(i64), None) < 0
    L.ygzorb_sparse_align_create.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    assert L.ygzorb_sparse_align_create(None, None) < 0
    L.ygzorb_pyramid_info.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    assert L.ygzorb_pyramid_info(None, None, None, None) < 0
    L.ygzorb_distinctive_descriptors.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int, ctypes.c_uint] + [ctypes.c_void_p] * 3
    assert L.ygzorb_distinctive_descriptors(None, None, None, 1, 0, None, None, None) < 0
    assert b"null" in L.ygzorb_last_error()


def test_direct_tracking_demo_builds():
    subprocess.run(["make", "-C", os.path.join(PKG, "host")], check=True, stdout=subprocess.DEVNULL)
    demo = os.path.join(PKG, "host", "host_demo3")
    res = subprocess.run([demo], capture_output=True, text=True)
    assert res.returncode == 2 and "usage" in res.stderr


def test_product_does_not_reference_the_oracle():
    for sub in ("csrc", "host"):
        d = os.path.join(PKG, sub)
        for fn in os.listdir(d):
            if fn.endswith((".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(d, fn)).read()
                assert "oracle/" not in src and "ygzo_" not in src, fn
    assert "oracle" not in open(os.path.join(PKG, "ygzorb_ctypes.py")).read()


def test_host_classes_compile_and_link():
    subprocess.run(["make", "-C", os.path.join(PKG, "host")], check=True, stdout=subprocess.DEVNULL)
    demo = os.path.join(PKG, "host", "host_demo")
    assert os.path.exists(demo)
    res = subprocess.run([demo], capture_output=True, text=True)
    assert res.returncode == 2 and "usage" in res.stderr
