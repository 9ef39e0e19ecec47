"""C-ABI surface checks that need no GPU: the library loads, exports every symbol that
include/b200_stitch.h declares, validates arguments, and refuses to compute without a device
(there is no CPU fallback to fall into)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, pkg

HEADER = os.path.join(ROOT, "include", "b200_stitch.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bs_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_something():
    syms = declared_symbols()
    assert "bs_create" in syms and "bs_submit" in syms and "bs_warp_perspective" in syms


def test_library_exports_every_declared_symbol():
    L = pkg().lib()
    for s in declared_symbols():
        assert hasattr(L, s), f"{s} declared in include/b200_stitch.h but not exported"


def test_python_mirror_lists_the_same_symbols():
    assert sorted(pkg().EXPORTS) == declared_symbols()


def test_version_and_device_count():
    p = pkg()
    assert "sm_100a" in p.version()
    assert p.device_count() >= 0


def test_argument_validation_needs_no_device():
    p = pkg()
    L = p.lib()
    h = ctypes.c_void_p()
    one = (ctypes.c_int * 1)(16)
    Hm = (ctypes.c_double * 9)(1, 0, 0, 0, 1, 0, 0, 0, 1)
    assert L.bs_create(0, 0, one, one, 8, 8, Hm, 0, ctypes.byref(h)) == -1       # n_cams
    assert b"n_cams" in L.bs_last_error(None)
    assert L.bs_create(0, 1, one, one, 0, 8, Hm, 0, ctypes.byref(h)) == -1       # canvas
    big = (ctypes.c_int * 1)(40000)
    assert L.bs_create(0, 1, big, one, 8, 8, Hm, 0, ctypes.byref(h)) == -1       # int16 source limit
    assert L.bs_create(0, 1, one, one, 8, 8, None, 0, ctypes.byref(h)) == -1     # NULL H
    assert L.bs_submit(None, None, None, None, 0, None) == -1
    assert L.bs_wait(None, 0) == -1
    assert L.bs_destroy(None) == 0


def test_no_silent_cpu_fallback():
    p = pkg()
    if p.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(p.StitchError) as e:
        p.Stitcher([(16, 16)], (8, 8), [np.eye(3)])
    assert e.value.code == -3
    src = np.zeros((4, 4, 3), np.uint8)
    dst = np.zeros((4, 4, 3), np.uint8)
    with pytest.raises(p.StitchError) as e:
        p.warpPerspective(src, dst, np.eye(3), (4, 4))
    assert e.value.code == -3


def test_mirror_rejects_other_modes():
    p = pkg()
    src = np.zeros((4, 4, 3), np.uint8)
    with pytest.raises(p.StitchError):
        p.warpPerspective(src, src.copy(), np.eye(3), (4, 4), flags=0)  # INTER_NEAREST
