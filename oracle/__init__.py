"""ctypes view of oracle/liboracle.so -- the CPU checker.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  See warp_oracle.h for what it restates and how it is pinned.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

BORDER_CV4 = 0
BORDER_LEGACY = 1
BORDER_CONSTANT = 2     # warp_perspective only: cv2.BORDER_CONSTANT, value 0, every pixel written
NOISE = 0
SMOOTH = 1


def build(force=False):
    src = os.path.join(_HERE, "warp_oracle.c")
    hdr = os.path.join(_HERE, "warp_oracle.h")
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(p) > os.path.getmtime(_LIB_PATH) for p in (src, hdr))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle.so"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        u8p = ctypes.POINTER(ctypes.c_uint8)
        dp = ctypes.POINTER(ctypes.c_double)
        L.oracle_invert3x3.argtypes = [dp, dp]
        L.oracle_invert3x3.restype = ctypes.c_int
        L.oracle_warp_coord.argtypes = [dp, ctypes.c_int, ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
        L.oracle_warp_perspective.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                              u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                              dp, ctypes.c_int, ctypes.c_int]
        L.oracle_warp_perspective.restype = None
        L.oracle_stitch_feather.argtypes = [ctypes.c_int, ctypes.POINTER(u8p), ctypes.POINTER(ctypes.c_int),
                                            ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_size_t), u8p,
                                            ctypes.c_int, ctypes.c_int, ctypes.c_size_t, dp, ctypes.c_int, ctypes.c_int]
        L.oracle_stitch_feather.restype = None
        L.oracle_undistort.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p,
                                       ctypes.c_size_t, dp, dp, ctypes.c_int, ctypes.c_int]
        L.oracle_undistort.restype = None
        L.oracle_gen_frame.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                       ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.oracle_gen_frame.restype = None
        L.oracle_count_texels.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int),
                                          ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_int,
                                          dp, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64),
                                          ctypes.POINTER(ctypes.c_uint64)]
        L.oracle_count_texels.restype = ctypes.c_uint64
        _lib = L
    return _lib


def _u8(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))


def _d(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def default_threads():
    return os.cpu_count() or 1


def invert3x3(H):
    H = np.ascontiguousarray(H, dtype=np.float64).reshape(9)
    M = np.zeros(9, dtype=np.float64)
    ok = lib().oracle_invert3x3(_d(H), _d(M))
    return M.reshape(3, 3), bool(ok)


def warp_perspective(src, H, canvas, border_rule=BORDER_CV4, nthreads=None):
    """In-place cv::warpPerspective(src, canvas, H, canvas.size, INTER_LINEAR, BORDER_TRANSPARENT)."""
    assert src.dtype == np.uint8 and src.ndim == 3 and src.shape[2] == 3
    assert canvas.dtype == np.uint8 and canvas.ndim == 3 and canvas.shape[2] == 3
    assert src.strides[2] == 1 and src.strides[1] == 3 and canvas.strides[2] == 1 and canvas.strides[1] == 3
    H = np.ascontiguousarray(H, dtype=np.float64).reshape(9)
    lib().oracle_warp_perspective(_u8(src), src.shape[1], src.shape[0], src.strides[0],
                                  _u8(canvas), canvas.shape[1], canvas.shape[0], canvas.strides[0],
                                  _d(H), border_rule, nthreads or default_threads())
    return canvas


def stitch(srcs, Hs, canvas_wh, border_rule=BORDER_CV4, nthreads=None):
    """Zeroed canvas, then cameras in index order (tools/stitch/main.cpp:144-152)."""
    W, Hc = canvas_wh
    canvas = np.zeros((Hc, W, 3), dtype=np.uint8)
    for s, H in zip(srcs, Hs):
        warp_perspective(s, H, canvas, border_rule, nthreads)
    return canvas


def stitch_feather(srcs, Hs, canvas_wh, border_rule=BORDER_CV4, nthreads=None):
    """BS_FLAG_FEATHER restated (PARITY UNPINNED: the reference does not blend; include/b200_stitch.h defines it)."""
    W, Hc = canvas_wh
    n = len(srcs)
    assert all(s.dtype == np.uint8 and s.ndim == 3 and s.shape[2] == 3 and s.strides[1] == 3 and s.strides[2] == 1 for s in srcs)
    canvas = np.zeros((Hc, W, 3), dtype=np.uint8)
    u8p = ctypes.POINTER(ctypes.c_uint8)
    ptrs = (u8p * n)(*[_u8(s) for s in srcs])
    sw = (ctypes.c_int * n)(*[s.shape[1] for s in srcs])
    sh = (ctypes.c_int * n)(*[s.shape[0] for s in srcs])
    sp = (ctypes.c_size_t * n)(*[s.strides[0] for s in srcs])
    Hflat = np.ascontiguousarray(np.asarray(Hs, dtype=np.float64).reshape(n * 9))
    lib().oracle_stitch_feather(n, ptrs, sw, sh, sp, _u8(canvas), W, Hc, canvas.strides[0], _d(Hflat), border_rule,
                                nthreads or default_threads())
    return canvas


def undistort(src, K, dist, nthreads=None):
    assert src.dtype == np.uint8 and src.ndim == 3 and src.shape[2] == 3 and src.flags.c_contiguous
    K = np.ascontiguousarray(K, dtype=np.float64).reshape(9)
    dist = np.ascontiguousarray(dist, dtype=np.float64).reshape(-1)
    out = np.empty_like(src)
    lib().oracle_undistort(_u8(src), src.shape[1], src.shape[0], src.strides[0], _u8(out),
                           out.strides[0], _d(K), _d(dist) if dist.size else None, int(dist.size),
                           nthreads or default_threads())
    return out


def gen_frame(w, h, cam, frame, kind=NOISE):
    out = np.empty((h, w, 3), dtype=np.uint8)
    lib().oracle_gen_frame(_u8(out), w, h, out.strides[0], cam, frame, kind)
    return out


def count_texels(src_whs, canvas_wh, Hs, border_rule=BORDER_CV4):
    n = len(src_whs)
    sw = (ctypes.c_int * n)(*[int(w) for w, _ in src_whs])
    sh = (ctypes.c_int * n)(*[int(h) for _, h in src_whs])
    Hs = np.ascontiguousarray(np.asarray(Hs, dtype=np.float64).reshape(n, 9))
    tex = (ctypes.c_uint64 * n)()
    wins = (ctypes.c_uint64 * n)()
    total = lib().oracle_count_texels(n, sw, sh, canvas_wh[0], canvas_wh[1], _d(Hs), border_rule,
                                      tex, wins)
    return int(total), [int(t) for t in tex], [int(v) for v in wins]
