"""B200-native compositing engine for the stitch stage of AlexeySachkov/image-stitching.

This module is a thin ctypes view of the C ABI in include/b200_stitch.h (built in-tree as
libb200stitch.so from csrc/).  It exists for the tests and the benchmark; the product boundary is
the C ABI, and the C++ host (host/) is the drop-in for the reference's `stitch` tool.

There is no CPU fallback: if the shared library is missing the import fails, and without a CUDA
device every compute call raises StitchError.
"""
import ctypes
import os
import weakref

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_STITCH_LIB") or os.path.join(_HERE, "libb200stitch.so")   # env: development override

FLAG_LEGACY_BORDER = 0x1
FLAG_KEEP_CANVAS = 0x2
FLAG_EXACT_ONLY = 0x4
FLAG_BORDER_CONSTANT = 0x8
FLAG_FEATHER = 0x10   # opt-in, not the reference's result: weighted mean where cameras overlap (include/b200_stitch.h)
SUBMIT_SOURCES_SETTLED = 0x1   # bs_submit_device_ex: no kernel directly before the call writes the source frames

# names of the cv2 constants the reference passes (tools/stitch/main.cpp:116-117)
INTER_LINEAR = 1
BORDER_TRANSPARENT = 5
BORDER_CONSTANT = 0

EXPORTS = [
    "bs_device_count", "bs_version", "bs_last_error", "bs_create", "bs_destroy", "bs_submit",
    "bs_wait", "bs_pipeline_depth", "bs_stitch", "bs_submit_device", "bs_sync",
    "bs_warp_perspective", "bs_host_alloc", "bs_host_free", "bs_get_stats", "bs_stitch_device_static",
    "bs_set_undistort", "bs_undistort", "bs_submit_device_ex", "bs_undistort_device",
    "bs_pool_create", "bs_pool_destroy", "bs_pool_size", "bs_pool_depth", "bs_pool_handle", "bs_pool_last_error",
    "bs_pool_set_undistort", "bs_pool_submit", "bs_pool_wait", "bs_pool_sync", "bs_submit_jpeg", "bs_jpeg_backend",
]


class StitchError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200_stitch error {code}: {msg}")
        self.code = code


class JpegOut(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("capacity", ctypes.c_size_t), ("size", ctypes.c_size_t), ("quality", ctypes.c_int)]


class Stats(ctypes.Structure):
    _fields_ = [("n_cams", ctypes.c_int), ("canvas_w", ctypes.c_int), ("canvas_h", ctypes.c_int),
                ("tile_w", ctypes.c_int), ("tile_h", ctypes.c_int),
                ("n_tiles", ctypes.c_int64), ("n_tiles_fast", ctypes.c_int64),
                ("n_tiles_exact", ctypes.c_int64), ("n_tiles_empty", ctypes.c_int64),
                ("kernel_launches", ctypes.c_int64), ("frames_submitted", ctypes.c_int64),
                ("device_bytes", ctypes.c_uint64), ("scratch_bytes", ctypes.c_uint64),
                ("h2d_bytes_last_submit", ctypes.c_uint64), ("frame_kernel", ctypes.c_int)]


_lib = None


def build():
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    import subprocess
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-s"])
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}/csrc` "
                          "(there is no fallback implementation)")
    L = ctypes.CDLL(LIB_PATH)
    vp, i, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    ip, dp = ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_double)
    L.bs_device_count.restype = i
    L.bs_version.restype = ctypes.c_char_p
    L.bs_last_error.argtypes = [vp]
    L.bs_last_error.restype = ctypes.c_char_p
    L.bs_create.argtypes = [i, i, ip, ip, i, i, dp, ctypes.c_uint, ctypes.POINTER(vp)]
    L.bs_destroy.argtypes = [vp]
    L.bs_submit.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz,
                            ctypes.POINTER(ctypes.c_int64)]
    L.bs_wait.argtypes = [vp, ctypes.c_int64]
    L.bs_pipeline_depth.argtypes = [vp]
    L.bs_stitch.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz]
    L.bs_submit_device.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz, vp]
    L.bs_submit_device_ex.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz, vp, ctypes.c_uint]
    L.bs_sync.argtypes = [vp]
    L.bs_warp_perspective.argtypes = [vp, i, i, sz, vp, i, i, sz, dp, ctypes.c_uint]
    L.bs_host_alloc.argtypes = [sz]
    L.bs_host_alloc.restype = vp
    L.bs_host_free.argtypes = [vp]
    L.bs_get_stats.argtypes = [vp, ctypes.POINTER(Stats)]
    L.bs_set_undistort.argtypes = [vp, i, dp, dp, i]
    L.bs_undistort_device.argtypes = [vp, i, vp, sz, vp, sz, vp, ctypes.c_uint]
    L.bs_undistort.argtypes = [vp, i, i, sz, vp, sz, dp, dp, i]
    L.bs_stitch_device_static.argtypes = [i, ip, ip, ctypes.POINTER(vp), ctypes.POINTER(sz), i, i, vp, sz, dp,
                                          ctypes.c_uint, vp]
    L.bs_submit_jpeg.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz, ctypes.POINTER(JpegOut),
                                 ctypes.POINTER(ctypes.c_int64)]
    L.bs_jpeg_backend.argtypes = [vp]
    L.bs_jpeg_backend.restype = ctypes.c_char_p
    L.bs_pool_create.argtypes = [i, ip, i, ip, ip, i, i, dp, ctypes.c_uint, ctypes.POINTER(vp)]
    L.bs_pool_destroy.argtypes = [vp]
    L.bs_pool_size.argtypes = [vp]
    L.bs_pool_depth.argtypes = [vp]
    L.bs_pool_handle.argtypes = [vp, i]
    L.bs_pool_handle.restype = vp
    L.bs_pool_last_error.argtypes = [vp]
    L.bs_pool_last_error.restype = ctypes.c_char_p
    L.bs_pool_set_undistort.argtypes = [vp, i, dp, dp, i]
    L.bs_pool_submit.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, sz, ctypes.POINTER(ctypes.c_int64)]
    L.bs_pool_wait.argtypes = [vp, ctypes.c_int64]
    L.bs_pool_sync.argtypes = [vp]
    _lib = L
    return L


def device_count():
    return lib().bs_device_count()


def version():
    return lib().bs_version().decode()


def _check(rc, handle=None):
    if rc != 0:
        raise StitchError(rc, lib().bs_last_error(handle).decode())


def _check_img(a, name):
    if not (isinstance(a, np.ndarray) and a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 3
            and a.strides[2] == 1 and a.strides[1] == 3):
        raise ValueError(f"{name} must be an HxWx3 uint8 array with packed pixels")


def pinned_empty(shape):
    """uint8 ndarray backed by pinned host memory (bs_host_alloc); freed with the array."""
    n = int(np.prod(shape))
    p = lib().bs_host_alloc(n)
    if not p:
        raise StitchError(-5, lib().bs_last_error(None).decode())
    buf = (ctypes.c_uint8 * n).from_address(p)
    weakref.finalize(buf, lib().bs_host_free, p)   # freed when the last view of buf dies
    return np.frombuffer(buf, dtype=np.uint8).reshape(shape)


class Stitcher:
    """One calibration on one GPU: bs_create ... bs_destroy.

    src_whs: [(w, h)] per camera; canvas_wh: (W, H); Hs: n x 3 x 3 forward homographies
    (as in stitch.conf.xml, /root/reference/tools/stitch/main.cpp:60-73).
    """

    def __init__(self, src_whs, canvas_wh, Hs, flags=0, device=0):
        L = lib()
        n = len(src_whs)
        self.n_cams = n
        self.src_whs = [(int(w), int(h)) for w, h in src_whs]
        self.canvas_wh = (int(canvas_wh[0]), int(canvas_wh[1]))
        Hs = np.ascontiguousarray(np.asarray(Hs, dtype=np.float64).reshape(n, 9))
        sw = (ctypes.c_int * n)(*[w for w, _ in self.src_whs])
        sh = (ctypes.c_int * n)(*[h for _, h in self.src_whs])
        self._h = ctypes.c_void_p()
        _check(L.bs_create(device, n, sw, sh, self.canvas_wh[0], self.canvas_wh[1],
                           Hs.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), flags,
                           ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            lib().bs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # --- host buffers -------------------------------------------------------------------
    def _frame_args(self, frames):
        if len(frames) != self.n_cams:
            raise ValueError(f"expected {self.n_cams} frames, got {len(frames)}")
        ptrs = (ctypes.c_void_p * self.n_cams)()
        pitches = (ctypes.c_size_t * self.n_cams)()
        for k, (f, (w, h)) in enumerate(zip(frames, self.src_whs)):
            _check_img(f, f"frames[{k}]")
            if f.shape[0] != h or f.shape[1] != w:
                raise ValueError(f"frames[{k}] is {f.shape[1]}x{f.shape[0]}, expected {w}x{h}")
            ptrs[k] = f.ctypes.data
            pitches[k] = f.strides[0]
        return ptrs, pitches

    def new_canvas(self, pinned=False):
        shape = (self.canvas_wh[1], self.canvas_wh[0], 3)
        return pinned_empty(shape) if pinned else np.empty(shape, np.uint8)

    def submit(self, frames, out):
        ptrs, pitches = self._frame_args(frames)
        _check_img(out, "out")
        t = ctypes.c_int64(-1)
        _check(lib().bs_submit(self._h, ptrs, pitches, out.ctypes.data, out.strides[0],
                               ctypes.byref(t)), self._h)
        return t.value

    def wait(self, ticket):
        _check(lib().bs_wait(self._h, ticket), self._h)

    def stitch(self, frames, out=None):
        if out is None:
            out = self.new_canvas()
        self.wait(self.submit(frames, out))
        return out

    # --- compressed frames ----------------------------------------------------------------
    def submit_jpeg(self, jpegs, out=None, jpeg_capacity=0, quality=0):
        """bs_submit_jpeg: `jpegs` = one bytes-like JPEG bitstream per camera, decoded on the GPU.  The panorama
        comes back raw in `out` and/or, when jpeg_capacity > 0, as a JPEG (see jpeg_result).  Returns the ticket."""
        bufs = [j if isinstance(j, np.ndarray) and j.dtype == np.uint8 else np.frombuffer(bytes(j), dtype=np.uint8) for j in jpegs]
        ptrs = (ctypes.c_void_p * self.n_cams)(*[b.ctypes.data for b in bufs])
        sizes = (ctypes.c_size_t * self.n_cams)(*[b.size for b in bufs])
        jout = None
        if jpeg_capacity:
            store = np.empty(int(jpeg_capacity), np.uint8)
            jout = JpegOut(store.ctypes.data, store.size, 0, int(quality))
        if out is not None:
            _check_img(out, "out")
        t = ctypes.c_int64(-1)
        _check(lib().bs_submit_jpeg(self._h, ptrs, sizes, out.ctypes.data if out is not None else None,
                                    out.strides[0] if out is not None else 0, ctypes.byref(jout) if jout else None,
                                    ctypes.byref(t)), self._h)
        self._jpeg_pending = getattr(self, "_jpeg_pending", {})
        self._jpeg_pending[t.value] = (bufs, jout, store if jpeg_capacity else None)     # keep the buffers alive until wait()
        return t.value

    def jpeg_backend(self):
        return lib().bs_jpeg_backend(self._h).decode()

    def jpeg_result(self, ticket):
        """After wait(ticket): the panorama's JPEG bitstream (bytes) of a submit_jpeg with jpeg_capacity > 0."""
        bufs, jout, store = self._jpeg_pending.pop(ticket)
        return bytes(store[:jout.size]) if jout else None

    # --- device buffers -------------------------------------------------------------------
    def submit_device(self, frame_ptrs, frame_pitches, canvas_ptr, canvas_pitch, stream=0, submit_flags=None):
        """bs_submit_device (submit_flags None) or bs_submit_device_ex."""
        ptrs = (ctypes.c_void_p * self.n_cams)(*[int(p) for p in frame_ptrs])
        pitches = (ctypes.c_size_t * self.n_cams)(*[int(p) for p in frame_pitches])
        st = int(stream) if stream else None
        if submit_flags is None:
            _check(lib().bs_submit_device(self._h, ptrs, pitches, int(canvas_ptr), int(canvas_pitch), st), self._h)
        else:
            _check(lib().bs_submit_device_ex(self._h, ptrs, pitches, int(canvas_ptr), int(canvas_pitch), st,
                                             int(submit_flags)), self._h)

    def sync(self):
        _check(lib().bs_sync(self._h), self._h)

    def set_undistort(self, cam, K, dist):
        """Lens model of camera `cam` (cv::undistort before the warp, tools/stitch/main.cpp:148-149)."""
        K = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(9))
        dist = np.ascontiguousarray(np.asarray(dist, dtype=np.float64).reshape(-1))
        dptr = dist.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if dist.size else None
        _check(lib().bs_set_undistort(self._h, cam, K.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), dptr,
                                      int(dist.size)), self._h)

    def undistort_device(self, cam, src_ptr, src_pitch, dst_ptr, dst_pitch, stream=0, submit_flags=0):
        """bs_undistort_device: cv::undistort of one device-resident frame of camera `cam`."""
        _check(lib().bs_undistort_device(self._h, int(cam), int(src_ptr), int(src_pitch), int(dst_ptr), int(dst_pitch),
                                         int(stream) if stream else None, int(submit_flags)), self._h)

    def stats(self):
        s = Stats()
        _check(lib().bs_get_stats(self._h, ctypes.byref(s)), self._h)
        return {k: getattr(s, k) for k, _ in Stats._fields_}


class Pool:
    """bs_pool_*: one replica of the calibration per GPU, frame f -> device f mod G, results in frame order
    (the multi-GPU form of the video loop, /root/reference/tools/stitch/main.cpp:143-155)."""

    def __init__(self, devices, src_whs, canvas_wh, Hs, flags=0):
        L = lib()
        n = len(src_whs)
        self.n_cams = n
        self.src_whs = [(int(w), int(h)) for w, h in src_whs]
        self.canvas_wh = (int(canvas_wh[0]), int(canvas_wh[1]))
        Hs = np.ascontiguousarray(np.asarray(Hs, dtype=np.float64).reshape(n, 9))
        sw = (ctypes.c_int * n)(*[w for w, _ in self.src_whs])
        sh = (ctypes.c_int * n)(*[h for _, h in self.src_whs])
        devs = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
        self._p = ctypes.c_void_p()
        rc = L.bs_pool_create(len(devices), devs, n, sw, sh, self.canvas_wh[0], self.canvas_wh[1],
                              Hs.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), flags, ctypes.byref(self._p))
        if rc != 0:
            raise StitchError(rc, L.bs_last_error(None).decode())
        self.depth = L.bs_pool_depth(self._p)
        self.size = L.bs_pool_size(self._p)

    def _check(self, rc):
        if rc != 0:
            raise StitchError(rc, lib().bs_pool_last_error(self._p).decode())

    def set_undistort(self, cam, K, dist):
        K = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(9))
        dist = np.ascontiguousarray(np.asarray(dist, dtype=np.float64).reshape(-1))
        dptr = dist.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if dist.size else None
        self._check(lib().bs_pool_set_undistort(self._p, cam, K.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), dptr, int(dist.size)))

    def submit(self, frames, out):
        ptrs = (ctypes.c_void_p * self.n_cams)()
        pitches = (ctypes.c_size_t * self.n_cams)()
        for k, f in enumerate(frames):
            _check_img(f, f"frames[{k}]")
            ptrs[k] = f.ctypes.data
            pitches[k] = f.strides[0]
        _check_img(out, "out")
        idx = ctypes.c_int64(-1)
        self._check(lib().bs_pool_submit(self._p, ptrs, pitches, out.ctypes.data, out.strides[0], ctypes.byref(idx)))
        return idx.value

    def wait(self, frame):
        self._check(lib().bs_pool_wait(self._p, frame))

    def sync(self):
        self._check(lib().bs_pool_sync(self._p))

    def replica_stats(self, i):
        s = Stats()
        _check(lib().bs_get_stats(lib().bs_pool_handle(self._p, i), ctypes.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def close(self):
        if getattr(self, "_p", None):
            lib().bs_pool_destroy(self._p)
            self._p = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def stitch_device_static(src_whs, frame_ptrs, frame_pitches, canvas_wh, canvas_ptr, canvas_pitch, Hs, flags=0,
                         stream=0):
    """bs_stitch_device_static: one frame on device buffers, no allocation anywhere (the plan lives
    in the library's fixed static device array).  Pointers are raw device addresses."""
    n = len(src_whs)
    Hs = np.ascontiguousarray(np.asarray(Hs, dtype=np.float64).reshape(n, 9))
    sw = (ctypes.c_int * n)(*[int(w) for w, _ in src_whs])
    sh = (ctypes.c_int * n)(*[int(h) for _, h in src_whs])
    ptrs = (ctypes.c_void_p * n)(*[int(p) for p in frame_ptrs])
    pitches = (ctypes.c_size_t * n)(*[int(p) for p in frame_pitches])
    _check(lib().bs_stitch_device_static(n, sw, sh, ptrs, pitches, int(canvas_wh[0]), int(canvas_wh[1]),
                                         int(canvas_ptr), int(canvas_pitch),
                                         Hs.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), flags,
                                         int(stream) if stream else None))


def undistort(src, K, dist):
    """Mirror of cv::undistort(src, dst, cameraMatrix, distCoeffs) (tools/stitch/main.cpp:148-149)."""
    _check_img(src, "src")
    K = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(9))
    dist = np.ascontiguousarray(np.asarray(dist, dtype=np.float64).reshape(-1))
    dst = np.empty((src.shape[0], src.shape[1], 3), np.uint8)
    dptr = dist.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if dist.size else None
    _check(lib().bs_undistort(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], dst.ctypes.data,
                              dst.strides[0], K.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), dptr, int(dist.size)))
    return dst


def warpPerspective(src, dst, H, dsize, flags=INTER_LINEAR, borderMode=BORDER_TRANSPARENT,
                    legacy_border=False):
    """Mirror of the call the reference makes (tools/stitch/main.cpp:116-117):
    warpPerspective(src, dst, H, dsize, INTER_LINEAR, BORDER_TRANSPARENT), dst updated in place.
    borderMode=BORDER_CONSTANT (value 0; lib/Calibrate/utils.cpp:390) overwrites all of dst."""
    if flags != INTER_LINEAR or borderMode not in (BORDER_TRANSPARENT, BORDER_CONSTANT):
        raise StitchError(-4, "only INTER_LINEAR with BORDER_TRANSPARENT or BORDER_CONSTANT 0 (the reference's calls) is implemented")
    _check_img(src, "src")
    _check_img(dst, "dst")
    if (dst.shape[1], dst.shape[0]) != tuple(dsize):
        raise ValueError("dst does not have size dsize")
    H = np.ascontiguousarray(np.asarray(H, dtype=np.float64).reshape(9))
    _check(lib().bs_warp_perspective(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0],
                                     dst.ctypes.data, dst.shape[1], dst.shape[0], dst.strides[0],
                                     H.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                     (FLAG_LEGACY_BORDER if legacy_border else 0) |
                                     (FLAG_BORDER_CONSTANT if borderMode == BORDER_CONSTANT else 0)))
    return dst
