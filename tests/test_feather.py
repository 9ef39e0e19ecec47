"""BS_FLAG_FEATHER: opt-in blending where cameras overlap.

PARITY UNPINNED: the reference only overwrites (tools/stitch/main.cpp:117,151), so no OpenCV call defines this
result.  The CPU restatement oracle_stitch_feather follows include/b200_stitch.h's definition; the CPU tests below
check that restatement against an independent numpy computation and against the properties the definition implies,
the GPU tests check the kernels against the restatement bit for bit.  None of this counts in the parity metric."""
import ctypes

import numpy as np
import pytest

import oracle
from conftest import pkg, workloads


def shift_h(dx, dy):
    return np.array([[1.0, 0, dx], [0, 1.0, dy], [0, 0, 1.0]])


def overlap_rig(sw, sh, n, step, tilt=1.5e-6):
    """n cameras side by side, each `step` canvas pixels right of the previous one, mild perspective."""
    Hs = []
    for i in range(n):
        Hs.append(np.array([[1.01 + 0.004 * i, 0.006 * (i % 2), i * step + 3.3],
                            [-0.005, 0.99 + 0.003 * i, 2.0 + 1.7 * i],
                            [tilt * (1 - 2 * (i % 2)), -tilt / 2, 1.0]]))
    return Hs


# ---------------------------------------------------------------------------- CPU: the restatement itself
def test_feather_integer_shift_against_numpy():
    """Pure translations by whole pixels: coordinates are exact, every camera's pixel is a plain copy, so the
    whole definition can be recomputed with numpy integer arithmetic."""
    rng = np.random.default_rng(5)
    sw, sh, W, Hc = 40, 24, 80, 32
    shifts = [(0, 0), (17, 3), (30, 5)]
    srcs = [rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8) for _ in shifts]
    got = oracle.stitch_feather(srcs, [shift_h(*s) for s in shifts], (W, Hc), nthreads=1)
    acc = np.zeros((Hc, W, 3), np.int64)
    wsum = np.zeros((Hc, W), np.int64)
    for s, (dx, dy) in zip(srcs, shifts):
        ys, xs = np.mgrid[0:sh, 0:sw]
        d = 1 + 32 * np.minimum(np.minimum(xs, sw - 1 - xs), np.minimum(ys, sh - 1 - ys))
        d = np.maximum(d, 1)
        acc[dy:dy + sh, dx:dx + sw] += d[..., None] * s.astype(np.int64)
        wsum[dy:dy + sh, dx:dx + sw] += d
    want = np.where(wsum[..., None] > 0, (acc + wsum[..., None] // 2) // np.maximum(wsum[..., None], 1), 0).astype(np.uint8)
    assert np.array_equal(got, want)


def test_feather_properties():
    sw, sh, W, Hc = 96, 64, 260, 80
    Hs = overlap_rig(sw, sh, 3, 70)
    srcs = [oracle.gen_frame(sw, sh, i, 0, oracle.NOISE) for i in range(3)]
    hard = oracle.stitch(srcs, Hs, (W, Hc))
    soft = oracle.stitch_feather(srcs, Hs, (W, Hc))
    # coverage count per pixel from single-camera stitches of an all-255 frame
    ones = [np.full((sh, sw, 3), 255, np.uint8) for _ in range(3)]
    cover = sum((oracle.stitch([ones[i]], [Hs[i]], (W, Hc)).any(axis=2)).astype(int) for i in range(3))
    assert (cover >= 2).sum() > 1000 and (cover == 1).sum() > 1000 and (cover == 0).sum() > 100
    assert np.array_equal(soft[cover <= 1], hard[cover <= 1]), "pixels one camera writes must be that camera's pixels"
    assert (soft[cover == 0] == 0).all()
    assert (soft[cover >= 2] != hard[cover >= 2]).any()
    # a weighted mean of the contributions lies between their minimum and maximum
    singles = [oracle.stitch([srcs[i]], [Hs[i]], (W, Hc)).astype(int) for i in range(3)]
    masks = [oracle.stitch([ones[i]], [Hs[i]], (W, Hc)).any(axis=2) for i in range(3)]
    lo = np.full((Hc, W, 3), 255)
    hi = np.zeros((Hc, W, 3), int)
    for s, m in zip(singles, masks):
        lo = np.where(m[..., None], np.minimum(lo, s), lo)
        hi = np.where(m[..., None], np.maximum(hi, s), hi)
    cov = cover >= 1
    assert (soft[cov] >= lo[cov]).all() and (soft[cov] <= hi[cov]).all()
    # cameras that agree leave nothing to blend
    flat = [np.full((sh, sw, 3), (10, 128, 250), np.uint8) for _ in range(3)]
    out = oracle.stitch_feather(flat, Hs, (W, Hc))
    assert (out[cov] == (10, 128, 250)).all()
    # one camera: the reference's result
    assert np.array_equal(oracle.stitch_feather(srcs[:1], Hs[:1], (W, Hc)), oracle.stitch(srcs[:1], Hs[:1], (W, Hc)))
    # camera order does not matter (last-camera-wins does)
    perm = [2, 0, 1]
    assert np.array_equal(oracle.stitch_feather([srcs[i] for i in perm], [Hs[i] for i in perm], (W, Hc)), soft)


def test_feather_seam_is_continuous():
    """What the flag is for: across the boundary of an overlap the hard cut jumps from one camera to the other,
    the feathered panorama does not."""
    sw, sh, W, Hc = 128, 64, 220, 64
    a = np.full((sh, sw, 3), 40, np.uint8)
    b = np.full((sh, sw, 3), 200, np.uint8)
    Hs = [shift_h(0, 0), shift_h(80, 0)]
    hard = oracle.stitch([a, b], Hs, (W, Hc)).astype(int)
    soft = oracle.stitch_feather([a, b], Hs, (W, Hc)).astype(int)
    row_h, row_s = hard[32, :207, 0], soft[32, :207, 0]
    assert np.abs(np.diff(row_h)).max() == 160
    assert np.abs(np.diff(row_s)).max() <= 12


# ---------------------------------------------------------------------------- GPU: kernels vs the restatement
def _gpu_case(sw, sh, n, step, W, Hc, flags=0, kind=None, border=None):
    p = pkg()
    Hs = overlap_rig(sw, sh, n, step)
    srcs = [oracle.gen_frame(sw, sh, i, 1, oracle.NOISE if kind is None else kind) for i in range(n)]
    want = oracle.stitch_feather(srcs, Hs, (W, Hc), border_rule=oracle.BORDER_CV4 if border is None else border)
    with p.Stitcher([(sw, sh)] * n, (W, Hc), Hs, flags=p.FLAG_FEATHER | flags) as st:
        got = st.stitch(srcs)
        stats = st.stats()
    return got, want, stats


@pytest.mark.gpu
@pytest.mark.parametrize("flags", ["plan", "exact_only", "legacy"])
def test_gpu_feather_matches_restatement(flags):
    p = pkg()
    f = {"plan": 0, "exact_only": p.FLAG_EXACT_ONLY, "legacy": p.FLAG_LEGACY_BORDER}[flags]
    got, want, stats = _gpu_case(640, 360, 4, 300, 1700, 420, flags=f,
                                 border=oracle.BORDER_LEGACY if flags == "legacy" else None)
    assert np.array_equal(got, want), f"{(got != want).any(axis=2).sum()} pixels differ"
    if flags == "plan":
        assert stats["n_tiles_fast"] > 0 and stats["n_tiles_exact"] > 0, "overlaps per pixel, the rest staged"


@pytest.mark.gpu
def test_gpu_feather_keeps_single_coverage_and_differs_in_overlaps():
    p = pkg()
    w = workloads().get("cfg3")
    sw, sh = w.src_wh
    srcs = [oracle.gen_frame(sw, sh, i, 0, oracle.NOISE) for i in range(4)]
    with p.Stitcher([w.src_wh] * 4, w.canvas_wh, w.H) as st:
        hard = st.stitch(srcs)
    with p.Stitcher([w.src_wh] * 4, w.canvas_wh, w.H, flags=p.FLAG_FEATHER) as st:
        soft = st.stitch(srcs)
    want = oracle.stitch_feather(srcs, w.H, w.canvas_wh)
    assert np.array_equal(soft, want)
    ones = [np.full((sh, sw, 3), 255, np.uint8)] * 4
    cover = sum(oracle.stitch([ones[i]], [w.H[i]], w.canvas_wh).any(axis=2).astype(np.int8) for i in range(4))
    assert np.array_equal(soft[cover <= 1], hard[cover <= 1])
    assert (soft[cover >= 2] != hard[cover >= 2]).any()


@pytest.mark.gpu
def test_gpu_feather_refusals():
    p = pkg()
    L = p.lib()
    H17 = [shift_h(4 * i, 0) for i in range(17)]
    with pytest.raises(p.StitchError) as e:
        p.Stitcher([(64, 48)] * 17, (256, 64), H17, flags=p.FLAG_FEATHER)
    assert e.value.code == -4 and "16" in str(e.value)
    with pytest.raises(p.StitchError) as e:
        p.Stitcher([(64, 48)] * 2, (256, 64), H17[:2], flags=p.FLAG_FEATHER | p.FLAG_BORDER_CONSTANT)
    assert e.value.code == -4
    with pytest.raises(p.StitchError) as e:
        p.Stitcher([(64, 48)] * 2, (256, 64), H17[:2], flags=0x80)
    assert e.value.code == -1 and "unknown flags" in str(e.value)
    src = np.zeros((48, 64, 3), np.uint8)
    dst = np.zeros((64, 256, 3), np.uint8)
    Hm = np.ascontiguousarray(H17[1], dtype=np.float64)
    rc = L.bs_warp_perspective(src.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)), 64, 48, src.strides[0],
                               dst.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)), 256, 64, dst.strides[0],
                               Hm.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), p.FLAG_FEATHER)
    assert rc == -4
