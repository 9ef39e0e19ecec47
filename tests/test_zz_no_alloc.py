"""bs_stitch_device_static: the entry point for callers that may not allocate.

Some users of this library own all of their memory and cannot allocate while frames are flowing:
a capture or render loop with a fixed device-memory budget, a host application that forbids
cudaMalloc after start-up (it stalls the whole GPU), a plug-in running inside somebody else's
allocator.  For them the library keeps the staged kernel's scratch (block plan + per-warp work split)
in one fixed static device array and offers one call that allocates nothing.

What is checked here, through the C ABI, on the GPU:
  * the static entry point produces exactly the bytes the regular entry points produce
    (bs_create + bs_submit_device), on two calls in a row -- the second call finds the array full of
    the first call's plan and counters and must not be disturbed by them;
  * a different calibration right afterwards (other canvas, seams, ragged edges) is also exact, so
    nothing stale survives in the shared array;
  * a canvas that needs more blocks than the array holds is refused with a clear message, one block over the
    limit included, and a panorama that nearly fills the array (7680x2160: 32400 of 32768 blocks) is exact;
  * repeated calls leave the amount of free device memory unchanged (the property in the name), and flags the
    entry point does not implement are refused instead of being ignored.
This file sorts last on purpose: with `pytest -x` every other test has already run if it fails.
"""
import numpy as np
import pytest

import oracle
from conftest import pkg, workloads

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def regular_path(p, src_whs, canvas_wh, Hs, d_src, flags=0, prior=None):
    """The existing entry points: handle + device-resident submit."""
    d_out = torch.zeros((canvas_wh[1], canvas_wh[0], 3), dtype=torch.uint8, device="cuda")
    if prior is not None:
        d_out.copy_(torch.from_numpy(prior))
    with p.Stitcher(src_whs, canvas_wh, Hs, flags=flags) as st:
        st.submit_device([t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], d_out.data_ptr(), d_out.stride(0), 0)
        torch.cuda.synchronize()
    return d_out.cpu().numpy()


def static_path(p, src_whs, canvas_wh, Hs, d_src, flags=0, prior=None):
    d_out = torch.full((canvas_wh[1], canvas_wh[0], 3), 201, dtype=torch.uint8, device="cuda")
    if prior is not None:
        d_out.copy_(torch.from_numpy(prior))
    p.stitch_device_static(src_whs, [t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], canvas_wh,
                           d_out.data_ptr(), d_out.stride(0), Hs, flags=flags, stream=0)
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


def test_static_entry_matches_regular_entry_twice_in_a_row():
    p = pkg()
    w = workloads().get("cfg2")
    whs = [w.src_wh] * w.n_cams
    for frame in (0, 1):        # two calls in a row share the static array
        srcs = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, frame, oracle.NOISE) for i in range(w.n_cams)]
        d_src = [torch.from_numpy(s).cuda() for s in srcs]
        want = regular_path(p, whs, w.canvas_wh, w.H, d_src)
        got = static_path(p, whs, w.canvas_wh, w.H, d_src)
        assert np.array_equal(got, want), f"frame {frame}: {(got != want).any(axis=2).sum()} pixels differ"
        assert np.array_equal(got, oracle.stitch(srcs, w.H, w.canvas_wh))


def test_other_calibration_right_after_is_not_disturbed_by_stale_scratch():
    p = pkg()
    rng = np.random.default_rng(77)
    # first fill the array with a large plan ...
    w = workloads().get("cfg2")
    big = [torch.from_numpy(oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, 3, oracle.NOISE)).cuda() for i in range(2)]
    static_path(p, [w.src_wh] * 2, w.canvas_wh, w.H, big)
    # ... then run a small, ragged, three-camera canvas with seams, holes and last-row/column pixels
    whs = [(200, 120), (160, 130), (96, 64)]
    canvas = (531, 203)
    Hs = [np.array([[1.1, 0.03, 5.0], [-0.02, 1.2, 8.0], [2e-5, 1e-5, 1.0]]),
          np.array([[0.9, -0.05, 230.0], [0.04, 1.05, 20.0], [-1e-5, 2e-5, 1.0]]),
          np.array([[2.5, 0.0, 300.0], [0.0, 2.5, 30.0], [0.0, 0.0, 1.0]])]
    srcs = [rng.integers(0, 256, (h, wd, 3), dtype=np.uint8) for wd, h in whs]
    d_src = [torch.from_numpy(s).cuda() for s in srcs]
    for flags, prior in ((0, None), (p.FLAG_LEGACY_BORDER, None),
                         (p.FLAG_KEEP_CANVAS, rng.integers(0, 256, (canvas[1], canvas[0], 3), dtype=np.uint8))):
        want = regular_path(p, whs, canvas, Hs, d_src, flags, prior)
        got = static_path(p, whs, canvas, Hs, d_src, flags, prior)
        assert np.array_equal(got, want), f"flags={flags}"


def test_canvas_beyond_the_static_array_is_refused():
    p = pkg()
    w = workloads().get("cfg1")        # 6085 x 8353 needs 191 x 523 = 99893 blocks of 32x16, the array holds 32768
    d_src = [torch.zeros((8, 8, 3), dtype=torch.uint8, device="cuda")]
    d_out = torch.zeros((16, 16, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(p.StitchError) as e:
        p.stitch_device_static([(8, 8)], [d_src[0].data_ptr()], [d_src[0].stride(0)], w.canvas_wh, d_out.data_ptr(),
                               w.canvas_wh[0] * 3, [np.eye(3)])
    assert e.value.code == -4
    assert "99893 blocks" in str(e.value) and "32768" in str(e.value)


def test_near_limit_canvas_is_exact_and_nothing_is_allocated():
    """cfg3 needs 32400 of the array's 32768 blocks.  Ten calls in a row: same bytes as the regular entry points,
    and cudaMemGetInfo reports the same free memory before and after (calls 2..10; the first call on a device
    may make the driver load the kernels)."""
    p = pkg()
    w = workloads().get("cfg3")
    whs = [w.src_wh] * w.n_cams
    srcs = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, 0, oracle.NOISE) for i in range(w.n_cams)]
    d_src = [torch.from_numpy(s).cuda() for s in srcs]
    want = regular_path(p, whs, w.canvas_wh, w.H, d_src)
    d_out = torch.zeros((w.canvas_wh[1], w.canvas_wh[0], 3), dtype=torch.uint8, device="cuda")
    args = (whs, [t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], w.canvas_wh, d_out.data_ptr(), d_out.stride(0), w.H)
    p.stitch_device_static(*args)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    import time
    t0 = time.perf_counter()
    for _ in range(9):
        p.stitch_device_static(*args)
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0) / 9
    assert torch.cuda.mem_get_info()[0] == free0, "bs_stitch_device_static changed the amount of free device memory"
    assert np.array_equal(d_out.cpu().numpy(), want)
    print(f"bs_stitch_device_static at cfg3: {ms:.2f} ms per call (same calibration: the plan stays in the scratch)")


def test_plan_kept_between_calls_only_for_the_same_calibration():
    """A call with the previous call's calibration reuses the plan in the static array; anything that changes the
    plan (homography, flags, sizes) rebuilds it; calibrations A, B, A, A in turn, new frames and new buffers every
    time, an unaligned canvas in between (per-pixel kernel, scratch untouched): every result exact."""
    p = pkg()
    rng = np.random.default_rng(5)
    whs = [(320, 200), (320, 200)]
    canvas = (600, 230)
    HA = [np.array([[1.0, 0.01, 3.0], [-0.01, 1.0, 5.0], [1e-5, 0, 1.0]]), np.array([[1.02, 0.0, 270.0], [0.01, 0.98, 9.0], [0, 1e-5, 1.0]])]
    HB = [HA[0].copy(), HA[1].copy()]
    HB[1][0, 2] += 2.0 ** -20           # differs in one low bit of one entry: a different calibration
    seq = [(HA, 0), (HB, 0), (HA, 0), (HA, 0), (HA, p.FLAG_LEGACY_BORDER), (HA, 0)]
    for k, (Hs, flags) in enumerate(seq):
        srcs = [rng.integers(0, 256, (h, wd, 3), dtype=np.uint8) for wd, h in whs]
        d_src = [torch.from_numpy(s_).cuda() for s_ in srcs]
        want = oracle.stitch(srcs, Hs, canvas, border_rule=oracle.BORDER_LEGACY if flags else oracle.BORDER_CV4)
        got = static_path(p, whs, canvas, Hs, d_src, flags)
        assert np.array_equal(got, want), f"call {k}"
        if k == 2:                      # same calibration, canvas 1 byte off alignment: the per-pixel kernel, then staged again
            buf = torch.zeros(canvas[1] * canvas[0] * 3 + 16, dtype=torch.uint8, device="cuda")
            off = buf[1:1 + canvas[1] * canvas[0] * 3]
            p.stitch_device_static(whs, [t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], canvas, off.data_ptr(),
                                   canvas[0] * 3, Hs, flags=flags, stream=0)
            torch.cuda.synchronize()
            assert np.array_equal(off.cpu().numpy().reshape(canvas[1], canvas[0], 3), want)


def test_one_block_over_the_limit_and_unknown_flags_are_refused():
    p = pkg()
    d_src = [torch.zeros((8, 8, 3), dtype=torch.uint8, device="cuda")]
    d_out = torch.zeros((16, 16, 3), dtype=torch.uint8, device="cuda")
    # 99 x 331 = 32769 blocks of 32x16: one more than the array holds
    with pytest.raises(p.StitchError) as e:
        p.stitch_device_static([(8, 8)], [d_src[0].data_ptr()], [24], (99 * 32, 331 * 16), d_out.data_ptr(), 99 * 32 * 3, [np.eye(3)])
    assert e.value.code == -4 and "32769 blocks" in str(e.value)
    # 128 x 256 = 32768 blocks: accepted (a tiny source in a huge canvas: everything but one corner is empty)
    big = torch.full((256 * 16, 128 * 32, 3), 9, dtype=torch.uint8, device="cuda")
    src = torch.full((8, 8, 3), 200, dtype=torch.uint8, device="cuda")
    p.stitch_device_static([(8, 8)], [src.data_ptr()], [24], (128 * 32, 256 * 16), big.data_ptr(), big.stride(0), [np.eye(3)])
    torch.cuda.synchronize()
    got = big.cpu().numpy()
    assert (got[:7, :7] == 200).all() and got[8:, :].max() == 0 and got[:, 8:].max() == 0
    for bad in (p.FLAG_BORDER_CONSTANT, p.FLAG_EXACT_ONLY, 0x100):
        with pytest.raises(p.StitchError) as e:
            p.stitch_device_static([(8, 8)], [src.data_ptr()], [24], (64, 32), d_out.data_ptr(), 64 * 3, [np.eye(3)], flags=bad)
        assert e.value.code == -4 and "flags" in str(e.value)
