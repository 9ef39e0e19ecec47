"""GPU tests of the C-ABI entry points around the frame kernel that the round-1 review found unexercised:
device submits of a handle with lens models (caller's stream, two frames in flight), the ordering flag of
bs_submit_device_ex, bs_undistort_device, a rejected submit between two good ones, real image content, and both
staged kernels on the same input."""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from conftest import GOLDEN, ROOT, pkg, workloads

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def lens(sw, sh, k):
    Ks = [np.array([[0.9 * sw, 0, sw / 2 + 3.3], [0, 0.92 * sw, sh / 2 - 1.7], [0, 0, 1]]),
          np.array([[1.1 * sw, 0, sw / 2 - 7.0], [0, 1.1 * sw, sh / 2 + 2.5], [0, 0, 1]])]
    ds = [np.array([-0.2, 0.05, 0.001, -0.002, 0.01, 0, 0, 0]), np.array([0.12, -0.03, -0.001, 0.0007, 0.0])]
    return Ks[k % 2], ds[k % 2]


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_device_submit_with_lens_models_two_frames_in_flight():
    """tools/stitch/main.cpp:145-152 on device buffers: undistort + warp of frame A and of frame B queued back
    to back on the caller's stream, no synchronisation in between; both panoramas must be exact."""
    p = pkg()
    w = workloads().get("cfg2")
    sw, sh = w.src_wh
    frames = [[oracle.gen_frame(sw, sh, i, f, oracle.NOISE) for i in range(2)] for f in range(2)]
    d_frames = [[dev(x) for x in fs] for fs in frames]
    d_out = [torch.full((w.canvas_wh[1], w.canvas_wh[0], 3), 7, dtype=torch.uint8, device="cuda") for _ in range(2)]
    stream = torch.cuda.Stream()
    with p.Stitcher([w.src_wh] * 2, w.canvas_wh, w.H) as st:
        for i in range(2):
            st.set_undistort(i, *lens(sw, sh, i))
        free0 = torch.cuda.mem_get_info()[0]
        for f in range(2):
            st.submit_device([t.data_ptr() for t in d_frames[f]], [t.stride(0) for t in d_frames[f]], d_out[f].data_ptr(),
                             d_out[f].stride(0), stream.cuda_stream)
        stream.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "a device submit allocated memory (buffers belong to bs_set_undistort)"
    for f in range(2):
        und = [oracle.undistort(frames[f][i], *lens(sw, sh, i)) for i in range(2)]
        want = oracle.stitch(und, w.H, w.canvas_wh)
        got = d_out[f].cpu().numpy()
        assert np.array_equal(got, want), f"frame {f}: {(got != want).any(axis=2).sum()} pixels differ"


def test_submit_device_ex_flag_and_current_device():
    p = pkg()
    w = workloads().get("cfg2")
    srcs = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, 4, oracle.NOISE) for i in range(2)]
    d_src = [dev(s) for s in srcs]
    want = oracle.stitch(srcs, w.H, w.canvas_wh)
    with p.Stitcher([w.src_wh] * 2, w.canvas_wh, w.H) as st:
        for flags in (None, 0, p.SUBMIT_SOURCES_SETTLED):
            d_out = torch.zeros((w.canvas_wh[1], w.canvas_wh[0], 3), dtype=torch.uint8, device="cuda")
            for _ in range(3):       # back to back: with the flag, consecutive frames overlap
                st.submit_device([t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], d_out.data_ptr(), d_out.stride(0),
                                 0, flags)
            torch.cuda.synchronize()
            assert np.array_equal(d_out.cpu().numpy(), want), f"submit flags {flags}"
        with pytest.raises(p.StitchError) as e:
            st.submit_device([t.data_ptr() for t in d_src], [t.stride(0) for t in d_src], d_out.data_ptr(), d_out.stride(0), 0, 0x80)
        assert e.value.code == -1 and "submit flags" in str(e.value)
        assert st.stats()["frame_kernel"] == 2          # cfg2's boxes are compact: the compact-box kernel ran
    assert torch.cuda.current_device() == 0


def test_undistort_device_matches_host_entry():
    p = pkg()
    sw, sh = 1280, 720
    src = oracle.gen_frame(sw, sh, 1, 2, oracle.NOISE)
    K, d = lens(sw, sh, 0)
    want = oracle.undistort(src, K, d)
    d_src = dev(src)
    d_dst = torch.zeros_like(d_src)
    with p.Stitcher([(sw, sh)], (64, 32), [np.eye(3)]) as st:
        st.undistort_device(0, d_src.data_ptr(), d_src.stride(0), d_dst.data_ptr(), d_dst.stride(0))     # no model yet: a copy
        torch.cuda.synchronize()
        assert np.array_equal(d_dst.cpu().numpy(), src)
        st.set_undistort(0, K, d)
        st.undistort_device(0, d_src.data_ptr(), d_src.stride(0), d_dst.data_ptr(), d_dst.stride(0))
        torch.cuda.synchronize()
    assert np.array_equal(d_dst.cpu().numpy(), want)
    assert np.array_equal(p.undistort(src, K, d), want)


def test_rejected_submit_between_two_good_frames():
    """A submit that fails leaves no half-issued slot behind: the frames before and after it are exact and the
    tickets keep counting."""
    p = pkg()
    w = workloads().get("cfg2")
    frames = [[oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, f, oracle.NOISE) for i in range(2)] for f in range(3)]
    outs = [np.empty((w.canvas_wh[1], w.canvas_wh[0], 3), np.uint8) for _ in range(3)]
    with p.Stitcher([w.src_wh] * 2, w.canvas_wh, w.H) as st:
        t0 = st.submit(frames[0], outs[0])
        narrow = [f[:, : w.src_wh[0] // 2] for f in frames[1]]       # wrong width: rejected
        with pytest.raises(ValueError):
            st.submit(narrow, outs[1])
        import ctypes
        L = p.lib()
        ptrs = (ctypes.c_void_p * 2)(*[f.ctypes.data for f in frames[1]])
        bad = (ctypes.c_size_t * 2)(w.src_wh[0] * 3, w.src_wh[0] * 3 - 1)   # pitch below width * 3, straight at the C ABI
        t = ctypes.c_int64(-1)
        rc = L.bs_submit(st._h, ptrs, bad, outs[1].ctypes.data, outs[1].strides[0], ctypes.byref(t))
        assert rc == -1 and b"pitch[1]" in L.bs_last_error(st._h)
        t1 = st.submit(frames[1], outs[1])
        t2 = st.submit(frames[2], outs[2])
        assert (t0, t1, t2) == (0, 1, 2)
        for t in (t0, t1, t2):
            st.wait(t)
    for f in range(3):
        assert np.array_equal(outs[f], oracle.stitch(frames[f], w.H, w.canvas_wh)), f"frame {f}"


def test_real_image_content():
    """Crops of the reference's own photos (test/Inputs/1a.jpg, 1b.jpg) under cfg1-like homographies: the GPU
    result against the OpenCV hash recorded by tests/golden/make_real_crop.py and against the oracle."""
    p = pkg()
    z = np.load(os.path.join(GOLDEN, "real_crop.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    srcs = [np.ascontiguousarray(z["src0"]), np.ascontiguousarray(z["src1"])]
    whs = [(s.shape[1], s.shape[0]) for s in srcs]
    for flags in (0, p.FLAG_EXACT_ONLY):
        with p.Stitcher(whs, tuple(meta["canvas_wh"]), z["H"], flags=flags) as st:
            got = st.stitch(srcs)
        assert hashlib.sha256(got.tobytes()).hexdigest() == meta["canvas_sha256"], f"flags={flags}"
    assert np.array_equal(got, oracle.stitch(srcs, z["H"], tuple(meta["canvas_wh"])))
    want_legacy = oracle.stitch(srcs, z["H"], tuple(meta["canvas_wh"]), oracle.BORDER_LEGACY)
    with p.Stitcher(whs, tuple(meta["canvas_wh"]), z["H"], flags=p.FLAG_LEGACY_BORDER) as st:
        assert np.array_equal(st.stitch(srcs), want_legacy)


def test_general_staged_kernel_still_exact_on_compact_plans():
    """BS_DEBUG_NO_COMPACT routes a compact plan through the general staged kernel (the one strong perspective
    takes): both kernels must give the recorded OpenCV canvas.  Runs in a child process (the switch is read once)."""
    code = ("import sys, hashlib, json, importlib; sys.path.insert(0, %r); sys.path.insert(0, %r + '/tests')\n"
            "import oracle\n"
            "p = importlib.import_module('image-stitching_b200'); wl = importlib.import_module('image-stitching_b200.workloads')\n"
            "w = wl.get('cfg2'); srcs = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, 0, oracle.NOISE) for i in range(2)]\n"
            "st = p.Stitcher([w.src_wh] * 2, w.canvas_wh, w.H); got = st.stitch(srcs); k = st.stats()['frame_kernel']; st.close()\n"
            "print(json.dumps({'kernel': k, 'sha': hashlib.sha256(got.tobytes()).hexdigest()}))\n") % (ROOT, ROOT)
    ka = json.load(open(os.path.join(GOLDEN, "known_answers.json")))
    want = [c for c in ka["cases"] if (c["config"], c["content"], c["frame"]) == ("cfg2", "noise", 0)][0]["canvas_sha256"]
    for env_extra, kernel in (({}, 2), ({"BS_DEBUG_NO_COMPACT": "1"}, 1)):
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ, **env_extra))
        assert out.returncode == 0, out.stderr[-1500:]
        res = json.loads(out.stdout.strip().splitlines()[-1])
        assert res["kernel"] == kernel and res["sha"] == want, res


@pytest.mark.parametrize("devices", [[0, 0, 0], "all"])
def test_pool_returns_frames_in_order(devices):
    """bs_pool_*: frame f -> replica f mod G, panoramas in frame order.  [0, 0, 0] runs three replicas on one GPU
    (the dispatch logic without needing a multi-GPU box); "all" uses every visible GPU and needs at least two."""
    p = pkg()
    if devices == "all":
        n = p.device_count()
        if n < 2:
            pytest.skip("needs at least 2 GPUs")
        devices = list(range(n))
    w = workloads().get("cfg2")
    n_frames = 4 * len(devices) + 3
    frames = [[oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, f, oracle.NOISE) for i in range(2)] for f in range(n_frames)]
    with p.Pool(devices, [w.src_wh] * 2, w.canvas_wh, w.H) as pool:
        assert pool.size == len(devices) and pool.depth == 3 * len(devices)
        outs = [np.empty((w.canvas_wh[1], w.canvas_wh[0], 3), np.uint8) for _ in range(pool.depth)]
        got_sha, waited = [], 0
        for f in range(n_frames):
            if f - waited >= pool.depth:          # the buffer of frame f - depth is about to be reused
                pool.wait(waited)
                got_sha.append(hashlib.sha256(outs[waited % pool.depth].tobytes()).hexdigest())
                waited += 1
            assert pool.submit(frames[f], outs[f % pool.depth]) == f
        while waited < n_frames:
            pool.wait(waited)
            got_sha.append(hashlib.sha256(outs[waited % pool.depth].tobytes()).hexdigest())
            waited += 1
        pool.sync()
        per_replica = [pool.replica_stats(i)["frames_submitted"] for i in range(pool.size)]
    assert sum(per_replica) == n_frames and max(per_replica) - min(per_replica) <= 1
    want = [hashlib.sha256(oracle.stitch(fr, w.H, w.canvas_wh).tobytes()).hexdigest() for fr in frames]
    assert got_sha == want


def test_pool_reports_a_failed_frame():
    p = pkg()
    w = workloads().get("cfg2")
    fr = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, 0, oracle.NOISE) for i in range(2)]
    out = np.empty((w.canvas_wh[1], w.canvas_wh[0], 3), np.uint8)
    with p.Pool([0, 0], [w.src_wh] * 2, w.canvas_wh, w.H) as pool:
        f0 = pool.submit(fr, out)
        pool.wait(f0)
        import ctypes
        L = p.lib()
        ptrs = (ctypes.c_void_p * 2)(*[f.ctypes.data for f in fr])
        bad = (ctypes.c_size_t * 2)(10, 10)              # pitch below the frame width: the replica's bs_submit fails
        idx = ctypes.c_int64(-1)
        assert L.bs_pool_submit(pool._p, ptrs, bad, out.ctypes.data, out.strides[0], ctypes.byref(idx)) == 0
        with pytest.raises(p.StitchError) as e:
            pool.wait(idx.value)
        assert "frame 1 on device 0" in str(e.value) and "pitch" in str(e.value)


def test_lens_set_mixed_cameras():
    """All cameras' lens plans run as one launch: cameras of different sizes and models, one camera without a model
    in between, a model replaced and one removed afterwards (the set is rebuilt each time)."""
    p = pkg()
    whs = [(640, 360), (512, 300), (704, 400)]
    Hs = [np.array([[1.02, 0.004, 5.0 + 420 * i], [-0.003, 0.99, 3.0 + 2 * i], [1e-6 * (i - 1), -5e-7, 1.0]]) for i in range(3)]
    W, Hc = 1400, 440
    frames = [oracle.gen_frame(w_, h_, i, 3, oracle.NOISE) for i, (w_, h_) in enumerate(whs)]
    models = {0: lens(*whs[0], 0), 2: lens(*whs[2], 1)}

    def want():
        und = [oracle.undistort(frames[i], *models[i]) if i in models else frames[i] for i in range(3)]
        return oracle.stitch(und, Hs, (W, Hc))

    with p.Stitcher(whs, (W, Hc), Hs) as st:
        for i, m in models.items():
            st.set_undistort(i, *m)
        got = st.stitch(frames)
        assert np.array_equal(got, want())
        launches = st.stats()["kernel_launches"]
        st.stitch(frames)
        assert st.stats()["kernel_launches"] - launches == 2, "one lens launch for both cameras + the frame kernel"
        models[1] = lens(*whs[1], 0)             # a third member
        st.set_undistort(1, *models[1])
        assert np.array_equal(st.stitch(frames), want())
        models[0] = lens(*whs[0], 1)             # replaced
        st.set_undistort(0, *models[0])
        assert np.array_equal(st.stitch(frames), want())
        del models[2]                            # removed (zero model = copy)
        st.set_undistort(2, np.eye(3), np.zeros(5))
        assert np.array_equal(st.stitch(frames), want())
        d_in = [dev(f) for f in frames]          # device submit through the same set
        d_out = torch.zeros((Hc, W, 3), dtype=torch.uint8, device="cuda")
        st.submit_device([t.data_ptr() for t in d_in], [t.stride(0) for t in d_in], d_out.data_ptr(), d_out.stride(0), 0)
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy(), want())


def test_lens_launch_trimmed_to_what_the_panorama_reads():
    """cfg3 (no per-pixel blocks in the panorama plan) with a lens model on every camera: the lens launch covers only
    the blocks of each undistorted frame inside the rectangle where that camera wins, and bs_submit uploads only the
    rectangle those blocks read -- the panorama must still equal the full-frame pipeline (undistort everything, then
    warp) bit for bit, from host buffers and from device buffers."""
    p = pkg()
    w = workloads().get("cfg3")
    sw, sh = w.src_wh
    n = w.n_cams
    frames = [oracle.gen_frame(sw, sh, i, 2, oracle.NOISE) for i in range(n)]
    und = [oracle.undistort(frames[i], *lens(sw, sh, i)) for i in range(n)]
    want = oracle.stitch(und, w.H, w.canvas_wh)
    with p.Stitcher([w.src_wh] * n, w.canvas_wh, w.H) as st:
        assert st.stats()["n_tiles_exact"] == 0
        for i in range(n):
            st.set_undistort(i, *lens(sw, sh, i))
        got = st.stitch(frames)
        assert np.array_equal(got, want), f"host path: {(got != want).any(axis=2).sum()} pixels differ"
        up = st.stats()["h2d_bytes_last_submit"]
        assert up < 0.9 * n * sw * sh * 3, f"uploaded {up} bytes: the source rectangles were not trimmed"
        d_in = [dev(f) for f in frames]
        d_out = torch.zeros((w.canvas_wh[1], w.canvas_wh[0], 3), dtype=torch.uint8, device="cuda")
        for _ in range(2):          # twice: the second launch finds the undistorted buffers holding the first one's pixels
            st.submit_device([t.data_ptr() for t in d_in], [t.stride(0) for t in d_in], d_out.data_ptr(), d_out.stride(0), 0)
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy(), want)
