#!/usr/bin/env python
"""bench.py -- panorama frames/sec of the warp + composite hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload cfg3]

A "step" is one pass of the hot path over one batch of synthetic input: `--frames-per-step`
panorama frames (default 64), cycling through a ring of 8 distinct frame sets per camera, so the
inputs of consecutive launches never sit in L2 (ring of 8 x (sources + canvas) = 1.19 GB at cfg3,
L2 is 126 MB).

Printed JSON (one line, rank 0):
  value      frames/sec, whole job (all ranks), inputs and canvases resident in HBM;
  e2e        the same metric through bs_submit/bs_wait with pinned HOST buffers, H2D and D2H copies
             inside the timed region;
  roofline   algorithmic bytes per frame / kernel time against the measured HBM peak of
             MEASURED_PEAKS.json.  Two kernel times are printed: `kernel_us_pipelined` = timed region /
             frames with the launches back to back (programmatic dependent launch lets the head of frame
             n+1 overlap the tail of frame n; this is the rate a video stream sees, and what `frac` and
             `achieved` use), and `kernel_us_isolated` = one launch between two events after a device
             synchronise (best and median of 10);
  cpu_baseline  the reference's OpenCV path (cv2.warpPerspective, same call as
             /root/reference/tools/stitch/main.cpp:150-151) on this box's host cores;
  clocks     SM clock and throttle reasons sampled during the timed regions.

--impl reference times that OpenCV CPU path alone (a step = one panorama frame).
Multi-GPU: one process per GPU under torchrun, frames sharded frame-parallel, no data-path
collective; torch.distributed is used only for the barrier and the max-over-ranks time.
"""
import argparse
import hashlib
import importlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RING = 8
METRIC = "panorama_frames_per_sec"
UNIT = "frames/s"
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=["cfg2", "cfg3", "cfg5", "cfg1"])
    ap.add_argument("--content", default="noise", choices=["noise", "smooth"])
    ap.add_argument("--frames-per-step", type=int, default=64)
    ap.add_argument("--e2e-frames-per-step", type=int, default=8)
    ap.add_argument("--cpu-seconds", type=float, default=12.0,
                    help="bound on the cpu_baseline sample (wall seconds)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--flags", type=int, default=0, help="BS_FLAG_* for the handle (debug)")
    ap.add_argument("--lens", action="store_true",
                    help="every camera gets a lens model (cv::undistort before the warp, the video-mode loop "
                         "of tools/stitch/main.cpp:148-151); adds the `undistort` record")
    ap.add_argument("--no-host-ceiling", action="store_true")
    ap.add_argument("--no-compressed", action="store_true", help="skip the e2e_compressed record (JPEG frames in)")
    return ap.parse_args()


def common_config(w, content, lens=False):
    """`config` is the same dict in both arms (the driver compares them); arm-specific facts go under `run`."""
    return {"workload": workload_text(w), "content": content, "lens_models": bool(lens)}


def workload_text(w):
    return (f"{w.name}: {w.n_cams}x{w.src_wh[0]}x{w.src_wh[1]} BGR u8 -> {w.canvas_wh[0]}x{w.canvas_wh[1]} "
            f"panorama (BASELINE.json configs, SURVEY.md App. B.2 homographies)")


# ------------------------------------------------------------------------------------------
# CPU reference: the OpenCV call the reference makes, through cv2; the C port only if cv2 is
# not importable on this box.
# ------------------------------------------------------------------------------------------
def lens_model(w):
    """The lens model of tests/test_gpu_undistort.py (8 coefficients, as calibrate writes them)."""
    import numpy as np
    sw, sh = w.src_wh
    K = np.array([[0.9 * sw, 0, sw / 2 + 3.3], [0, 0.92 * sw, sh / 2 - 1.7], [0, 0, 1]])
    return K, np.array([-0.2, 0.05, 0.001, -0.002, 0.01, 0.0, 0.0, 0.0])


class CpuReference:
    def __init__(self, w, content, lens=False):
        wl = importlib.import_module("image-stitching_b200.workloads")
        import numpy as np
        self.np = np
        self.w = w
        self.lens = lens_model(w) if lens else None
        self.frames = [[wl.gen_frame_numpy(w.src_wh[0], w.src_wh[1], i, f, content)
                        for i in range(w.n_cams)] for f in range(2)]
        self.canvas = np.zeros((w.canvas_wh[1], w.canvas_wh[0], 3), np.uint8)
        try:
            import cv2
            self.cv2 = cv2
            cv2.setNumThreads(cv2.getNumberOfCPUs())
            self.cores = int(cv2.getNumThreads())
            self.kind = "reference"
            self.what = f"cv2 {cv2.__version__} warpPerspective(INTER_LINEAR, BORDER_TRANSPARENT)"
        except Exception:   # pragma: no cover - cv2 ships in this image
            import oracle
            self.cv2 = None
            self.oracle = oracle
            self.cores = oracle.default_threads()
            self.kind = "port"
            self.what = "oracle/warp_oracle.c (OpenMP)"

    def one_frame(self, f):
        w = self.w
        self.canvas[:] = 0
        srcs = self.frames[f % len(self.frames)]
        if self.cv2 is not None:
            for s, H in zip(srcs, w.H):
                if self.lens is not None:      # tools/stitch/main.cpp:148-149
                    s = self.cv2.undistort(s, self.lens[0], self.lens[1])
                self.cv2.warpPerspective(s, H, w.canvas_wh, dst=self.canvas, flags=self.cv2.INTER_LINEAR,
                                         borderMode=self.cv2.BORDER_TRANSPARENT)
        else:
            for s, H in zip(srcs, w.H):
                if self.lens is not None:
                    s = self.oracle.undistort(s, self.lens[0], self.lens[1])
                self.oracle.warp_perspective(s, H, self.canvas, nthreads=self.cores)
        return self.canvas

    def time_frames(self, n_warm, n, budget_s=None):
        for i in range(n_warm):
            self.one_frame(i)
        t0 = time.perf_counter()
        done = 0
        for i in range(n):
            self.one_frame(i)
            done += 1
            if budget_s is not None and time.perf_counter() - t0 > budget_s:
                break
        dt = time.perf_counter() - t0
        return done, dt


def _one_thread(self, n):
    """frames/sec of the same path on ONE host thread (a short sample: n frames)."""
    if self.cv2 is None:
        return None
    self.cv2.setNumThreads(1)
    try:
        done, dt = self.time_frames(1, n)
    finally:
        self.cv2.setNumThreads(self.cores)
    return done / dt


CpuReference.time_one_thread = _one_thread


def run_reference(args, rank):
    if rank != 0:
        return
    wl = importlib.import_module("image-stitching_b200.workloads")
    w = wl.get(args.workload)
    ref = CpuReference(w, args.content, args.lens)
    done, dt = ref.time_frames(args.warmup, args.steps)
    fps = done / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": done, "warmup": args.warmup, "ms_per_step": 1e3 * dt / done, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": common_config(w, args.content, args.lens),
        "run": {"step": "one panorama frame on the host cores (bounded sample of the workload)"},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": ref.cores, "kind": ref.kind,
                         "sample": f"{done} frames of {w.name}, {ref.what}"},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._h = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.005)

    def __enter__(self):
        if self._h is not None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------
def pin_near_gpu(index):
    """One process per GPU: run this rank on the host cores next to its GPU (NVML's ideal CPU set,
    intersected with what the container allows), so that its pinned frame buffers are allocated
    on that NUMA node and the copies do not cross sockets.  Returns the CPU list, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        ideal = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus = sorted(ideal & os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return cpus
    except Exception:
        pass
    return None


def run_b200(args, rank, local_rank, world):
    import numpy as np
    import torch
    pkg = importlib.import_module("image-stitching_b200")
    wl = importlib.import_module("image-stitching_b200.workloads")
    if not torch.cuda.is_available() or pkg.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback "
                         "(use --impl reference for the host baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = pin_near_gpu(local_rank) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w = wl.get(args.workload)
    n = w.n_cams
    sw, sh = w.src_wh
    W, H = w.canvas_wh
    B = args.frames_per_step

    # ring of RING frame sets per camera, distinct per rank (frame f -> GPU f mod world)
    torch.cuda.synchronize()
    def padded(h_, w_):
        """(h_, w_, 3) view of a buffer whose rows start on 128-byte boundaries (any width then takes
        the staged kernel; a tight pitch that is not a multiple of 16 falls back to the exact kernel)."""
        pitch = (3 * w_ + 127) // 128 * 128
        buf = torch.empty((h_, pitch), dtype=torch.uint8, device=dev)
        return buf[:, :3 * w_].unflatten(1, (w_, 3))

    d_frames = []
    for f in range(RING):
        fs = []
        for i in range(n):
            t = padded(sh, sw)
            t.copy_(wl.gen_frame_torch(sw, sh, i, rank + world * f, args.content, dev))
            fs.append(t)
        d_frames.append(fs)
    d_canvas = [padded(H, W) for _ in range(RING)]
    st = pkg.Stitcher([w.src_wh] * n, w.canvas_wh, w.H, flags=args.flags, device=local_rank)
    if args.lens:
        K_lens, d_lens = lens_model(w)
        for i in range(n):
            st.set_undistort(i, K_lens, d_lens)
    stream = torch.cuda.Stream(device=dev)       # the kernels and the timing events share this stream
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream
    f_ptrs = [[t.data_ptr() for t in fs] for fs in d_frames]
    f_pitch = [t.stride(0) for t in d_frames[0]]
    c_ptrs = [t.data_ptr() for t in d_canvas]
    c_pitch = d_canvas[0].stride(0)

    # the ring's frames were written long before the timed region: no kernel in the stream produces them
    settled = pkg.SUBMIT_SOURCES_SETTLED

    def device_step():
        for j in range(B):
            k = j % RING
            st.submit_device(f_ptrs[k], f_pitch, c_ptrs[k], c_pitch, sp, settled)

    # ---- parity spot check (untimed): frame 0 of rank 0 against the recorded OpenCV hash --------
    parity = None
    if args.lens:
        parity = "not checked here (lens models on: tests/test_gpu_undistort.py and test_gpu_api.py hold the parity checks)"
    elif rank == 0 and args.content == "noise":
        try:
            ka = json.load(open(os.path.join(ROOT, "tests", "golden", "known_answers.json")))
            want = [c for c in ka["cases"] if (c["config"], c["content"], c["frame"]) == (w.name, "noise", 0)]
            st.submit_device(f_ptrs[0], f_pitch, c_ptrs[0], c_pitch, sp, settled)
            stream.synchronize()
            got = hashlib.sha256(d_canvas[0].cpu().numpy().tobytes()).hexdigest()
            if want:
                parity = "sha256 == OpenCV golden" if got == want[0]["canvas_sha256"] else "MISMATCH vs OpenCV golden"
        except Exception as e:   # pragma: no cover
            parity = f"not checked ({e})"

    # ---- device-resident: `value` and the roofline ------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        device_step()
    launches0 = st.stats()["kernel_launches"]
    sampler = ClockSampler(physical_gpu_index(local_rank))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with sampler:
        ev0.record(stream)
        for _ in range(args.steps):
            device_step()
        ev1.record(stream)
        torch.cuda.synchronize()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = st.stats()["kernel_launches"] - launches0
    frames_dev = args.steps * B
    value = world * frames_dev / (dev_ms * 1e-3)
    kernel_us = 1e3 * dev_ms / frames_dev          # one frame = the launches of one submit

    # one frame alone: device idle before, two events around the launch(es) of a single submit
    iso = []
    for j in range(10):
        torch.cuda.synchronize()
        ev0.record(stream)
        st.submit_device(f_ptrs[j % RING], f_pitch, c_ptrs[j % RING], c_pitch, sp, settled)
        ev1.record(stream)
        torch.cuda.synchronize()
        iso.append(1e3 * ev0.elapsed_time(ev1))
    iso.sort()
    launches_per_frame = launches / frames_dev

    # row f1 alone: cv::undistort of one camera frame on device buffers (bs_undistort_device)
    undist = None
    if args.lens:
        d_tmp = padded(sh, sw)
        for _ in range(3):
            st.undistort_device(0, f_ptrs[0][0], f_pitch[0], d_tmp.data_ptr(), d_tmp.stride(0), sp, settled)
        torch.cuda.synchronize()
        reps = 8 * RING
        ev0.record(stream)
        for j in range(reps):
            st.undistort_device(j % n, f_ptrs[j % RING][j % n], f_pitch[j % n], d_tmp.data_ptr(), d_tmp.stride(0), sp, settled)
        ev1.record(stream)
        torch.cuda.synchronize()
        und_us = 1e3 * ev0.elapsed_time(ev1) / reps
        und_bytes = 2 * sw * sh * 3            # every output pixel written once, every source texel read about once
        undist = {"us_per_camera_frame": und_us, "algorithmic_bytes": und_bytes,
                  "achieved_gbs": und_bytes / (und_us * 1e-6) / 1e9, "cameras_per_panorama": n,
                  "what": "cv::undistort of one 8UC3 frame on device buffers, back to back over the frame ring"}
        # inside a submit the lens plans of all cameras run as ONE launch: its cost per camera frame is the
        # pipelined frame time with lens models minus the same handle's frame time without them
        with pkg.Stitcher([w.src_wh] * n, w.canvas_wh, w.H, flags=args.flags, device=local_rank) as st2:
            for j in range(RING):
                st2.submit_device(f_ptrs[j], f_pitch, c_ptrs[j], c_pitch, sp, settled)
            torch.cuda.synchronize()
            ev0.record(stream)
            for j in range(reps):
                st2.submit_device(f_ptrs[j % RING], f_pitch, c_ptrs[j % RING], c_pitch, sp, settled)
            ev1.record(stream)
            torch.cuda.synchronize()
            plain_us = 1e3 * ev0.elapsed_time(ev1) / reps
        set_us = (kernel_us - plain_us) / n
        undist.update({"us_per_camera_frame_in_submit": set_us, "frame_us_without_lens": plain_us,
                       "in_submit_how": "(pipelined frame time with lens models - without) / cameras; a submit runs all "
                                        "cameras' lens plans as one launch and only the blocks of each undistorted frame "
                                        "that the panorama plan can read (the part of the frame where that camera wins)"})

    # ---- what the host side can carry: the e2e copies alone, all ranks at once ----------------------
    def host_ceiling(h2d_bytes, d2h_bytes, reps=12):
        """Frames/s if a frame cost nothing but its copies: the same byte counts as one e2e frame, H2D and
        D2H concurrently on two side streams from/to pinned memory, every rank at the same time."""
        up_h = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
        up_d = torch.empty(h2d_bytes, dtype=torch.uint8, device=dev)
        dn_h = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
        dn_d = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
        s_up, s_dn = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

        def go(k):
            for _ in range(k):
                with torch.cuda.stream(s_up):
                    up_d.copy_(up_h, non_blocking=True)
                with torch.cuda.stream(s_dn):
                    dn_h.copy_(dn_d, non_blocking=True)
            s_up.synchronize()
            s_dn.synchronize()
        go(2)
        barrier()
        t0 = time.perf_counter()
        go(reps)
        t1 = time.perf_counter()
        barrier()
        return world * reps / max_over_ranks(t1 - t0)

    # ---- end to end: pinned host buffers, copies inside the timed region ------------------------
    e2e = None
    if not args.no_e2e:
        Be = args.e2e_frames_per_step
        depth = pkg.lib().bs_pipeline_depth(st._h)
        h_frames = []
        for f in range(RING):
            fs = []
            for i in range(n):
                a = pkg.pinned_empty((sh, sw, 3))
                a[...] = d_frames[f][i].cpu().numpy()
                fs.append(a)
            h_frames.append(fs)
        h_out = [pkg.pinned_empty((H, W, 3)) for _ in range(depth + 1)]

        last = {"j": 0}

        def e2e_steps(k_steps):
            tickets = []
            j = 0
            for _ in range(k_steps):
                for _ in range(Be):
                    if len(tickets) >= depth:
                        st.wait(tickets[len(tickets) - depth])
                    tickets.append(st.submit(h_frames[j % RING], h_out[j % len(h_out)]))
                    j += 1
            for t in tickets[-depth:]:
                st.wait(t)
            last["j"] = j - 1

        e2e_steps(max(2, min(args.warmup, 3)))
        barrier()
        with sampler:
            t0 = time.perf_counter()
            e2e_steps(args.steps)
            st.sync()
            t1 = time.perf_counter()
        barrier()
        e2e_s = max_over_ranks(t1 - t0)
        h2d_per_frame = int(st.stats()["h2d_bytes_last_submit"])     # the plan uploads only what it can read
        e2e = {"value": world * args.steps * Be / e2e_s, "unit": UNIT,
               "h2d_bytes_per_step": Be * h2d_per_frame, "d2h_bytes_per_step": Be * w.canvas_bytes,
               "h2d_bytes_per_step_full_frames": Be * w.src_bytes,
               "frames_per_step": Be, "pipeline_depth": depth, "ms_per_frame": 1e3 * e2e_s / (args.steps * Be),
               "host_buffers": "pinned (bs_host_alloc)"}
        if not args.no_host_ceiling:
            e2e["host_ceiling_fps"] = host_ceiling(h2d_per_frame, w.canvas_bytes)
            e2e["of_host_ceiling"] = e2e["value"] / e2e["host_ceiling_fps"]
        if rank == 0:   # the last panorama read back must equal the device-resident result
            j = last["j"]
            e2e["matches_device_path"] = bool(np.array_equal(h_out[j % len(h_out)],
                                                             d_canvas[j % RING].cpu().numpy()))

    # ---- end to end with COMPRESSED frames: JPEG bitstreams up, decoded on the device (bs_submit_jpeg) ----
    # Two records: plain JPEGs as cv2.imencode writes them (no restart markers) and the same frames written with a
    # restart interval of one MCU row.  Both are baseline 4:2:0, so both are decoded by the library's own GPU decoder
    # (stitch_jpegdec.cuh: entropy stage on the GPU, pixels equal cv2.imdecode's); `decoder` in the record says which
    # decoder ran (nvJPEG takes what that decoder does not: progressive, grey ...).
    e2e_jpeg = None
    e2e_jpeg_rst = None
    jpeg_seconds = []
    if not args.no_e2e and not args.no_compressed:
        try:
            import cv2
            # hash noise does not compress (its JPEG is as large as the raw frame), so these records use the smooth
            # content plus a little texture; the panorama comes back raw, as in `e2e`
            def textured(i, f):
                a = wl.gen_frame_numpy(sw, sh, i, rank + world * f, "smooth").astype(np.int16)
                a += (wl.gen_frame_numpy(sw, sh, i, f, "noise") >> 5).astype(np.int16) - 4
                return np.clip(a, 0, 255).astype(np.uint8)
            pics = [[textured(i, f) for i in range(n)] for f in range(3)]
            depth = pkg.lib().bs_pipeline_depth(st._h)
            hj_out = [pkg.pinned_empty((H, W, 3)) for _ in range(depth + 1)]

            def compressed_record(rst):
                enc_args = [cv2.IMWRITE_JPEG_QUALITY, 90] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, rst] if rst else [])
                jf = [[np.frombuffer(cv2.imencode(".jpg", pic, enc_args)[1].tobytes(), dtype=np.uint8).copy() for pic in fs] for fs in pics]
                jbytes = int(sum(b.size for b in jf[0]))

                def jpeg_frames(k):
                    tickets = []
                    for j in range(k):
                        if len(tickets) >= depth:
                            st.wait(tickets[len(tickets) - depth])
                        tickets.append(st.submit_jpeg(jf[j % len(jf)], hj_out[j % len(hj_out)]))
                    for t in tickets[-depth:]:
                        st.wait(t)
                jpeg_frames(4)
                # parity of the compressed path (rank 0, untimed): the panorama from the bitstreams must equal the panorama the
                # raw entry point makes from the pixels cv2.imdecode (the reference's imread) gives for the same bitstreams
                same = None
                if rank == 0:
                    chk = pkg.pinned_empty((H, W, 3))
                    st.wait(st.submit_jpeg(jf[0], chk))
                    dec = [np.ascontiguousarray(cv2.imdecode(b, cv2.IMREAD_COLOR)) for b in jf[0]]
                    same = bool(np.array_equal(chk, st.stitch(dec)))
                nj = max(8, min(48, args.steps * 2))
                # no collective in here: a rank that fails (and reports "unavailable") must not leave the others waiting;
                # the ranks start together within the spread of the last barrier, and every rank's own time is reduced
                # to the slowest one after the records (jpeg_seconds below)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                jpeg_frames(nj)
                st.sync()
                t1 = time.perf_counter()
                jpeg_seconds.append(t1 - t0)
                js = t1 - t0
                return {"value": world * nj / js, "unit": UNIT, "h2d_bytes_per_frame": jbytes, "d2h_bytes_per_frame": w.canvas_bytes,
                        "frames": nj, "restart_interval_mcus": rst, "equals_raw_path_on_cv2_imdecode_pixels": same,
                        "content": "smooth ramp + 3 bits of texture, JPEG quality 90, 4:2:0 (cv2.imencode)",
                        "decoder": f"{st.jpeg_backend()}; into the device frame ring (bs_submit_jpeg), panorama returned raw",
                        "raw_h2d_bytes_per_frame_for_comparison": h2d_per_frame if e2e else None}
            rst_env = int(os.environ.get("BS_BENCH_JPEG_RST", "0"))     # experiment: another restart interval for the second record
            e2e_jpeg = compressed_record(0)
            e2e_jpeg_rst = compressed_record(rst_env if rst_env else (sw + 15) // 16)
        except Exception as ex:   # pragma: no cover
            e2e_jpeg = e2e_jpeg or {"unavailable": f"{type(ex).__name__}: {ex}"}
        # (every rank gets here, whatever happened above) the slowest rank's time decides the whole-job rate
        for k, rec in enumerate((e2e_jpeg, e2e_jpeg_rst)):
            t = max_over_ranks(jpeg_seconds[k] if k < len(jpeg_seconds) else 0.0)
            if rec and "value" in rec and t > 0:
                rec["value"] = world * rec["frames"] / t

    stats = st.stats()
    stats_kernel = ["stitch_exact_kernel", "stitch_fast_kernel (general staged)", "stitch_fast2_kernel (compact boxes)"][stats["frame_kernel"]]
    st.close()

    # ---- CPU baseline: rank 0, N = 1 only ------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ref = CpuReference(w, args.content, args.lens)
        done, dt = ref.time_frames(2, 10 ** 6, budget_s=args.cpu_seconds)
        cpu = {"value": done / dt, "unit": UNIT, "cores": ref.cores, "kind": ref.kind,
               "sample": f"{done} frames of {w.name} in {dt:.1f} s, {ref.what}, all host threads"}
        one = ref.time_one_thread(3)
        if one is not None:
            cpu["one_thread"] = one           # also report the 1-thread figure

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"
    alg = w.algorithmic_bytes
    roofline = None
    if alg:
        achieved = alg / (kernel_us * 1e-6) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg, "kernel_us": kernel_us,
                    "kernel_us_pipelined": kernel_us, "kernel_us_isolated": iso[0],
                    "kernel_us_isolated_median": iso[len(iso) // 2],
                    "frac_uses": "kernel_us_pipelined (back-to-back launches, the rate a video stream sees)",
                    "frac_isolated": alg / (iso[0] * 1e-6) / 1e9 / peak,
                    "launches_per_frame": launches_per_frame,
                    "frac_of_nominal_8TBs": achieved / 8000.0}
        if args.lens:
            roofline["note"] = ("lens models on: a frame is one launch over all cameras' lens plans + the frame kernel; "
                                "algorithmic bytes are the frame kernel's only")
        tpath = os.path.join(ROOT, "profiles", "dram_traffic.json")
        if os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))
                roofline["traffic"] = tj.get(w.name)
                roofline["traffic_how"] = tj.get("how")
            except Exception:
                pass

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": common_config(w, args.content, args.lens),
        "run": {"frames_per_step": B, "ring_frames": RING,
                "l2": f"inputs larger than L2: ring of {RING} frame sets + canvases = "
                      f"{RING * (w.src_bytes + w.canvas_bytes) / 1e9:.2f} GB cycled per GPU",
                "parallelism": f"frame-parallel x{world}, no collective",
                "host_cpus_per_rank": (len(numa) if numa else None),
                "tiles_fast_exact_empty": [stats["n_tiles_fast"], stats["n_tiles_exact"], stats["n_tiles_empty"]],
                "kernel": stats_kernel},
        "undistort": undist,
        "e2e": e2e, "e2e_compressed": e2e_jpeg, "e2e_compressed_restart": e2e_jpeg_rst, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
        "clocks": sampler.summary(), "parity": parity,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
