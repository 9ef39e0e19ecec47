#!/usr/bin/env python
"""End-to-end frames/s of the in-product pool (bs_pool_*: ONE process, one host thread per GPU) at cfg3, from pinned
host buffers -- the counterpart of bench.py's one-process-per-GPU e2e number.  usage: python tools/pool_bench.py [n_gpus]"""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
p = importlib.import_module("image-stitching_b200"); wl = importlib.import_module("image-stitching_b200.workloads")
n_gpus = int(sys.argv[1]) if len(sys.argv) > 1 else p.device_count()
w = wl.get("cfg3"); sw, sh = w.src_wh; W, H = w.canvas_wh
frames = []
for f in range(4):
    fs = []
    for i in range(w.n_cams):
        a = p.pinned_empty((sh, sw, 3)); a[...] = wl.gen_frame_numpy(sw, sh, i, f, "noise"); fs.append(a)
    frames.append(fs)
pool = p.Pool(list(range(n_gpus)), [w.src_wh] * w.n_cams, w.canvas_wh, w.H)
outs = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory().numpy() for _ in range(pool.depth)]
def run(n):
    waited = 0
    for f in range(n):
        if f - waited >= pool.depth:
            pool.wait(base + waited); waited += 1
        pool.submit(frames[f % 4], outs[f % pool.depth])
    while waited < n:
        pool.wait(base + waited); waited += 1
base = 0; run(4 * n_gpus); base += 4 * n_gpus
n = 60 * n_gpus
t0 = time.perf_counter(); run(n); pool.sync(); t1 = time.perf_counter()
print(json.dumps({"what": "bs_pool e2e at cfg3, one process", "n_gpus": n_gpus, "frames": n, "frames_per_s": n / (t1 - t0),
                  "write_combined_uploads": bool(os.environ.get("BS_HOST_ALLOC_WC"))}))
pool.close()
