#!/usr/bin/env python
"""Device time per panorama with and without BS_FLAG_FEATHER (opt-in blending, not the reference's result) at one
configuration, frames resident in HBM.  Not part of `value`.  usage: python tools/feather_bench.py [cfg3]"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
p = importlib.import_module("image-stitching_b200"); wl = importlib.import_module("image-stitching_b200.workloads")
w = wl.get(sys.argv[1] if len(sys.argv) > 1 else "cfg3"); sw, sh = w.src_wh; W, H = w.canvas_wh
sets = [[torch.from_numpy(wl.gen_frame_numpy(sw, sh, i, f, "noise")).cuda() for i in range(w.n_cams)] for f in range(12)]
outs = [torch.empty((H, W, 3), dtype=torch.uint8, device="cuda") for _ in range(12)]
stream = torch.cuda.Stream()
res = {"what": "device us per panorama, frames in HBM", "workload": w.name}
for name, flags in (("last_camera_wins", 0), ("feather", p.FLAG_FEATHER)):
    with p.Stitcher([w.src_wh] * w.n_cams, w.canvas_wh, w.H, flags=flags) as st:
        def go(n):
            for k in range(n):
                fs = sets[k % 12]
                st.submit_device([t.data_ptr() for t in fs], [t.stride(0) for t in fs], outs[k % 12].data_ptr(), outs[k % 12].stride(0),
                                 stream.cuda_stream, submit_flags=p.SUBMIT_SOURCES_SETTLED)
        go(6); stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(); go(48); e1.record()
        stream.synchronize()
        s = st.stats()
        res[name] = {"us": 1000 * e0.elapsed_time(e1) / 48, "blocks_staged": s["n_tiles_fast"], "blocks_per_pixel": s["n_tiles_exact"]}
print(json.dumps(res))
