"""Hand-run timing of the warp with and without cv::undistort in front (4x4K -> 7680x2160, device-resident):
python tests/time_undistort.py on a GPU box; the ncu summary under profiles/ was captured from it."""
import sys, importlib, numpy as np, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
p = importlib.import_module('image-stitching_b200'); wl = importlib.import_module('image-stitching_b200.workloads')
w = wl.get('cfg3'); sw, sh = w.src_wh; W, H = w.canvas_wh
dev = torch.device('cuda:0')
frames = [torch.from_numpy(wl.gen_frame_numpy(sw, sh, i, 0)).to(dev) for i in range(w.n_cams)]
canvas = torch.empty((H, W, 3), dtype=torch.uint8, device=dev)
st = p.Stitcher([w.src_wh] * w.n_cams, w.canvas_wh, w.H)
s = torch.cuda.Stream(); torch.cuda.set_stream(s)
ptrs = [f.data_ptr() for f in frames]; pit = [sw * 3] * w.n_cams
def run(n):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    for _ in range(5): st.submit_device(ptrs, pit, canvas.data_ptr(), W * 3, s.cuda_stream)
    torch.cuda.synchronize(); e0.record(s)
    for _ in range(n): st.submit_device(ptrs, pit, canvas.data_ptr(), W * 3, s.cuda_stream)
    e1.record(s); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n
print('warp only us/frame', run(50))
K = np.array([[0.9 * sw, 0, sw / 2 + 3.3], [0, 0.92 * sw, sh / 2 - 1.7], [0, 0, 1]])
for i in range(w.n_cams): st.set_undistort(i, K, [-0.2, 0.05, 0.001, -0.002, 0.01])
print('undistort x4 + warp us/frame', run(50))
