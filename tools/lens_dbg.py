"""Debug aid: which 32x16 blocks of a lens-plan undistort differ from the oracle."""
import sys, importlib, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oracle, torch
p = importlib.import_module('image-stitching_b200')
w, h = 640, 360
src = oracle.gen_frame(w, h, 1, 0, oracle.NOISE)
K = np.array([[0.9 * w, 0, w / 2 + 3.3], [0, 0.92 * w, h / 2 - 1.7], [0, 0, 1]])
dist = [-0.2, 0.05, 0.001, -0.002, 0.01]
want = oracle.undistort(src, K, np.array(dist))
d_src = torch.from_numpy(src).cuda(); d_dst = torch.zeros_like(d_src)
with p.Stitcher([(w, h)], (64, 32), [np.eye(3)]) as st:
    st.set_undistort(0, K, dist)
    for rep in range(2):
        d_dst.zero_()
        st.undistort_device(0, d_src.data_ptr(), d_src.stride(0), d_dst.data_ptr(), d_dst.stride(0))
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        diff = (got != want).any(axis=2)
        print("rep", rep, "differ:", int(diff.sum()))
        bad = {}
        ys, xs = np.nonzero(diff)
        for y, x in zip(ys, xs):
            bad[(x // 32, y // 16)] = bad.get((x // 32, y // 16), 0) + 1
        print(sorted(bad.items())[:60])
        if diff.any():
            y, x = ys[0], xs[0]
            print("first", x, y, got[y, x], want[y, x], "zero?", int((got[diff] == 0).all()))
    print("--- host path: upload, lens plan kernel, frame kernel (identity warp)")
with p.Stitcher([(w, h)], (w, h), [np.eye(3)]) as st:
    st.set_undistort(0, K, dist)
    for rep in range(3):
        got = st.stitch([src])
        diff = (got != want).any(axis=2)
        bad = {}
        ys, xs = np.nonzero(diff)
        for y, x in zip(ys, xs):
            bad[(x // 32, y // 16)] = bad.get((x // 32, y // 16), 0) + 1
        print("rep", rep, "differ:", int(diff.sum()), sorted(bad.items())[:40])
        if diff.any():
            y, x = ys[0], xs[0]
            print("first", x, y, got[y, x], want[y, x], "src there", src[y, x], "all zero?", int((got[diff] == 0).all()))
