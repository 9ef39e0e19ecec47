#!/usr/bin/env python
"""Randomised check of the GPU JPEG decoder against cv2.imdecode (libjpeg-turbo): random sizes, contents (noise,
blurred noise, flat shapes, stripes, camera-like), qualities, restart intervals, 4:2:0 / 4:4:4 -- every frame through
bs_submit_jpeg with an identity homography must come out as cv2 decodes it, bit for bit.
usage (GPU box): python tools/fuzz_jpeg.py [seconds] [seed]"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2
import numpy as np
p = importlib.import_module("image-stitching_b200")
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
t0 = time.time(); n = 0; kinds = {}
while time.time() - t0 < budget:
    h, w = int(rng.integers(16, 1100)), int(rng.integers(16, 2000))
    kind = int(rng.integers(0, 5))
    if kind == 0: img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    elif kind == 1: img = cv2.GaussianBlur(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), (0, 0), float(rng.uniform(0.5, 6)))
    elif kind == 2:
        img = np.zeros((h, w, 3), np.uint8); img[:] = rng.integers(0, 256, 3)
        cv2.circle(img, (w // 2, h // 2), max(2, min(w, h) // 3), tuple(int(v) for v in rng.integers(0, 256, 3)), -1)
    elif kind == 3:
        img = np.dstack([np.tile(np.linspace(0, 255, w).astype(np.uint8), (h, 1))] * 3).copy(); img[::7] = 255 - img[::7]
    else:
        base = cv2.GaussianBlur(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), (0, 0), 8)
        img = np.clip(base.astype(int) + rng.integers(-6, 7, (h, w, 3)), 0, 255).astype(np.uint8)
    args = [cv2.IMWRITE_JPEG_QUALITY, int(rng.integers(20, 101))]
    rst = int(rng.choice([0, 0, 1, 5, (w + 15) // 16, 3000]))
    if rst: args += [cv2.IMWRITE_JPEG_RST_INTERVAL, rst]
    if rng.random() < 0.3: args += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444]
    bits = cv2.imencode(".jpg", img, args)[1].tobytes()
    ref = cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)
    with p.Stitcher([(w, h)], (w, h), [np.eye(3)]) as st:
        out = np.empty((h, w, 3), np.uint8)
        st.wait(st.submit_jpeg([bits], out))
        be = st.jpeg_backend()
    if not be.startswith("b200") or not np.array_equal(out, ref):
        print("MISMATCH", dict(case=n, w=w, h=h, kind=kind, args=args, backend=be, bad=int((out != ref).any(axis=2).sum())))
        open(os.path.join(ROOT, "gpurun_out", f"fuzz_jpeg_fail_{n}.jpg"), "wb").write(bits)
        sys.exit(1)
    kinds[kind] = kinds.get(kind, 0) + 1
    n += 1
print(f"{n} JPEGs decoded on the GPU equal cv2.imdecode bit for bit in {time.time() - t0:.0f} s; per kind (noise, blurred, flat shape, stripes, camera-like): {sorted(kinds.items())}")
