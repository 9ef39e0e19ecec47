"""Wider random sweep than the test-suite (run by hand on a GPU box: python tests/fuzz_parity.py SEED N):
many seeds, mild to strong perspective, large canvases, near-identity maps (the staged kernel's
territory), both border rules, keep-canvas.  CUDA path through the C ABI against the CPU oracle."""
import sys, importlib, time, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import oracle
p = importlib.import_module('image-stitching_b200')

def rand_h(rng, dw, dh, kind):
    if kind == 0:      # near identity / panorama-like
        ang = rng.uniform(-0.06, 0.06); s = rng.uniform(0.9, 1.15); pp = rng.choice([0, 2e-6, 2e-5])
        tx, ty = rng.uniform(-50, dw * 0.7), rng.uniform(-40, 40)
    elif kind == 1:    # moderate
        ang = rng.uniform(-0.4, 0.4); s = rng.uniform(0.6, 1.7); pp = 1e-4
        tx, ty = rng.uniform(-40, dw / 2), rng.uniform(-40, dh / 2)
    else:              # strong
        ang = rng.uniform(-0.8, 0.8); s = rng.uniform(0.4, 2.5); pp = 6e-4
        tx, ty = rng.uniform(-40, dw / 2), rng.uniform(-40, dh / 2)
    return np.array([[s * np.cos(ang), -s * np.sin(ang), tx], [s * np.sin(ang), s * np.cos(ang), ty],
                     [rng.uniform(-pp, pp), rng.uniform(-pp, pp), rng.uniform(0.9, 1.1)]])

seed0 = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n_trials = int(sys.argv[2]) if len(sys.argv) > 2 else 150
bad = 0; t0 = time.time()
for trial in range(n_trials):
    rng = np.random.default_rng(seed0 * 100000 + trial)
    kind = trial % 3
    legacy = bool((trial // 3) & 1)
    keep = (trial % 7) == 0
    n = int(rng.integers(1, 6))
    big = (trial % 5) == 0
    dw, dh = (int(rng.integers(1500, 5000)), int(rng.integers(600, 2300))) if big else (int(rng.integers(1, 1800)), int(rng.integers(1, 1000)))
    srcs, Hs = [], []
    for _ in range(n):
        sw, sh = (int(rng.integers(800, 2600)), int(rng.integers(500, 1500))) if big else (int(rng.integers(1, 1000)), int(rng.integers(1, 800)))
        srcs.append(rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8))
        Hs.append(rand_h(rng, dw, dh, kind))
    rule = oracle.BORDER_LEGACY if legacy else oracle.BORDER_CV4
    flags = (p.FLAG_LEGACY_BORDER if legacy else 0) | (p.FLAG_KEEP_CANVAS if keep else 0)
    prior = rng.integers(0, 256, (dh, dw, 3), dtype=np.uint8) if keep else None
    if keep:
        want = prior.copy()
        for s_, H_ in zip(srcs, Hs): oracle.warp_perspective(s_, H_, want, rule)
    else:
        want = oracle.stitch(srcs, Hs, (dw, dh), rule)
    whs = [(s.shape[1], s.shape[0]) for s in srcs]
    with p.Stitcher(whs, (dw, dh), Hs, flags=flags) as st:
        out = prior.copy() if keep else None
        got = st.stitch(srcs, out) if keep else st.stitch(srcs)
        stt = st.stats()
    if not np.array_equal(got, want):
        bad += 1
        print(f"MISMATCH seed {seed0} trial {trial}: kind {kind} legacy {legacy} keep {keep} n {n} canvas {dw}x{dh}: {(got != want).any(axis=2).sum()} px", flush=True)
print(f"seed {seed0}: {n_trials} trials, {bad} mismatches, {time.time() - t0:.1f} s")
