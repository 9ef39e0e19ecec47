#!/usr/bin/env python
"""Randomised parity run, longer than the test-suite's: random camera counts, sizes, homographies (near-unit-scale
panorama rigs that take the compact kernel, and strong ones that take the general kernel), both border rules,
KEEP_CANVAS, lens models, BS_FLAG_FEATHER -- every panorama compared bit for bit with the oracle.
usage (GPU box): python tools/fuzz.py [seconds] [seed] [size scale, default 1]"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import oracle
p = importlib.import_module("image-stitching_b200")
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
SC = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0

def rig(n, sw, sh):      # side-by-side cameras with overlap, mild rotation/perspective: the compact kernel's case
    step = rng.uniform(0.45, 0.95) * sw
    Hs = []
    for i in range(n):
        a = rng.uniform(-0.02, 0.02); s = rng.uniform(0.9, 1.12)
        Hs.append(np.array([[s * np.cos(a), -s * np.sin(a), i * step + rng.uniform(-8, 8)],
                            [s * np.sin(a), s * np.cos(a), rng.uniform(-10, 10)],
                            [rng.uniform(-3e-6, 3e-6), rng.uniform(-3e-6, 3e-6), rng.uniform(0.97, 1.03)]]))
    return Hs, (int(step * (n - 1) + sw * 1.05), int(sh * 1.05))

def strong(n, dw, dh):
    Hs = []
    for _ in range(n):
        ang = rng.uniform(-0.7, 0.7); s = rng.uniform(0.45, 2.2)
        Hs.append(np.array([[s * np.cos(ang), -s * np.sin(ang), rng.uniform(-40, dw / 2)],
                            [s * np.sin(ang), s * np.cos(ang), rng.uniform(-40, dh / 2)],
                            [rng.uniform(-4e-4, 4e-4), rng.uniform(-4e-4, 4e-4), rng.uniform(0.8, 1.25)]]))
    return Hs

def lens(sw, sh):
    K = np.array([[rng.uniform(0.8, 1.2) * sw, 0, sw / 2 + rng.uniform(-9, 9)], [0, rng.uniform(0.8, 1.2) * sw, sh / 2 + rng.uniform(-9, 9)], [0, 0, 1]])
    d = np.array([rng.uniform(-0.25, 0.15), rng.uniform(-0.05, 0.05), rng.uniform(-2e-3, 2e-3), rng.uniform(-2e-3, 2e-3), rng.uniform(-0.02, 0.02)])
    return K, d

t0 = time.time(); n_cases = 0; kinds = {}
while time.time() - t0 < budget:
    n = int(rng.integers(1, 6))
    compact = rng.random() < 0.6
    if compact:
        sw, sh = int(rng.integers(160, 900 * SC)), int(rng.integers(64, 500 * SC))
        whs = [(sw, sh)] * n
        Hs, (dw, dh) = rig(n, sw, sh)
    else:
        dw, dh = int(rng.integers(1, 1200 * SC)), int(rng.integers(1, 700 * SC))
        whs = [(int(rng.integers(1, 700 * SC)), int(rng.integers(1, 500 * SC))) for _ in range(n)]
        Hs = strong(n, dw, dh)
    srcs = [rng.integers(0, 256, (h_, w_, 3), dtype=np.uint8) for w_, h_ in whs]
    mode = rng.choice(["plain", "legacy", "keep", "feather", "lens", "feather_legacy", "lens_legacy", "lens_legacy"])
    flags = {"plain": 0, "legacy": p.FLAG_LEGACY_BORDER, "keep": p.FLAG_KEEP_CANVAS, "feather": p.FLAG_FEATHER, "lens": 0,
             "feather_legacy": p.FLAG_FEATHER | p.FLAG_LEGACY_BORDER, "lens_legacy": p.FLAG_LEGACY_BORDER}[mode]
    rule = oracle.BORDER_LEGACY if "legacy" in mode else oracle.BORDER_CV4
    models = {}
    if mode.startswith("lens"):
        for i in range(n):
            if rng.random() < 0.7 and whs[i][0] >= 16 and whs[i][1] >= 16: models[i] = lens(*whs[i])
    und = [oracle.undistort(np.ascontiguousarray(srcs[i]), *models[i]) if i in models else srcs[i] for i in range(n)]
    prior = rng.integers(0, 256, (dh, dw, 3), dtype=np.uint8)
    if "feather" in mode:
        want = oracle.stitch_feather(und, Hs, (dw, dh), border_rule=rule)
    elif mode == "keep":
        want = prior.copy()
        for s_, H_ in zip(und, Hs): oracle.warp_perspective(s_, H_, want, rule)
    else:
        want = oracle.stitch(und, Hs, (dw, dh), rule)
    with p.Stitcher(whs, (dw, dh), Hs, flags=flags) as st:
        for i, m in models.items(): st.set_undistort(i, *m)
        out = prior.copy()
        st.wait(st.submit(srcs, out))
        stt = st.stats()
        kern = stt["frame_kernel"]
        if stt["h2d_bytes_last_submit"] < sum(w_ * h_ * 3 for w_, h_ in whs): mode = mode + "+trimmed"      # only part of the frames was uploaded
    bad = int((out != want).any(axis=2).sum())
    kinds[(mode, kern)] = kinds.get((mode, kern), 0) + 1
    if bad:
        print("MISMATCH", dict(mode=mode, n=n, whs=whs, canvas=(dw, dh), kernel=kern, bad_pixels=bad, case=n_cases))
        np.savez(os.path.join(ROOT, "gpurun_out", f"fuzz_fail_{n_cases}.npz"), Hs=np.array(Hs), whs=np.array(whs), canvas=np.array([dw, dh]), flags=flags)
        sys.exit(1)
    n_cases += 1
print(f"{n_cases} cases bit-exact in {time.time() - t0:.0f} s; (mode, frame kernel) counts: {sorted(kinds.items())}")
