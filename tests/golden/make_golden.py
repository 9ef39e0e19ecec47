#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ from the reference's own OpenCV path.

Run in the BUILD container only (needs cv2 and, for cfg1, /root/reference/test/Inputs):

    python tests/golden/make_golden.py

Every expected output here comes from the same OpenCV entry points the reference calls
(/root/reference/tools/stitch/main.cpp:116-117, :148-151), reached through the cv2 binding:

    cv2.warpPerspective(src, H, (W, H), dst=canvas, flags=INTER_LINEAR, borderMode=BORDER_TRANSPARENT)
    cv2.undistort(src, K, dist)

Neither the oracle (oracle/) nor the CUDA library is imported: the vectors are independent of
both, and both are tested against them.  Frames are made with a numpy copy of the generator
of SURVEY.md section 8(d) so the C generator in the oracle is cross-checked as well.
"""
import hashlib
import importlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
workloads = importlib.import_module("image-stitching_b200.workloads")

REF_INPUTS = "/root/reference/test/Inputs"


def gen_frame(w, h, cam, frame, kind):
    x = np.arange(w, dtype=np.uint32)[None, :, None]
    y = np.arange(h, dtype=np.uint32)[:, None, None]
    c = np.arange(3, dtype=np.uint32)[None, None, :]
    if kind == "smooth":
        return ((x + 2 * y + 64 * c + np.uint32(7 * frame + 31 * cam)) & 255).astype(np.uint8)
    with np.errstate(over="ignore"):
        v = (x * np.uint32(0x9E3779B1)) ^ (y * np.uint32(0x85EBCA77)) ^ (c * np.uint32(0xC2B2AE3D))
        v = v ^ (np.uint32(cam * 131 + frame) * np.uint32(0x27D4EB2F))
        v ^= v >> 15
        v *= np.uint32(0x2C1B3C6D)
        v ^= v >> 12
        v *= np.uint32(0x297A2D39)
        v ^= v >> 15
    return (v >> 24).astype(np.uint8)


def cv_stitch(srcs, Hs, canvas_wh, canvas=None):
    W, Hc = canvas_wh
    if canvas is None:
        canvas = np.zeros((Hc, W, 3), np.uint8)
    for s, H in zip(srcs, Hs):
        cv2.warpPerspective(s, np.asarray(H, np.float64).reshape(3, 3), (W, Hc), dst=canvas,
                            flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_TRANSPARENT)
    return canvas


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def known_answers():
    out = {"opencv": cv2.__version__, "cases": []}
    plan = [("cfg2", "noise", 0), ("cfg2", "noise", 1), ("cfg2", "smooth", 0), ("cfg2", "smooth", 1),
            ("cfg3", "noise", 0), ("cfg3", "noise", 1), ("cfg3", "smooth", 0), ("cfg3", "smooth", 1),
            ("cfg5", "noise", 0),
            # cfg1's strong-perspective homographies on generated frames of the photos' size,
            # so the GPU box (which has no /root/reference) can check them at full size
            ("cfg1", "noise", 0), ("cfg1", "smooth", 0)]
    for name, kind, frame in plan:
        w = workloads.get(name)
        srcs = [gen_frame(w.src_wh[0], w.src_wh[1], i, frame, kind) for i in range(w.n_cams)]
        canvas = cv_stitch(srcs, w.H, w.canvas_wh)
        out["cases"].append({"config": name, "content": kind, "frame": frame,
                             "src0_sha256": sha(srcs[0]), "canvas_sha256": sha(canvas),
                             "covered_fraction": float((canvas != 0).any(axis=2).mean())})
        print(name, kind, frame, out["cases"][-1]["canvas_sha256"][:16])
    # configs[0] proper: the two chessboard photos
    if os.path.isdir(REF_INPUTS):
        w = workloads.get("cfg1")
        imgs = [cv2.imread(os.path.join(REF_INPUTS, n)) for n in ("1a.jpg", "1b.jpg")]
        canvas = cv_stitch(imgs, w.H, w.canvas_wh)
        out["cfg1_photos"] = {"inputs": ["test/Inputs/1a.jpg", "test/Inputs/1b.jpg"],
                              "decoded_sha256": [sha(i) for i in imgs],
                              "canvas_sha256": sha(canvas),
                              "covered_fraction": float((canvas != 0).any(axis=2).mean())}
        print("cfg1 photos", out["cfg1_photos"]["canvas_sha256"][:16])
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump(out, f, indent=1)


def rand_h(rng, dw, dh, strong):
    ang = rng.uniform(-0.7, 0.7)
    s = rng.uniform(0.45, 2.2)
    p = 4e-4 if strong else 2e-5
    return np.array([[s * np.cos(ang), -s * np.sin(ang), rng.uniform(-40, dw / 2)],
                     [s * np.sin(ang), s * np.cos(ang), rng.uniform(-40, dh / 2)],
                     [rng.uniform(-p, p), rng.uniform(-p, p), rng.uniform(0.8, 1.25)]])


def small_cases():
    """Small full-array fixtures: random homographies, ragged sizes, non-zero prior canvas,
    multi-camera overlaps, degenerate H."""
    rng = np.random.default_rng(20261106)
    arrays = {}
    meta = []
    shapes = [((64, 48), (96, 64), 1), ((37, 53), (131, 71), 2), ((120, 90), (222, 97), 3),
              ((16, 16), (1, 1), 1), ((9, 150), (257, 33), 2), ((128, 96), (64, 16), 1),
              ((50, 50), (63, 15), 2), ((160, 100), (200, 120), 4), ((3, 3), (40, 40), 1),
              ((2, 2), (30, 20), 1), ((1, 1), (8, 8), 1), ((100, 80), (65, 17), 2)]
    for k, ((sw, sh), (dw, dh), n) in enumerate(shapes):
        srcs = [rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8) for _ in range(n)]
        Hs = np.stack([rand_h(rng, dw, dh, strong=(k % 2 == 1)) for _ in range(n)])
        if k == 8:   # exact integer translation + unit scale: every pixel on a=b=0
            Hs[0] = np.array([[1, 0, 5], [0, 1, 7], [0, 0, 1]], float)
        if k == 9:   # magnification by 8: many pixels on the last row/column rule
            Hs[0] = np.array([[8, 0, 1.5], [0, 8, 2.25], [0, 0, 1]], float)
        if k == 10:  # a single source pixel
            Hs[0] = np.array([[3, 0, 2], [0, 3, 2], [0, 0, 1]], float)
        prior = rng.integers(0, 256, (dh, dw, 3), dtype=np.uint8)
        exp_zero = cv_stitch(srcs, Hs, (dw, dh))
        exp_prior = cv_stitch(srcs, Hs, (dw, dh), canvas=prior.copy())
        for i, s in enumerate(srcs):
            arrays[f"c{k}_src{i}"] = s
        arrays[f"c{k}_H"] = Hs
        arrays[f"c{k}_prior"] = prior
        arrays[f"c{k}_exp_zero"] = exp_zero
        arrays[f"c{k}_exp_prior"] = exp_prior
        meta.append({"case": k, "n_cams": n, "src_wh": [sw, sh], "canvas_wh": [dw, dh]})
    # singular homography: cv::invert leaves M = 0 -> every pixel maps to source (0,0)
    src = rng.integers(0, 256, (12, 10, 3), dtype=np.uint8)
    Hs = np.array([[[1, 2, 3], [2, 4, 6], [0, 0, 1]]], float)
    k = len(shapes)
    arrays[f"c{k}_src0"] = src
    arrays[f"c{k}_H"] = Hs
    arrays[f"c{k}_prior"] = np.zeros((20, 24, 3), np.uint8)
    arrays[f"c{k}_exp_zero"] = cv_stitch([src], Hs, (24, 20))
    arrays[f"c{k}_exp_prior"] = arrays[f"c{k}_exp_zero"]
    meta.append({"case": k, "n_cams": 1, "src_wh": [10, 12], "canvas_wh": [24, 20], "singular": True})
    arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "small_cases.npz"), **arrays)
    print("small cases:", len(meta))


def undistort_cases():
    rng = np.random.default_rng(7)
    arrays = {}
    meta = []
    dists = [[-0.2, 0.05, 0.001, -0.002, 0.01],
             [-0.3, 0.1, 0.002, 0.001, -0.02, 0, 0, 0],
             [0, 0, 0, 0, 0],
             [0.1, -0.05, 0, 0],
             [-0.25, 0.08, 0.001, 0.0005, 0.01, 0.02, -0.01, 0.003]]
    for k, (w, h) in enumerate([(96, 64), (131, 71), (640, 48), (5000, 6), (64, 64)]):
        src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        K = np.array([[0.9 * w, 0, w / 2 + 3.3], [0, 0.92 * w, h / 2 - 1.7], [0, 0, 1]])
        d = np.array(dists[k], float)
        arrays[f"u{k}_src"] = src
        arrays[f"u{k}_K"] = K
        arrays[f"u{k}_dist"] = d
        arrays[f"u{k}_exp"] = cv2.undistort(src, K, d)
        meta.append({"case": k, "wh": [w, h], "n_dist": len(d)})
    arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "undistort_cases.npz"), **arrays)
    print("undistort cases:", len(meta))


def border_constant_cases():
    """cv::warpPerspective with BORDER_CONSTANT (value 0), the variant projectToTheFloor uses
    (lib/Calibrate/utils.cpp:390)."""
    rng = np.random.default_rng(11)
    arrays = {}
    meta = []
    for k, ((sw, sh), (dw, dh)) in enumerate([((90, 60), (200, 120)), ((333, 77), (300, 260)), ((2, 2), (70, 40)),
                                               ((320, 240), (500, 350)), ((50, 50), (33, 9))]):
        src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        H = rand_h(rng, dw, dh, strong=bool(k & 1))
        arrays[f"b{k}_src"] = src
        arrays[f"b{k}_H"] = H
        arrays[f"b{k}_exp"] = cv2.warpPerspective(src, H, (dw, dh), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_CONSTANT,
                                                  borderValue=0)
        meta.append({"case": k, "src_wh": [sw, sh], "canvas_wh": [dw, dh]})
    arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "border_constant_cases.npz"), **arrays)
    print("border-constant cases:", len(meta))


if __name__ == "__main__":
    cv2.setNumThreads(cv2.getNumberOfCPUs())
    small_cases()
    undistort_cases()
    border_constant_cases()
    known_answers()
