"""Real image content for the GPU parity tests (the GPU box has no /root/reference, so it only ever saw
synthetic frames).  Run in the build container:  python tests/golden/make_real_crop.py

Takes a 320x240 crop out of each of the reference's two chessboard photos (test/Inputs/1a.jpg, 1b.jpg, decoded
with cv2.imread as the reference does, tools/stitch/main.cpp:107,114), warps them with cfg1-like homographies
(strong perspective, rotation, 2.4x magnification: SURVEY.md App. B.1 rescaled to the crop) through the OpenCV
call the reference makes, and records the crops plus the SHA-256 of the resulting canvases under both border rules'
defining library (cv2 4.x rule only: the 2.4 rule has no library here).  Imports neither the oracle nor the CUDA
library."""
import hashlib
import json
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/test/Inputs"
CANVAS = (900, 1100)      # (W, H)
# cfg1's two homographies with the translation rescaled to a 320x240 crop and a 900x1100 canvas
H = np.array([
    [[1.1120443247359568, 0.2321184108605204, 60.5], [-0.7385466102701779, 2.359969315805939, 330.25],
     [-0.0013831019805894855, 0.0018804770166660635, 1.0]],
    [[1.1355734419545565, 0.16166738072381795, 20.0], [-0.5899936756162424, 2.554407720031592, 520.75],
     [-0.0009468468147649584, 0.0018031156578367973, 1.0]],
])


def main():
    crops = []
    for name, (x0, y0) in (("1a.jpg", (1300, 1900)), ("1b.jpg", (1500, 2000))):
        img = cv2.imread(os.path.join(SRC, name))
        assert img is not None and img.shape == (4032, 3016, 3), img.shape if img is not None else None
        crops.append(np.ascontiguousarray(img[y0:y0 + 240, x0:x0 + 320]))
    canvas = np.zeros((CANVAS[1], CANVAS[0], 3), np.uint8)
    for s, h in zip(crops, H):
        cv2.warpPerspective(s, h, CANVAS, dst=canvas, flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_TRANSPARENT)
    meta = {"canvas_wh": list(CANVAS), "canvas_sha256": hashlib.sha256(canvas.tobytes()).hexdigest(),
            "covered_fraction": float((canvas.any(axis=2)).mean()), "cv2": cv2.__version__,
            "sources": "320x240 crops of /root/reference/test/Inputs/1a.jpg at (1300,1900) and 1b.jpg at (1500,2000)"}
    np.savez_compressed(os.path.join(HERE, "real_crop.npz"), src0=crops[0], src1=crops[1], H=H,
                        meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8))
    print(meta, os.path.getsize(os.path.join(HERE, "real_crop.npz")))


if __name__ == "__main__":
    main()
