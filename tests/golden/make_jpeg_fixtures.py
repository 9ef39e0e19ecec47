"""JPEG fixtures for the compressed-frame entry point (bs_submit_jpeg).  Run in the build container:
    python tests/golden/make_jpeg_fixtures.py
Encodes the two real-content crops of real_crop.npz (made by make_real_crop.py from the reference's photos) as
baseline JPEG with cv2.imencode -- the encoder behind the reference's cv::imwrite -- at quality 92, 4:2:0, and
stores cv2's own decode of them for the decoder-difference report.  Imports neither the oracle nor the CUDA library."""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
z = np.load(os.path.join(HERE, "real_crop.npz"))
out = {}
for k in ("src0", "src1"):
    ok, buf = cv2.imencode(".jpg", z[k], [cv2.IMWRITE_JPEG_QUALITY, 92])
    assert ok
    with open(os.path.join(HERE, f"jpeg_{k}.jpg"), "wb") as f:
        f.write(buf.tobytes())
    out[k + "_cv2"] = cv2.imdecode(buf, cv2.IMREAD_COLOR)
    print(k, len(buf), "bytes; max |decoded - original| =", int(np.abs(out[k + "_cv2"].astype(int) - z[k].astype(int)).max()))
np.savez_compressed(os.path.join(HERE, "jpeg_cv2_decode.npz"), **out)
