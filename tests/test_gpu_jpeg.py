"""Row f2 of SURVEY.md section 8: compressed frames in, compressed panorama out (bs_submit_jpeg), the device-side
form of imread / imwrite at /root/reference/tools/stitch/main.cpp:107,114,125.  nvJPEG is the codec (libjpeg is the
reference's); two JPEG decoders may differ by a unit or two per pixel, so the parity statement here is: the
panorama made from compressed frames is BIT-EQUAL to the oracle applied to the pixels nvJPEG decodes (obtained
through the same entry point with an identity homography); how far those pixels are from cv2's decode is reported
and loosely bounded, not asserted as parity."""
import json
import os

import numpy as np
import pytest

import oracle
from conftest import GOLDEN, pkg

pytestmark = pytest.mark.gpu


def fixtures():
    jpgs = [open(os.path.join(GOLDEN, f"jpeg_src{i}.jpg"), "rb").read() for i in range(2)]
    z = np.load(os.path.join(GOLDEN, "real_crop.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    cv2_dec = np.load(os.path.join(GOLDEN, "jpeg_cv2_decode.npz"))
    return jpgs, z["H"], tuple(meta["canvas_wh"]), [cv2_dec["src0_cv2"], cv2_dec["src1_cv2"]]


def nvjpeg_pixels(p, jpg, wh):
    """The frame as nvJPEG decodes it: identity homography, canvas = frame, raw output (an exact copy)."""
    with p.Stitcher([wh], wh, [np.eye(3)]) as st:
        out = np.empty((wh[1], wh[0], 3), np.uint8)
        st.wait(st.submit_jpeg([jpg], out))
    return out


def test_compressed_frames_give_the_oracle_panorama_of_the_decoded_pixels():
    p = pkg()
    jpgs, H, canvas_wh, cv2_dec = fixtures()
    whs = [(d.shape[1], d.shape[0]) for d in cv2_dec]
    decoded = [nvjpeg_pixels(p, j, wh) for j, wh in zip(jpgs, whs)]
    for d, c in zip(decoded, cv2_dec):
        diff = np.abs(d.astype(int) - c.astype(int))
        print(f"nvJPEG vs cv2 decode: max |d| = {diff.max()}, {100.0 * (diff > 0).mean():.2f} % of samples differ")
        assert diff.max() <= 4           # two decoders of the same bitstream, not a parity claim
    want = oracle.stitch(decoded, H, canvas_wh)
    with p.Stitcher(whs, canvas_wh, H) as st:
        outs = [np.empty((canvas_wh[1], canvas_wh[0], 3), np.uint8) for _ in range(4)]
        tickets = [st.submit_jpeg(jpgs, o) for o in outs]          # more frames than pipeline slots: slots are reused
        for t in tickets:
            st.wait(t)
        raw_too = st.stitch(decoded)                                # the raw entry point on the same pixels
    for o in outs:
        assert np.array_equal(o, want)
    assert np.array_equal(raw_too, want)


def test_compressed_panorama_out_and_errors():
    p = pkg()
    cv2 = pytest.importorskip("cv2")
    jpgs, H, canvas_wh, cv2_dec = fixtures()
    whs = [(d.shape[1], d.shape[0]) for d in cv2_dec]
    with p.Stitcher(whs, canvas_wh, H) as st:
        raw = np.empty((canvas_wh[1], canvas_wh[0], 3), np.uint8)
        t = st.submit_jpeg(jpgs, raw, jpeg_capacity=4 << 20, quality=95)      # raw and JPEG of the same panorama
        st.wait(t)
        bits = st.jpeg_result(t)
        t2 = st.submit_jpeg(jpgs, None, jpeg_capacity=4 << 20)                # JPEG only: no raw canvas comes down
        st.wait(t2)
        bits2 = st.jpeg_result(t2)
        assert bits[:2] == b"\xff\xd8" and bits[-2:] == b"\xff\xd9" and 1000 < len(bits) < raw.nbytes // 4
        back = cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)
        assert back.shape == raw.shape
        mse = float(((back.astype(float) - raw.astype(float)) ** 2).mean())
        psnr = 10 * np.log10(255.0 ** 2 / max(mse, 1e-9))
        print(f"panorama JPEG: {len(bits)} bytes for {raw.nbytes} raw, PSNR {psnr:.1f} dB")
        assert psnr > 30.0
        assert len(bits2) > 1000 and cv2.imdecode(np.frombuffer(bits2, np.uint8), cv2.IMREAD_COLOR).shape == raw.shape
        # a buffer that is too small is reported with the size needed; the handle stays usable
        t3 = st.submit_jpeg(jpgs, None, jpeg_capacity=64)
        with pytest.raises(p.StitchError) as e:
            st.wait(t3)
        assert "needs" in str(e.value)
        with pytest.raises(p.StitchError) as e:
            st.submit_jpeg([b"not a jpeg at all", jpgs[1]], raw)
        assert e.value.code == -1 and "frame 0" in str(e.value)
        with pytest.raises(p.StitchError) as e:
            st.submit_jpeg([jpgs[0], jpgs[0][:200]], raw)                     # truncated bitstream
        assert "frame 1" in str(e.value)
        t4 = st.submit_jpeg(jpgs, raw)
        st.wait(t4)
    decoded = [nvjpeg_pixels(p, j, wh) for j, wh in zip(jpgs, whs)]
    assert np.array_equal(raw, oracle.stitch(decoded, H, canvas_wh))
