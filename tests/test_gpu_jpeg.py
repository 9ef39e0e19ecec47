"""Row f2 of SURVEY.md section 8: compressed frames in, compressed panorama out (bs_submit_jpeg), the device-side
form of imread / imwrite at /root/reference/tools/stitch/main.cpp:107,114,125.  Baseline 4:2:0 / 4:4:4 JPEGs are
decoded by the library's own GPU decoder, whose pixels equal cv2.imdecode's (tests/test_jpegdec.py pins that); other
JPEGs, and the panorama encoder, are nvJPEG's, and two JPEG decoders may differ by a unit or two per pixel.  The
parity statements here therefore come in two strengths: the panorama made from compressed frames is BIT-EQUAL to the
oracle applied to the pixels the same entry point decodes (obtained with an identity homography) -- whatever the
decoder -- and, for the fixtures (baseline 4:2:0), also to the oracle applied to cv2's decode, i.e. to the
reference's own imread + warpPerspective."""
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
        t5 = st.submit_jpeg([jpgs[0], jpgs[0][:200]], raw)                    # truncated bitstream: the header is fine, so it is
        with pytest.raises(p.StitchError) as e:                               # the decode that fails -- reported by bs_wait
            st.wait(t5)
        assert "frame 1" in str(e.value)
        t4 = st.submit_jpeg(jpgs, raw)
        st.wait(t4)
    decoded = [nvjpeg_pixels(p, j, wh) for j, wh in zip(jpgs, whs)]
    assert np.array_equal(raw, oracle.stitch(decoded, H, canvas_wh))


def test_decoders_agree():
    """The library's own GPU decoder gives cv2's pixels; nvJPEG behind it (BS_DEBUG_JPEG_NO_GPU_ENTROPY: entropy stages
    on worker threads; additionally BS_DEBUG_JPEG_THREADS=0: nvJPEG's one-call decoders on the calling thread -- child
    processes, the switches are read once) gives the same pixels either way; a damaged bitstream leaves the handle
    usable."""
    import hashlib
    import subprocess
    import sys
    from conftest import ROOT
    p = pkg()
    jpgs, H, canvas_wh, cv2_dec = fixtures()
    whs = [(d.shape[1], d.shape[0]) for d in cv2_dec]
    out = np.empty((canvas_wh[1], canvas_wh[0], 3), np.uint8)
    with p.Stitcher(whs, canvas_wh, H) as st:
        st.wait(st.submit_jpeg(jpgs, out))
        assert st.jpeg_backend().startswith("b200")
        assert np.array_equal(out, oracle.stitch(cv2_dec, H, canvas_wh)), "JPEG in = the reference's imread + warpPerspective"
        want = hashlib.sha256(out.tobytes()).hexdigest()
        broken = bytearray(jpgs[1])
        broken[len(broken) // 2:len(broken) // 2 + 4096] = b"\xff\x00" * 2048      # header intact, scan data destroyed
        try:
            st.wait(st.submit_jpeg([jpgs[0], bytes(broken)], out))                  # decoders may conceal the damage or refuse it
        except p.StitchError as e:
            assert e.code in (-2, -1)
        st.wait(st.submit_jpeg(jpgs, out))
        assert hashlib.sha256(out.tobytes()).hexdigest() == want
    code = ("import sys, json, hashlib, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r + '/tests');"
            "from test_gpu_jpeg import fixtures; from conftest import pkg; p = pkg(); jpgs, H, cwh, dec = fixtures();"
            "whs = [(d.shape[1], d.shape[0]) for d in dec]; out = np.empty((cwh[1], cwh[0], 3), np.uint8);"
            "st = p.Stitcher(whs, cwh, H); st.wait(st.submit_jpeg(jpgs, out));"
            "print(json.dumps({'sha': hashlib.sha256(out.tobytes()).hexdigest(), 'backend': st.jpeg_backend()}))") % (ROOT, ROOT)
    got = []
    for extra in ({"BS_DEBUG_JPEG_NO_GPU_ENTROPY": "1"}, {"BS_DEBUG_JPEG_NO_GPU_ENTROPY": "1", "BS_DEBUG_JPEG_THREADS": "0"}):
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=dict(os.environ, **extra))
        assert r.returncode == 0, r.stderr[-2000:]
        got.append(json.loads(r.stdout.strip().splitlines()[-1]))
    assert "worker threads" in got[0]["backend"] and "worker threads" not in got[1]["backend"] and not got[1]["backend"].startswith("b200")
    assert got[0]["sha"] == got[1]["sha"]


def test_compressed_and_raw_submits_interleave_in_ticket_order():
    """bs_submit_jpeg leaves the host decode stages running and issues the GPU side later; a raw submit, bs_sync or
    a lens-model change right behind it must still see the frames in ticket order."""
    p = pkg()
    jpgs, H, canvas_wh, cv2_dec = fixtures()
    whs = [(d.shape[1], d.shape[0]) for d in cv2_dec]
    decoded = [nvjpeg_pixels(p, j, wh) for j, wh in zip(jpgs, whs)]
    want = oracle.stitch(decoded, H, canvas_wh)
    other = [np.ascontiguousarray(d[::-1]) for d in decoded]            # different pixels for the raw frames
    want_other = oracle.stitch(other, H, canvas_wh)
    outs = [np.full((canvas_wh[1], canvas_wh[0], 3), 99, np.uint8) for _ in range(5)]
    with p.Stitcher(whs, canvas_wh, H) as st:
        t0 = st.submit_jpeg(jpgs, outs[0])
        t1 = st.submit(other, outs[1])                                   # raw frame behind a compressed one, no wait in between
        t2 = st.submit_jpeg(jpgs, outs[2])
        t3 = st.submit_jpeg(jpgs, outs[3])                               # fourth frame of a three-slot pipeline
        for t in (t3, t0, t2, t1):                                       # waits in any order
            st.wait(t)
        assert np.array_equal(outs[0], want) and np.array_equal(outs[2], want) and np.array_equal(outs[3], want)
        assert np.array_equal(outs[1], want_other)
        st.submit_jpeg(jpgs, outs[4])
        st.sync()                                                        # bs_sync alone completes a compressed frame
        assert np.array_equal(outs[4], want)
        K = np.array([[0.9 * whs[0][0], 0, whs[0][0] / 2], [0, 0.9 * whs[0][0], whs[0][1] / 2], [0, 0, 1]])
        dist = np.array([-0.15, 0.03, 0.001, -0.001, 0.0])
        t5 = st.submit_jpeg(jpgs, outs[0])                               # submitted before the model: must not see it
        st.set_undistort(0, K, dist)
        st.wait(t5)
        assert np.array_equal(outs[0], want)
        st.wait(st.submit_jpeg(jpgs, outs[1]))                           # submitted after: sees it
        und = [oracle.undistort(decoded[0], K, dist), decoded[1]]
        assert np.array_equal(outs[1], oracle.stitch(und, H, canvas_wh))
