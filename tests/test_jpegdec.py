"""The JPEG decoder whose entropy stage runs on the GPU (image-stitching_b200/csrc/stitch_jpegdec.cuh, row f2 of
SURVEY.md section 8: imread / VideoCapture at /root/reference/tools/stitch/main.cpp:107,114,147).

Its arithmetic is libjpeg's (islow IDCT, fancy h2v2 up-sampling, fixed-point YCbCr -> RGB), so the pin is the
reference's own decoder: cv2.imdecode (libjpeg-turbo).  The CPU tests build the decoder's functions for the host
(tests/aux/jpegdec_host.cpp compiles the same header) and compare with cv2.imdecode bit for bit; the GPU tests run
the kernels through bs_submit_jpeg and require (a) the same pixels as cv2.imdecode and (b) a panorama equal to the
oracle applied to those pixels -- i.e. JPEG in gives exactly what the reference's imread + warpPerspective gives."""
import os
import subprocess

import numpy as np
import pytest

import oracle
from conftest import ROOT, pkg

cv2 = pytest.importorskip("cv2")


def encode(img, quality=90, rst=4, sampling=None):
    args = [cv2.IMWRITE_JPEG_QUALITY, quality] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, rst] if rst else [])
    if sampling is not None:
        args += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sampling]
    ok, enc = cv2.imencode(".jpg", img, args)
    assert ok
    return enc.tobytes()


def pictures():
    rng = np.random.default_rng(11)
    smooth = cv2.GaussianBlur(rng.integers(0, 256, (200, 333, 3), dtype=np.uint8), (0, 0), 2)
    noise = rng.integers(0, 256, (97, 131, 3), dtype=np.uint8)
    ramp = np.dstack([np.tile(np.arange(256, dtype=np.uint8), (64, 1))] * 3)
    ramp[:, :, 1] = ramp[::-1, ::-1, 1]
    return {"smooth": smooth, "noise": noise, "ramp": ramp, "tiny": noise[:9, :17].copy(), "mcu": noise[:32, :48].copy()}


@pytest.fixture(scope="module")
def host_decoder(tmp_path_factory):
    exe = tmp_path_factory.mktemp("jpegdec") / "jpegdec_host"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "aux", "jpegdec_host.cpp")])
    return str(exe)


def host_decode(exe, bits, tmp_path, chunks=0):
    """chunks = T chunks per interval of the self-synchronising decoder the kernels run (files without restart markers
    are one interval); 0 or 1: one chunk, i.e. the sequential decode."""
    src, dst = tmp_path / "in.jpg", tmp_path / "out.bgr"
    src.write_bytes(bits)
    r = subprocess.run([exe, str(src), str(dst)] + ([str(chunks)] if chunks else []), capture_output=True, text=True)
    if r.returncode != 0:
        return r.returncode, None
    w, h = [int(v) for v in r.stdout.split()[:2]]
    return 0, np.fromfile(dst, dtype=np.uint8).reshape(h, w, 3)


@pytest.mark.parametrize("name", ["smooth", "noise", "ramp", "tiny", "mcu"])
@pytest.mark.parametrize("quality,rst", [(90, 4), (50, 1), (100, 7), (75, 1000)])
def test_host_build_of_the_decoder_equals_cv2(host_decoder, tmp_path, name, quality, rst):
    img = pictures()[name]
    bits = encode(img, quality, rst)
    rc, got = host_decode(host_decoder, bits, tmp_path)
    assert rc == 0
    assert np.array_equal(got, cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR))


@pytest.mark.parametrize("chunks", [1, 3, 16, 200])
@pytest.mark.parametrize("rst", [0, 6, 1000])
def test_chunked_decoder_equals_cv2(host_decoder, tmp_path, chunks, rst):
    """The self-synchronising form: chunks start from guesses and converge to the sequential decode, with and without
    restart markers, from one chunk (sequential) to chunks too short to synchronise inside themselves."""
    for name in ("smooth", "noise", "tiny"):
        img = pictures()[name]
        args = [cv2.IMWRITE_JPEG_QUALITY, 85] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, rst] if rst else [])
        bits = cv2.imencode(".jpg", img, args)[1].tobytes()
        rc, got = host_decode(host_decoder, bits, tmp_path, chunks)
        assert rc == 0, name
        assert np.array_equal(got, cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)), name


def test_444_and_refusals(host_decoder, tmp_path):
    img = pictures()["smooth"]
    try:
        bits = encode(img, 92, 3, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444)
    except AttributeError:
        pytest.skip("this cv2 cannot choose the sampling factor")
    rc, got = host_decode(host_decoder, bits, tmp_path)
    assert rc == 0 and np.array_equal(got, cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR))
    # progressive, grey: not for this decoder (bs_submit_jpeg then uses nvJPEG)
    prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1, cv2.IMWRITE_JPEG_RST_INTERVAL, 4])[1].tobytes()
    grey = cv2.imencode(".jpg", img[:, :, 0], [cv2.IMWRITE_JPEG_RST_INTERVAL, 4])[1].tobytes()
    for bits in (prog, grey, b"\xff\xd8 garbage"):
        assert host_decode(host_decoder, bits, tmp_path, 8)[0] == 2


@pytest.mark.gpu
def test_gpu_decoder_gives_cv2_pixels_and_the_reference_panorama():
    p = pkg()
    rng = np.random.default_rng(3)
    sw, sh = 640, 360
    frames = [cv2.GaussianBlur(rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8), (0, 0), 1.5 + i) for i in range(3)]
    Hs = [np.array([[1.01, 0.004 * i, 5.0 + 560 * i], [-0.003, 0.99, 3.0 + i], [1e-6 * (i - 1), 0, 1.0]]) for i in range(3)]
    canvas = (1800, 380)
    for rst in (40, 3, 2000, 0):                   # one MCU row, tiny intervals, one interval for the whole frame, no markers
        bits = [encode(f, 88, rst) for f in frames]
        ref = [cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR) for b in bits]
        with p.Stitcher([(sw, sh)], (sw, sh), [np.eye(3)]) as st:      # identity warp: the decoded frame itself
            out = np.empty((sh, sw, 3), np.uint8)
            st.wait(st.submit_jpeg([bits[0]], out))
            assert st.jpeg_backend().startswith("b200")
            assert np.array_equal(out, ref[0]), f"rst {rst}: {(out != ref[0]).any(axis=2).sum()} pixels differ from cv2.imdecode"
        want = oracle.stitch(ref, Hs, canvas)      # what the reference computes: imread, then warpPerspective per camera
        with p.Stitcher([(sw, sh)] * 3, canvas, Hs) as st:
            outs = [np.empty((canvas[1], canvas[0], 3), np.uint8) for _ in range(5)]
            tickets = [st.submit_jpeg(bits, o) for o in outs]
            for t in tickets:
                st.wait(t)
        for o in outs:
            assert np.array_equal(o, want)


@pytest.mark.gpu
def test_gpu_decoder_odd_sizes_444_and_mixed_streams():
    p = pkg()
    rng = np.random.default_rng(4)
    for (w, h) in ((333, 200), (131, 97), (48, 32)):
        img = cv2.GaussianBlur(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), (0, 0), 1.2)
        variants = [encode(img, 90, 5), encode(img, 60, 1)]
        try:
            variants.append(encode(img, 95, 2, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444))
        except AttributeError:
            pass
        plain = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 90])[1].tobytes()      # no restart markers: one interval
        prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])[1].tobytes()    # progressive: nvJPEG
        with p.Stitcher([(w, h)], (w, h), [np.eye(3)]) as st:
            out = np.empty((h, w, 3), np.uint8)
            for bits in variants + [plain, prog] + variants[:1]:
                st.wait(st.submit_jpeg([bits], out))
                ref = cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)
                if bits is prog:
                    assert not st.jpeg_backend().startswith("b200")
                    mse = float(((out.astype(float) - ref.astype(float)) ** 2).mean())      # nvJPEG: another decoder
                    assert 10 * np.log10(255.0 ** 2 / max(mse, 1e-9)) > 30.0
                else:
                    assert st.jpeg_backend().startswith("b200")
                    assert np.array_equal(out, ref)


@pytest.mark.gpu
def test_gpu_decoder_on_streams_that_do_not_resynchronise():
    """Flat pictures are periodic bit patterns: chunks can follow them at the wrong phase for ever, the rounds then move
    one chunk at a time and the one-thread pass finishes the interval.  Noise at quality 100 has very long blocks.
    Every picture must still come out as cv2.imdecode decodes it."""
    p = pkg()
    rng = np.random.default_rng(8)
    w, h = 1280, 720
    flat = np.zeros((h, w, 3), np.uint8)
    flat[:] = (40, 90, 200)
    disc = flat.copy()
    cv2.circle(disc, (w // 2, h // 2), 200, (250, 20, 60), -1)
    stripes = np.dstack([np.tile(np.linspace(0, 255, w).astype(np.uint8), (h, 1))] * 3)
    stripes[::7] = 255 - stripes[::7]
    noise = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    cases = [("flat", flat, 90, 0), ("disc", disc, 90, 0), ("disc, one MCU row per interval", disc, 90, 80), ("stripes", stripes, 75, 0),
             ("noise q100", noise, 100, 0), ("noise q30", noise, 30, 40)]
    with p.Stitcher([(w, h)], (w, h), [np.eye(3)]) as st:
        out = np.empty((h, w, 3), np.uint8)
        for name, img, q, rst in cases:
            bits = encode(img, q, rst)
            st.wait(st.submit_jpeg([bits], out))
            assert st.jpeg_backend().startswith("b200"), name
            assert np.array_equal(out, cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)), name
