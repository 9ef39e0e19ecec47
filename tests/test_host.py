"""The C++ host side (image-stitching_b200/host): flag parser, stitch.conf.xml reader, PPM I/O and
the stitch tool's exit codes -- behaviours of /root/reference/lib/CommandLine/opts.cpp:6-106 and
/root/reference/tools/stitch/main.cpp:19-133 (SURVEY.md Appendices C and D).  CPU only, except the
two end-to-end runs of the tool at the bottom."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle
from conftest import ROOT, pkg, workloads

BIN = os.path.join(ROOT, "image-stitching_b200", "bin")
SELFTEST = os.path.join(BIN, "host_selftest")
STITCH = os.path.join(BIN, "stitch")

pytestmark = pytest.mark.skipif(not os.path.exists(SELFTEST), reason="host tools not built (make -C image-stitching_b200/host)")


def parse(*args):
    out = subprocess.run([SELFTEST, "opts", *args], capture_output=True, text=True, check=True).stdout
    return json.loads(out)


def test_defaults_match_reference_struct():
    o = parse()
    assert o == {"valid": True, "verbosity": 0, "interactive": False, "video": False, "delay": 300,
                 "number_of_frames": 30, "board_width": 5, "board_height": 3, "angle": 1, "mode": "chain",
                 "calibrate_config": "", "stitch_config": "stitch.conf.xml", "output": "", "device": 0, "devices": [],
                 "file_paths": []}


def test_flags_and_prefix_quirks():
    o = parse("--s-conf=my.xml", "-i", "-v=2", "--video", "-b=3x4", "--mode=extend", "--delay=5", "--num=9",
              "--angle=7", "--c-conf=c.xml", "a.jpg", "b.jpg")
    assert o["valid"] and o["stitch_config"] == "my.xml" and o["interactive"] and o["verbosity"] == 2
    assert o["video"] and (o["board_width"], o["board_height"]) == (3, 4) and o["mode"] == "extend"
    assert (o["delay"], o["number_of_frames"], o["angle"], o["calibrate_config"]) == (5, 9, 7, "c.xml")
    assert o["file_paths"] == ["a.jpg", "b.jpg"]
    # matching is by prefix: anything starting with -v is the verbosity flag
    assert parse("-verbose=4")["verbosity"] == 4
    # a valued flag without '=' makes the parse fail (stitch then exits 1)
    assert parse("-v")["valid"] is False
    assert parse("--s-conf")["valid"] is False
    assert parse("-b=5")["valid"] is False          # no 'x'
    assert parse("--mode=other")["valid"] is False
    # the usage text's "-c=..." is not a flag: it is swallowed as a path
    assert parse("-c=/x/conf.xml")["file_paths"] == ["-c=/x/conf.xml"]


CONF = """<?xml version="1.0"?>
<opencv_storage>
<video>1</video>
<file_paths>
  /abs/path/1a.jpg "/abs/with space/1b.jpg"</file_paths>
<result_size>
  6085 8353</result_size>
<H>
  <_ type_id="opencv-matrix">
    <rows>3</rows>
    <cols>3</cols>
    <dt>d</dt>
    <data>
      0.97999999999999998 0.02 1790.25 -0.014999999999999999 1.01 12.6
      -3.9999999999999998e-06 3.0000000000000001e-06 1.</data></_>
  <_ type_id="opencv-matrix">
    <rows>3</rows>
    <cols>3</cols>
    <dt>d</dt>
    <data>
      1. 0. 0. 0. 1. 0. 0. 0. 1.</data></_></H>
<cameraMatrix>
  <_ type_id="opencv-matrix"><rows>0</rows><cols>0</cols><dt>d</dt><data></data></_>
  <_ type_id="opencv-matrix"><rows>0</rows><cols>0</cols><dt>u</dt><data></data></_></cameraMatrix>
<distCoeffs>
  <_ type_id="opencv-matrix"><rows>0</rows><cols>0</cols><dt>d</dt><data></data></_>
  <_ type_id="opencv-matrix"><rows>0</rows><cols>0</cols><dt>d</dt><data></data></_></distCoeffs>
</opencv_storage>
"""


def conf(path):
    return json.loads(subprocess.run([SELFTEST, "conf", str(path)], capture_output=True, text=True, check=True).stdout)


def test_reads_opencv_storage_xml(tmp_path):
    p = tmp_path / "stitch.conf.xml"
    p.write_text(CONF)
    c = conf(p)
    assert c["opened"] and c["video"]["text"] == "1"
    assert c["file_paths"]["kind"] == 4 and c["file_paths"]["items"] == ["/abs/path/1a.jpg", "/abs/with space/1b.jpg"]
    assert c["result_size"]["items"] == ["6085", "8353"]
    assert c["H"]["size"] == 2 and c["H"]["items"][0]["data"] == [0.98, 0.02, 1790.25, -0.015, 1.01, 12.6, -4e-06, 3e-06, 1.0]
    assert c["cameraMatrix"]["size"] == 2 and c["cameraMatrix"]["items"][1]["rows"] == 0
    assert c["missing"]["kind"] == 0
    assert conf(tmp_path / "nope.xml")["opened"] is False


def test_node_kinds_follow_opencv(tmp_path):
    # cv::FileStorage reads a one-token text node as a scalar and an empty one as NONE; only <_> children or
    # two or more tokens make a sequence.  The reference's `type() != FileNode::SEQ` checks depend on it.
    p = tmp_path / "c.xml"
    p.write_text('<?xml version="1.0"?>\n<opencv_storage>\n<video>0</video>\n<file_paths>\n  "/a b/1a.jpg"</file_paths>\n'
                 '<result_size>\n  </result_size>\n<H>\n  <_ type_id="opencv-matrix"><rows>3</rows><cols>3</cols><dt>d</dt>'
                 '<data>1. 0. 0. 0. 1. 0. 0. 0. 1.</data></_></H>\n</opencv_storage>\n')
    c = conf(p)
    assert c["file_paths"]["kind"] == 3          # STRING, not SEQ: the reference exits 4 on a one-camera config
    assert c["result_size"]["kind"] == 0
    assert c["H"]["kind"] == 4 and c["H"]["size"] == 1


def test_matches_cv2_filestorage_when_available(tmp_path):
    cv2 = pytest.importorskip("cv2")
    w = workloads().get("cfg3")
    p = str(tmp_path / "w.xml")
    fs = cv2.FileStorage(p, cv2.FILE_STORAGE_WRITE)
    fs.write("video", 1)
    fs.startWriteStruct("file_paths", cv2.FileNode_SEQ)
    for i in range(4):
        fs.write("", f"/data/cam{i}.ppms")
    fs.endWriteStruct()
    fs.startWriteStruct("result_size", cv2.FileNode_SEQ | cv2.FileNode_FLOW)
    fs.write("", 7680)
    fs.write("", 2160)
    fs.endWriteStruct()
    for key, mats in (("H", list(w.H)), ("cameraMatrix", [np.zeros((0, 0))] * 4), ("distCoeffs", [np.zeros((0, 0))] * 4)):
        fs.startWriteStruct(key, cv2.FileNode_SEQ)
        for m in mats:
            fs.write("", m)
        fs.endWriteStruct()
    fs.release()
    c = conf(p)
    assert c["file_paths"]["items"] == [f"/data/cam{i}.ppms" for i in range(4)]
    got = np.array([m["data"] for m in c["H"]["items"]]).reshape(4, 3, 3)
    assert np.array_equal(got, w.H)               # 17 significant digits round-trip exactly


def write_ppm(path, bgr):
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (bgr.shape[1], bgr.shape[0]))
        f.write(np.ascontiguousarray(bgr[:, :, ::-1]).tobytes())


def read_ppm(path):
    data = open(path, "rb").read()
    parts = data.split(b"\n", 3)
    assert parts[0] == b"P6"
    w, h = map(int, parts[1].split())
    return np.frombuffer(parts[3], dtype=np.uint8).reshape(h, w, 3)[:, :, ::-1]


def test_ppm_round_trip(tmp_path):
    img = np.random.default_rng(1).integers(0, 256, (13, 17, 3), dtype=np.uint8)
    write_ppm(tmp_path / "a.ppm", img)
    r = subprocess.run([SELFTEST, "ppm", str(tmp_path / "a.ppm"), str(tmp_path / "b.ppm")], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split() == ["17", "13"]
    assert np.array_equal(read_ppm(tmp_path / "b.ppm"), img)


def make_config(path, video, paths, result_wh, Hs, Ks=None, dists=None):
    def mat(m):
        m = np.atleast_2d(np.asarray(m, float))
        return ('  <_ type_id="opencv-matrix">\n    <rows>%d</rows>\n    <cols>%d</cols>\n    <dt>d</dt>\n    <data>\n      ' % m.shape
                + " ".join(repr(float(v)) for v in m.reshape(-1)) + "</data></_>\n")
    mats = "".join(
        '  <_ type_id="opencv-matrix">\n    <rows>3</rows>\n    <cols>3</cols>\n    <dt>d</dt>\n    <data>\n      '
        + " ".join(repr(float(v)) for v in np.asarray(H).reshape(-1)) + "</data></_>\n" for H in Hs)
    empty = "".join('  <_ type_id="opencv-matrix"><rows>0</rows><cols>0</cols><dt>d</dt><data></data></_>\n' for _ in Hs)
    kmats = "".join(mat(K) for K in Ks) if Ks is not None else empty
    dmats = "".join(mat(d) for d in dists) if dists is not None else empty
    path.write_text('<?xml version="1.0"?>\n<opencv_storage>\n<video>%d</video>\n<file_paths>\n  %s</file_paths>\n'
                    '<result_size>\n  %d %d</result_size>\n<H>\n%s</H>\n<cameraMatrix>\n%s</cameraMatrix>\n<distCoeffs>\n%s</distCoeffs>\n'
                    '</opencv_storage>\n' % (video, " ".join(str(p) for p in paths), result_wh[0], result_wh[1], mats, kmats, dmats))


def run_stitch(cwd, *args):
    return subprocess.run([STITCH, *args], cwd=cwd, capture_output=True, text=True)


def test_exit_codes_without_touching_the_gpu(tmp_path):
    assert run_stitch(tmp_path, "-v").returncode == 1                                  # malformed flag
    r = run_stitch(tmp_path, "--s-conf=missing.xml")
    assert r.returncode == 2 and "Failed to open configuration file missing.xml" in r.stdout
    one = tmp_path / "one.xml"                                                          # one path: not a sequence
    one.write_text('<?xml version="1.0"?>\n<opencv_storage>\n<video>0</video>\n<file_paths>\n  a.ppm</file_paths>\n</opencv_storage>\n')
    r = run_stitch(tmp_path, f"--s-conf={one}")
    assert r.returncode == 4 and "file_paths is not a sequence!" in r.stdout
    noh = tmp_path / "noh.xml"
    noh.write_text('<?xml version="1.0"?>\n<opencv_storage>\n<video>0</video>\n<file_paths>\n  a.ppm b.ppm</file_paths>\n'
                   '<result_size>\n  8 8</result_size>\n</opencv_storage>\n')
    r = run_stitch(tmp_path, f"--s-conf={noh}")
    assert r.returncode == 4 and "H is not a sequence!" in r.stdout
    cfgv = tmp_path / "v.xml"                                                           # video inputs that do not exist
    make_config(cfgv, 1, ["/nonexistent/a.ppms", "/nonexistent/b.ppms"], (64, 32), [np.eye(3), np.eye(3)])
    r = run_stitch(tmp_path, f"--s-conf={cfgv}")
    assert r.returncode == 5 and "Failed to open file /nonexistent/a.ppms!" in r.stdout


@pytest.mark.gpu
def test_stitch_tool_image_mode_matches_oracle(tmp_path):
    rng = np.random.default_rng(3)
    imgs = [rng.integers(0, 256, (120, 160, 3), dtype=np.uint8) for _ in range(2)]
    Hs = [np.array([[1.05, 0.02, 4.0], [-0.01, 1.02, 3.0], [1e-5, 2e-5, 1.0]]),
          np.array([[0.97, -0.03, 150.0], [0.02, 1.01, 6.0], [-1e-5, 0, 1.0]])]
    for i, im in enumerate(imgs):
        write_ppm(tmp_path / f"{i}.ppm", im)
    cfg = tmp_path / "stitch.conf.xml"
    make_config(cfg, 0, [tmp_path / "0.ppm", tmp_path / "1.ppm"], (320, 140), Hs)
    r = run_stitch(tmp_path, "--out=final.ppm")            # default --s-conf=stitch.conf.xml in the CWD
    assert r.returncode == 0, r.stdout + r.stderr
    assert len([ln for ln in r.stdout.split() if ln.replace(".", "").replace("e-", "").isdigit()]) == 2   # one ms line per warp
    assert np.array_equal(read_ppm(tmp_path / "final.ppm"), oracle.stitch(imgs, Hs, (320, 140)))


@pytest.mark.gpu
def test_stitch_tool_video_mode_matches_oracle(tmp_path):
    rng = np.random.default_rng(4)
    n_frames = 5                                            # + the first frame, which the reference drops
    Hs = [np.array([[1.02, 0.0, 2.0], [0.0, 1.02, 1.0], [0, 0, 1.0]]), np.array([[1.0, 0.01, 90.0], [0.0, 1.0, 2.0], [0, 1e-5, 1.0]])]
    frames = [[rng.integers(0, 256, (64, 96, 3), dtype=np.uint8) for _ in range(n_frames + 1)] for _ in range(2)]
    for c in range(2):
        with open(tmp_path / f"cam{c}.ppms", "wb") as f:
            for im in frames[c]:
                f.write(b"P6\n96 64\n255\n" + np.ascontiguousarray(im[:, :, ::-1]).tobytes())
    cfg = tmp_path / "v.xml"
    make_config(cfg, 1, [tmp_path / "cam0.ppms", tmp_path / "cam1.ppms"], (200, 70), Hs)
    r = run_stitch(tmp_path, f"--s-conf={cfg}", "--out=pano_%03ld.ppm")
    assert r.returncode == 0, r.stdout + r.stderr
    for k in range(n_frames):
        want = oracle.stitch([frames[0][k + 1], frames[1][k + 1]], Hs, (200, 70))
        assert np.array_equal(read_ppm(tmp_path / f"pano_{k:03d}.ppm"), want), f"frame {k}"
    assert not os.path.exists(tmp_path / f"pano_{n_frames:03d}.ppm")


@pytest.mark.gpu
def test_stitch_tool_video_mode_with_lens_models(tmp_path):
    # cameraMatrix / distCoeffs as calibrate writes them (8 coefficients, column vector): undistort then warp
    rng = np.random.default_rng(6)
    Hs = [np.array([[1.02, 0.0, 2.0], [0.0, 1.02, 1.0], [0, 0, 1.0]]), np.array([[1.0, 0.01, 90.0], [0.0, 1.0, 2.0], [0, 1e-5, 1.0]])]
    Ks = [np.array([[90.0, 0, 48.5], [0, 91.0, 31.2], [0, 0, 1]]), np.array([[85.0, 0, 47.0], [0, 85.0, 33.0], [0, 0, 1]])]
    dists = [np.array([-0.2, 0.05, 0.001, -0.002, 0.01, 0, 0, 0]), np.zeros(8)]
    frames = [[rng.integers(0, 256, (64, 96, 3), dtype=np.uint8) for _ in range(4)] for _ in range(2)]
    for c in range(2):
        with open(tmp_path / f"cam{c}.ppms", "wb") as f:
            for im in frames[c]:
                f.write(b"P6\n96 64\n255\n" + np.ascontiguousarray(im[:, :, ::-1]).tobytes())
    cfg = tmp_path / "v.xml"
    make_config(cfg, 1, [tmp_path / "cam0.ppms", tmp_path / "cam1.ppms"], (200, 70), Hs, Ks, [d.reshape(-1, 1) for d in dists])
    r = run_stitch(tmp_path, f"--s-conf={cfg}", "--out=u_%03ld.ppm")
    assert r.returncode == 0, r.stdout + r.stderr
    for k in range(3):
        und = [oracle.undistort(frames[c][k + 1], Ks[c], dists[c]) for c in range(2)]
        assert np.array_equal(read_ppm(tmp_path / f"u_{k:03d}.ppm"), oracle.stitch(und, Hs, (200, 70))), f"frame {k}"


def test_devices_flag_forms():
    out = parse("--devices=0-2,5", "--device=3")
    assert out["devices"] == [0, 1, 2, 5] and out["device"] == 3
    assert parse("--devices=1")["devices"] == [1]
    assert parse("--devices=3-1")["valid"] is False and parse("--devices=")["valid"] is False


@pytest.mark.gpu
def test_stitch_tool_video_mode_over_a_pool_of_replicas(tmp_path):
    """--devices=0,0: two replicas (on one GPU here; any list of GPUs works), frame f on replica f mod 2,
    panoramas written in frame order."""
    rng = np.random.default_rng(9)
    n_frames = 9
    Hs = [np.array([[1.02, 0.0, 2.0], [0.0, 1.02, 1.0], [0, 0, 1.0]]), np.array([[1.0, 0.01, 90.0], [0.0, 1.0, 2.0], [0, 1e-5, 1.0]])]
    frames = [[rng.integers(0, 256, (64, 96, 3), dtype=np.uint8) for _ in range(n_frames + 1)] for _ in range(2)]
    for c in range(2):
        with open(tmp_path / f"cam{c}.ppms", "wb") as f:
            for im in frames[c]:
                f.write(b"P6\n96 64\n255\n" + np.ascontiguousarray(im[:, :, ::-1]).tobytes())
    cfg = tmp_path / "v.xml"
    make_config(cfg, 1, [tmp_path / "cam0.ppms", tmp_path / "cam1.ppms"], (200, 70), Hs)
    r = run_stitch(tmp_path, f"--s-conf={cfg}", "--out=p_%03ld.ppm", "--devices=0,0", "-v=1")
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"{n_frames} frames on 2 GPU(s)" in r.stdout
    for k in range(n_frames):
        want = oracle.stitch([frames[0][k + 1], frames[1][k + 1]], Hs, (200, 70))
        assert np.array_equal(read_ppm(tmp_path / f"p_{k:03d}.ppm"), want), f"frame {k}"


@pytest.mark.gpu
def test_stitch_tool_image_mode_with_jpeg_photos(tmp_path):
    """JPEG in (imread, tools/stitch/main.cpp:107,114) and JPEG out (imwrite, :125) through the tool: the bitstreams are
    decoded on the GPU by bs_submit_jpeg, so the panorama equals the oracle applied to the pixels nvJPEG decodes; with
    a .jpg output only the encoded panorama comes back."""
    import shutil
    from conftest import GOLDEN
    p = pkg()
    z = np.load(os.path.join(GOLDEN, "real_crop.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    H, canvas_wh = z["H"], tuple(meta["canvas_wh"])
    paths = []
    for i in range(2):
        shutil.copy(os.path.join(GOLDEN, f"jpeg_src{i}.jpg"), tmp_path / f"{i}.jpg")
        paths.append(tmp_path / f"{i}.jpg")
    decoded = []
    for pth in paths:                                   # the frames as nvJPEG decodes them (identity warp, raw output)
        bits = open(pth, "rb").read()
        dec = np.load(os.path.join(GOLDEN, "jpeg_cv2_decode.npz"))[f"src{len(decoded)}_cv2"]
        wh = (dec.shape[1], dec.shape[0])
        with p.Stitcher([wh], wh, [np.eye(3)]) as st:
            out = np.empty((wh[1], wh[0], 3), np.uint8)
            st.wait(st.submit_jpeg([bits], out))
        decoded.append(out)
    cfg = tmp_path / "stitch.conf.xml"
    make_config(cfg, 0, paths, canvas_wh, list(H))
    r = run_stitch(tmp_path, "--out=final.ppm")
    assert r.returncode == 0, r.stdout + r.stderr
    want = oracle.stitch(decoded, H, canvas_wh)
    assert np.array_equal(read_ppm(tmp_path / "final.ppm"), want)
    r = run_stitch(tmp_path)                            # default output: final.jpg, as the reference writes
    assert r.returncode == 0, r.stdout + r.stderr
    bits = open(tmp_path / "final.jpg", "rb").read()
    assert bits[:2] == b"\xff\xd8" and bits[-2:] == b"\xff\xd9"
    cv2 = pytest.importorskip("cv2")
    back = cv2.imdecode(np.frombuffer(bits, np.uint8), cv2.IMREAD_COLOR)
    assert back.shape == want.shape
    mse = float(((back.astype(float) - want.astype(float)) ** 2).mean())
    assert 10 * np.log10(255.0 ** 2 / max(mse, 1e-9)) > 30.0
    (tmp_path / "bad.jpg").write_bytes(b"\xff\xd8 not really a jpeg")
    make_config(cfg, 0, [paths[0], tmp_path / "bad.jpg"], canvas_wh, list(H))
    r = run_stitch(tmp_path)
    assert r.returncode == 5 and "bad.jpg" in r.stdout


@pytest.mark.gpu
def test_stitch_tool_video_mode_from_motion_jpeg(tmp_path):
    """Compressed video in (cv::VideoCapture at tools/stitch/main.cpp:127-147): Motion-JPEG streams, every frame handed to
    bs_submit_jpeg as a bitstream; lens models set; panoramas out raw and, in a second run, as JPEG files."""
    cv2 = pytest.importorskip("cv2")
    p = pkg()
    rng = np.random.default_rng(9)
    sw, sh, n_frames = 320, 192, 4
    Hs = [np.array([[1.01, 0.0, 3.0], [0.0, 1.01, 2.0], [0, 0, 1.0]]), np.array([[1.0, 0.01, 280.0], [0.0, 1.0, 4.0], [0, 1e-5, 1.0]])]
    Ks = [np.array([[300.0, 0, 160.0], [0, 300.0, 96.0], [0, 0, 1.0]])] * 2
    dists = [np.array([[-0.12, 0.02, 0.001, -0.001, 0.0, 0.0, 0.0, 0.0]]).T, np.zeros((8, 1))]
    bits = [[], []]
    for c in range(2):
        base = cv2.GaussianBlur(rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8), (0, 0), 3)
        with open(tmp_path / f"cam{c}.mjpeg", "wb") as f:
            for k in range(n_frames + 1):                    # + the first frame, which the reference drops
                img = np.roll(base, 7 * k, axis=1)
                enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 92] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, 5] if k % 2 else []))[1].tobytes()
                bits[c].append(enc)
                f.write(enc + (b"\x00\x00\x00" if k % 2 else b""))      # some streams pad between frames
    def decoded(b):
        with p.Stitcher([(sw, sh)], (sw, sh), [np.eye(3)]) as st:
            out = np.empty((sh, sw, 3), np.uint8)
            st.wait(st.submit_jpeg([b], out))
        return out
    cfg = tmp_path / "v.xml"
    make_config(cfg, 1, [tmp_path / "cam0.mjpeg", tmp_path / "cam1.mjpeg"], (620, 200), Hs, Ks, dists)
    r = run_stitch(tmp_path, f"--s-conf={cfg}", "--out=pano_%03ld.ppm")
    assert r.returncode == 0, r.stdout + r.stderr
    wants = []
    for k in range(n_frames):
        f0 = oracle.undistort(decoded(bits[0][k + 1]), Ks[0], dists[0].reshape(-1))
        wants.append(oracle.stitch([f0, decoded(bits[1][k + 1])], Hs, (620, 200)))
        assert np.array_equal(read_ppm(tmp_path / f"pano_{k:03d}.ppm"), wants[k]), f"frame {k}"
    assert not os.path.exists(tmp_path / f"pano_{n_frames:03d}.ppm")
    r = run_stitch(tmp_path, f"--s-conf={cfg}", "--out=enc_%03ld.jpg", "--frames=3")
    assert r.returncode == 0, r.stdout + r.stderr
    for k in range(3):
        back = cv2.imdecode(np.frombuffer(open(tmp_path / f"enc_{k:03d}.jpg", "rb").read(), np.uint8), cv2.IMREAD_COLOR)
        mse = float(((back.astype(float) - wants[k].astype(float)) ** 2).mean())
        assert back.shape == wants[k].shape and 10 * np.log10(255.0 ** 2 / max(mse, 1e-9)) > 30.0
    assert not os.path.exists(tmp_path / "enc_003.jpg")


def test_motion_jpeg_stream_is_split_at_frame_boundaries(tmp_path):
    """The stitch tool's view of a .mjpeg file (host/image_io.cpp JpegStream + jpeg_size): frames with and without
    restart markers, stuffed 0xFF bytes in the entropy data, an APP segment that contains 0xFFD9, padding between
    frames, a truncated last frame."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(2)
    frames = []
    for k, (w, h) in enumerate([(160, 96), (333, 200), (64, 48)]):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)          # noise: plenty of 0xFF 0x00 in the scan
        args = [cv2.IMWRITE_JPEG_QUALITY, 95] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, 3] if k == 1 else [])
        bits = cv2.imencode(".jpg", img, args)[1].tobytes()
        assert b"\xff\x00" in bits
        if k == 2:                                                       # an application segment holding an EOI look-alike
            bits = bits[:2] + b"\xff\xe1\x00\x08ab\xff\xd9cd" + bits[2:]
        frames.append((bits, w, h))
    path = tmp_path / "s.mjpeg"
    with open(path, "wb") as f:
        off, want = 0, []
        for k, (bits, w, h) in enumerate(frames):
            pad = b"\x00" * (3 * k)
            f.write(pad + bits)
            if k == 0:
                first = len(pad)
            want.append((off + len(pad) - first, len(bits), w, h))
            off += len(pad) + len(bits)
        f.write(frames[0][0][:500])                                      # truncated frame at the end: not reported
    r = subprocess.run([SELFTEST, "mjpeg", str(path)], capture_output=True, text=True)
    assert r.returncode == 0
    got = [tuple(int(v) for v in ln.split()) for ln in r.stdout.strip().splitlines()]
    assert got == want
