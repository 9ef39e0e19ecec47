"""Pin the CPU oracle (oracle/) against the golden vectors made by the reference's own OpenCV
path (tests/golden/make_golden.py) and against SURVEY.md Appendix B's known answers.
CPU only."""
import hashlib
import os

import numpy as np
import pytest

import oracle
from conftest import ROOT, workloads

# SURVEY.md Appendix B.3 (recorded when the reference was surveyed) -- must agree with the
# regenerated tests/golden/known_answers.json
SURVEY_B3 = {
    ("cfg2", "noise", 0): "b6a9f26e956a9dd1409ea1177b57f2e3031b5dbe75763551faf63b8ee459aa32",
    ("cfg2", "noise", 1): "792b56abb5943bf581954efdd4a98bccd5b6926d4e1aa50e44be75c48960f075",
    ("cfg2", "smooth", 0): "7de58817b7e7534a34d30eb41859ec09df3bb014b3b34768417c9d3fb3940edf",
    ("cfg2", "smooth", 1): "15279b9547845c688eca213b7759b49fc23e97abdcf7955286ddd7762c80c7f8",
    ("cfg3", "noise", 0): "2cda6f9b71766cffd1019f3a6ab3f200eee421ed1a11120375883ac96694dae6",
    ("cfg3", "noise", 1): "47c9179dc8bdf7a53189a44cc4bdbcce24a3b8d08f4f224e102175119f7a4e68",
    ("cfg3", "smooth", 0): "51246a29efa4b0fd94510295f3961279fad850b8fe4c437c7e476d892012e0ba",
    ("cfg3", "smooth", 1): "92b5edf1759b6164755988a0fd4505ed34f1f90a037c5139f471d65aa01addc3",
    ("cfg5", "noise", 0): "48619d9dd2d455f5ffd5927c5a94fd0a452769e9e3c544ce02b8da1db0d5b9a7",
}
CFG1_PHOTOS_CV4 = "dc70070aad9d1807333af6ab27a70a68dc96faa122e5f17c9d4318e6b003de20"
CFG1_PHOTOS_LEGACY = "c301099db99d7a6f1f178ae0a74ab8894e032b754d779e6f7e372c3e968eb122"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_generator_self_check():
    # noise, camera 0, frame 0, row 0, first 8 pixels
    f = oracle.gen_frame(1920, 1080, 0, 0, oracle.NOISE)
    assert f[0, :8].reshape(-1).tolist() == [0, 201, 69, 32, 31, 10, 63, 101, 57, 238, 106, 121, 80,
                                             124, 51, 242, 47, 53, 210, 226, 170, 231, 121, 130]
    assert sha(f).startswith("e3e753f85d4222ec")


def test_known_answers_file_matches_survey(known_answers):
    got = {(c["config"], c["content"], c["frame"]): c["canvas_sha256"] for c in known_answers["cases"]}
    for k, v in SURVEY_B3.items():
        assert got[k] == v
    assert known_answers["cfg1_photos"]["canvas_sha256"] == CFG1_PHOTOS_CV4


def test_small_cases_bit_exact(small_cases):
    z, meta = small_cases
    for m in meta:
        k = m["case"]
        srcs = [z[f"c{k}_src{i}"] for i in range(m["n_cams"])]
        Hs = z[f"c{k}_H"]
        got = oracle.stitch(srcs, Hs, tuple(m["canvas_wh"]))
        assert np.array_equal(got, z[f"c{k}_exp_zero"]), f"case {k} (zero canvas)"
        canvas = z[f"c{k}_prior"].copy()
        for s, H in zip(srcs, Hs):
            oracle.warp_perspective(s, H, canvas)
        assert np.array_equal(canvas, z[f"c{k}_exp_prior"]), f"case {k} (prior canvas kept)"


def test_legacy_rule_is_subset_of_cv4(small_cases):
    # Appendix A.4: the 2.4/3.x rule writes a subset of the pixels the 4.13 rule writes and
    # agrees on every pixel it does write.
    z, meta = small_cases
    for m in meta:
        k = m["case"]
        srcs = [z[f"c{k}_src{i}"] for i in range(m["n_cams"])][:1]
        Hs = z[f"c{k}_H"][:1]
        dw, dh = m["canvas_wh"]
        marker = np.full((dh, dw, 3), 7, np.uint8)
        a = marker.copy()
        b = marker.copy()
        oracle.warp_perspective(srcs[0], Hs[0], a, oracle.BORDER_CV4)
        oracle.warp_perspective(srcs[0], Hs[0], b, oracle.BORDER_LEGACY)
        wrote_b = (b != marker).any(axis=2)
        assert np.array_equal(a[wrote_b], b[wrote_b])


def test_undistort_cases_bit_exact(undistort_cases):
    z, meta = undistort_cases
    for m in meta:
        k = m["case"]
        got = oracle.undistort(z[f"u{k}_src"], z[f"u{k}_K"], z[f"u{k}_dist"])
        assert np.array_equal(got, z[f"u{k}_exp"]), f"undistort case {k}"


def test_undistort_identity_is_copy():
    # zero distortion -> exact copy
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, (40, 70, 3), dtype=np.uint8)
    K = np.array([[60.0, 0, 35.0], [0, 60.0, 20.0], [0, 0, 1]])
    assert np.array_equal(oracle.undistort(src, K, np.zeros(5)), src)


@pytest.mark.parametrize("cfg,content,frame", [("cfg2", "noise", 0), ("cfg2", "smooth", 1),
                                               ("cfg3", "noise", 0), ("cfg3", "smooth", 0),
                                               ("cfg5", "noise", 0), ("cfg1", "noise", 0)])
def test_full_size_known_answers(known_answers, cfg, content, frame):
    want = [c for c in known_answers["cases"]
            if (c["config"], c["content"], c["frame"]) == (cfg, content, frame)][0]
    w = workloads().get(cfg)
    kind = oracle.NOISE if content == "noise" else oracle.SMOOTH
    srcs = [oracle.gen_frame(w.src_wh[0], w.src_wh[1], i, frame, kind) for i in range(w.n_cams)]
    assert sha(srcs[0]) == want["src0_sha256"]
    assert sha(oracle.stitch(srcs, w.H, w.canvas_wh)) == want["canvas_sha256"]


@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg1"])
def test_algorithmic_bytes(cfg):
    # W*H*3 + 3*sum_i U_i
    w = workloads().get(cfg)
    total, tex, wins = oracle.count_texels([w.src_wh] * w.n_cams, w.canvas_wh, w.H)
    if cfg != "cfg1":       # the synthetic panoramas are fully covered; the two photos cover 56 % of their canvas
        assert sum(wins) == w.canvas_wh[0] * w.canvas_wh[1]
    assert w.canvas_bytes + 3 * total == w.algorithmic_bytes


@pytest.mark.skipif(not os.path.isdir("/root/reference/test/Inputs"),
                    reason="reference photos only exist in the build container")
def test_cfg1_photos_both_border_rules():
    cv2 = pytest.importorskip("cv2")
    w = workloads().get("cfg1")
    imgs = [cv2.imread(f"/root/reference/test/Inputs/{n}") for n in ("1a.jpg", "1b.jpg")]
    assert sha(oracle.stitch(imgs, w.H, w.canvas_wh, oracle.BORDER_CV4)) == CFG1_PHOTOS_CV4
    assert sha(oracle.stitch(imgs, w.H, w.canvas_wh, oracle.BORDER_LEGACY)) == CFG1_PHOTOS_LEGACY


def test_invert_matches_cv2_when_available():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for _ in range(500):
        H = np.eye(3) + rng.normal(size=(3, 3)) * np.array([[0.1, 0.1, 100], [0.1, 0.1, 100],
                                                            [1e-4, 1e-4, 0.01]])
        M, ok = oracle.invert3x3(H)
        assert np.array_equal(M, cv2.invert(H)[1])


def test_border_constant_golden_cases(border_constant_cases):
    """warpPerspective with BORDER_CONSTANT 0 (lib/Calibrate/utils.cpp:390), against
    cv2 outputs; the destination's previous contents must not matter."""
    z, meta = border_constant_cases
    for m in meta:
        k = m["case"]
        dw, dh = m["canvas_wh"]
        out = np.full((dh, dw, 3), 77, np.uint8)
        oracle.warp_perspective(np.ascontiguousarray(z[f"b{k}_src"]), z[f"b{k}_H"], out, oracle.BORDER_CONSTANT)
        assert np.array_equal(out, z[f"b{k}_exp"]), f"case {k}"


def test_real_crop_fixture_matches_oracle():
    """tests/golden/real_crop.npz: crops of the reference's photos warped by cv2 (make_real_crop.py)."""
    import json
    z = np.load(os.path.join(ROOT, "tests", "golden", "real_crop.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    got = oracle.stitch([np.ascontiguousarray(z["src0"]), np.ascontiguousarray(z["src1"])], z["H"], tuple(meta["canvas_wh"]))
    assert sha(got) == meta["canvas_sha256"]
