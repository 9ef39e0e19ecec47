"""Row f1 of SURVEY.md section 8: cv::undistort in front of the warp (tools/stitch/main.cpp:148-149),
on the GPU, bit-exact against the OpenCV goldens and against the CPU oracle."""
import numpy as np
import pytest

import oracle
from conftest import pkg, workloads

pytestmark = pytest.mark.gpu


def test_undistort_golden_cases(undistort_cases):
    p = pkg()
    z, meta = undistort_cases
    for m in meta:
        k = m["case"]
        got = p.undistort(np.ascontiguousarray(z[f"u{k}_src"]), z[f"u{k}_K"], z[f"u{k}_dist"])
        want = z[f"u{k}_exp"]
        assert np.array_equal(got, want), f"undistort case {k}: {(got != want).any(axis=2).sum()} pixels differ"


@pytest.mark.parametrize("wh", [(1920, 1080), (3840, 2160), (4097, 31), (333, 77)])
def test_undistort_full_size_vs_oracle(wh):
    # 3840 wide: OpenCV's stripes are one row high; 333 wide: 12 rows; 4097 wide: one row, > 4096
    p = pkg()
    w, h = wh
    src = oracle.gen_frame(w, h, 2, 1, oracle.NOISE)
    K = np.array([[0.9 * w, 0, w / 2 + 3.3], [0, 0.92 * w, h / 2 - 1.7], [0, 0, 1]])
    for dist in ([-0.25, 0.08, 0.001, 0.0005, 0.01, 0.02, -0.01, 0.003], [-0.2, 0.05, 0.001, -0.002, 0.01], [0.1, -0.05, 0, 0]):
        got = p.undistort(src, K, dist)
        want = oracle.undistort(src, K, np.array(dist))
        assert np.array_equal(got, want), f"{wh} n_dist={len(dist)}: {(got != want).any(axis=2).sum()} pixels differ"


def test_zero_model_is_a_copy_and_odd_matrices_are_refused():
    p = pkg()
    src = oracle.gen_frame(200, 100, 0, 0, oracle.NOISE)
    K = np.array([[180.0, 0, 100.0], [0, 180.0, 50.0], [0, 0, 1]])
    assert np.array_equal(p.undistort(src, K, np.zeros(5)), src)
    assert np.array_equal(p.undistort(src, K, []), src)
    Ks = K.copy()
    Ks[0, 1] = 0.5                     # skew: the row recurrence no longer factorises
    with pytest.raises(p.StitchError) as e:
        p.undistort(src, Ks, [0.1, 0, 0, 0])
    assert e.value.code == -4


def test_undistort_then_warp_matches_reference_order():
    """The video-mode loop body: undistort every frame, then warp it (main.cpp:145-152)."""
    p = pkg()
    w = workloads().get("cfg2")
    sw, sh = w.src_wh
    Ks = [np.array([[0.9 * sw, 0, sw / 2 + 3.3], [0, 0.92 * sw, sh / 2 - 1.7], [0, 0, 1]]),
          np.array([[1.1 * sw, 0, sw / 2 - 7.0], [0, 1.1 * sw, sh / 2 + 2.5], [0, 0, 1]])]
    dists = [np.array([-0.2, 0.05, 0.001, -0.002, 0.01, 0, 0, 0]), np.array([0.12, -0.03, -0.001, 0.0007, 0.0])]
    with p.Stitcher([w.src_wh] * 2, w.canvas_wh, w.H) as st:
        for i in range(2):
            st.set_undistort(i, Ks[i], dists[i])
        for frame in range(2):          # twice: the undistorted-frame buffers are reused
            srcs = [oracle.gen_frame(sw, sh, i, frame, oracle.NOISE) for i in range(2)]
            got = st.stitch(srcs)
            und = [oracle.undistort(srcs[i], Ks[i], dists[i]) for i in range(2)]
            want = oracle.stitch(und, w.H, w.canvas_wh)
            assert np.array_equal(got, want), f"frame {frame}: {(got != want).any(axis=2).sum()} pixels differ"
        st.set_undistort(1, Ks[1], np.zeros(5))      # switch camera 1 back to "no distortion"
        srcs = [oracle.gen_frame(sw, sh, i, 5, oracle.NOISE) for i in range(2)]
        got = st.stitch(srcs)
        want = oracle.stitch([oracle.undistort(srcs[0], Ks[0], dists[0]), srcs[1]], w.H, w.canvas_wh)
        assert np.array_equal(got, want)


@pytest.mark.parametrize("wh", [(333, 77), (1920, 1080), (4097, 31), (3840, 2160)])
def test_cached_map_path_matches_oracle(wh):
    """A handle evaluates the lens map once (bs_set_undistort) and resamples every frame from the
    packed map; identity homography, canvas = frame size, so the panorama is the undistorted frame.
    Mild and strong models (the strong ones push border pixels far outside the frame), odd widths
    (partial 32-pixel segments), two frames through the same map."""
    p = pkg()
    w, h = wh
    K = np.array([[0.9 * w, 0, w / 2 + 3.3], [0, 0.92 * w, h / 2 - 1.7], [0, 0, 1]])
    eye = [np.eye(3)]
    for dist in ([-0.25, 0.08, 0.001, 0.0005, 0.01, 0.02, -0.01, 0.003], [0.9, 2.5, 0.02, -0.03, 4.0], [-0.7, 0.3, 0, 0]):
        with p.Stitcher([wh], wh, eye) as st:
            st.set_undistort(0, K, dist)
            for frame in range(2):
                src = oracle.gen_frame(w, h, 1, frame, oracle.NOISE)
                got = st.stitch([src])
                want = oracle.stitch([oracle.undistort(src, K, np.array(dist, dtype=np.float64))], eye, wh)
                assert np.array_equal(got, want), f"{wh} dist={dist} frame {frame}: {(got != want).any(axis=2).sum()} pixels differ"
