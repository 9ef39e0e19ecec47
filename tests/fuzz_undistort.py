"""Random sweep of the cached-map undistort path (run by hand on a GPU box:
python tests/fuzz_undistort.py SEED N): random frame sizes, camera matrices and lens models through a
handle with an identity homography, against the CPU oracle's cv::undistort restatement."""
import sys, importlib, time, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
p = importlib.import_module('image-stitching_b200')

seed0 = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n_trials = int(sys.argv[2]) if len(sys.argv) > 2 else 100
bad = 0; t0 = time.time()
eye = [np.eye(3)]
for trial in range(n_trials):
    rng = np.random.default_rng(seed0 * 100000 + trial)
    big = trial % 6 == 0
    w, h = (int(rng.integers(1500, 4200)), int(rng.integers(700, 2300))) if big else (int(rng.integers(1, 1400)), int(rng.integers(1, 900)))
    K = np.array([[rng.uniform(0.3, 2.0) * w, 0, w * rng.uniform(0.3, 0.7)], [0, rng.uniform(0.3, 2.0) * w, h * rng.uniform(0.3, 0.7)], [0, 0, 1]])
    n_dist = int(rng.choice([4, 5, 8]))
    mag = float(rng.choice([0.02, 0.3, 1.5]))
    dist = rng.uniform(-mag, mag, n_dist)
    dist[2:4] *= 0.05                      # tangential terms are small in practice
    src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    want = oracle.stitch([oracle.undistort(src, K, dist)], eye, (w, h))
    with p.Stitcher([(w, h)], (w, h), eye) as st:
        st.set_undistort(0, K, dist)
        got = st.stitch([src])
    if not np.array_equal(got, want):
        bad += 1
        print(f"MISMATCH seed {seed0} trial {trial}: {w}x{h} n_dist {n_dist} mag {mag}: {(got != want).any(axis=2).sum()} px", flush=True)
print(f"seed {seed0}: {n_trials} trials, {bad} mismatches, {time.time() - t0:.1f} s")
