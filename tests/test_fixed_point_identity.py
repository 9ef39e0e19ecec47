"""CPU check of the arithmetic identity the staged kernel relies on (csrc/stitch_fast.cuh,
lut_entry / sample_staged): splitting OpenCV's bilinear weights into two byte planes and taking
byte 2 of 256*dot(S, hi) + dot(S, lo) + 2^15 reproduces (sum w*S + 2^14) >> 15 for every
fractional position and every 8-bit tap value that matters."""
import numpy as np


def lut():
    a = np.arange(32)[None, :]
    b = np.arange(32)[:, None]
    w = np.stack([(32 - a) * (32 - b), a * (32 - b), (32 - a) * b, a * b], axis=-1).reshape(1024, 4)
    w = w.copy()
    w[w[:, 0] == 1024, 0] = 1023
    return w >> 2, (w & 3) << 6


def reference(S, a, b):
    w = np.stack([(32 - a) * (32 - b) * 32, a * (32 - b) * 32, (32 - a) * b * 32, a * b * 32], axis=-1)
    return ((w * S).sum(-1) + (1 << 14)) >> 15


def test_weight_planes_fit_bytes():
    hi, lo = lut()
    assert hi.max() <= 255 and lo.max() <= 255 and hi.min() >= 0


def test_identity_all_fractions_random_taps():
    rng = np.random.default_rng(0)
    hi, lo = lut()
    idx = np.arange(1024)
    a, b = idx & 31, idx >> 5
    for _ in range(200):
        S = rng.integers(0, 256, (1024, 4))
        v = 256 * ((S * hi).sum(-1) + 128) + (S * lo).sum(-1)
        assert v.max() < 1 << 24
        assert np.array_equal((v >> 16) & 255, reference(S, a, b))


def test_identity_extreme_taps():
    hi, lo = lut()
    idx = np.arange(1024)
    a, b = idx & 31, idx >> 5
    for pattern in range(16):
        S = np.array([255 if (pattern >> t) & 1 else 0 for t in range(4)])[None, :].repeat(1024, 0)
        v = 256 * ((S * hi).sum(-1) + 128) + (S * lo).sum(-1)
        assert v.max() < 1 << 24
        assert np.array_equal((v >> 16) & 255, reference(S, a, b))
    for s in range(256):   # a = b = 0 uses 1023 instead of 1024
        assert ((1023 * s + 512) >> 10) == s
