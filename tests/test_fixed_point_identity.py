"""CPU check of the arithmetic identity the staged kernel relies on (csrc/stitch_fast.cuh,
tap_weights / blend_taps): with OpenCV's bilinear weights scaled to 16 bits (64 w', computed with the
same wrapping 32-bit products as the kernel) the two chained dp2a per channel leave the pixel in
byte 2 of  sum(64 w' S) + 2^15,  identical to (sum w*S + 2^14) >> 15 for every fractional position
and every 8-bit tap value that matters."""
import numpy as np

M32 = (1 << 32) - 1


def tap_weights(a, b):
    """Mirror of tap_weights() in stitch_fast.cuh, including the uint32 wrap-around."""
    a = a.astype(np.uint64)
    b = b.astype(np.uint64)
    a2 = (a * (0xFFFF * 64) + 2048) & M32
    w23 = (b * a2) & M32
    w01 = (a2 * 32 - w23) & M32
    w01 = np.where(w01 == 0x10000, 0xFFFF, w01)
    return w01, w23


def halves(w01, w23):
    return np.stack([w01 & 0xFFFF, w01 >> 16, w23 & 0xFFFF, w23 >> 16], axis=-1).astype(np.int64)


def reference(S, a, b):
    w = np.stack([(32 - a) * (32 - b) * 32, a * (32 - b) * 32, (32 - a) * b * 32, a * b * 32], axis=-1)
    return ((w * S).sum(-1) + (1 << 14)) >> 15


IDX = np.arange(1024)
A, B = IDX & 31, IDX >> 5


def test_packed_weights_are_the_scaled_opencv_weights():
    w = halves(*tap_weights(A, B))
    want = np.stack([(32 - A) * (32 - B), A * (32 - B), (32 - A) * B, A * B], axis=-1) * 64
    want[0, 0] = 65535                      # (a, b) = (0, 0): 65536 does not fit 16 bits
    assert np.array_equal(w, want)
    assert w.max() <= 0xFFFF and w.min() >= 0


def test_identity_all_fractions_random_taps():
    rng = np.random.default_rng(0)
    w = halves(*tap_weights(A, B))
    for _ in range(200):
        S = rng.integers(0, 256, (1024, 4))
        v = (S * w).sum(-1) + (1 << 15)
        assert v.max() < 1 << 24
        assert np.array_equal((v >> 16) & 255, reference(S, A, B))


def test_identity_extreme_taps():
    w = halves(*tap_weights(A, B))
    for pattern in range(16):
        S = np.array([255 if (pattern >> t) & 1 else 0 for t in range(4)])[None, :].repeat(1024, 0)
        v = (S * w).sum(-1) + (1 << 15)
        assert v.max() < 1 << 24
        assert np.array_equal((v >> 16) & 255, reference(S, A, B))
    for s in range(256):   # a = b = 0 uses 65535 instead of 65536
        assert ((65535 * s + (1 << 15)) >> 16) == s
