// Host build of the device JPEG decoder's functions (image-stitching_b200/csrc/stitch_jpegdec.cuh): decodes a file the
// way the GPU kernels do, block by block and pixel by pixel, and writes raw BGR.  Test infrastructure: lets
// tests/test_jpegdec.py pin the decoder's arithmetic against cv2.imdecode without a GPU.
// usage: jpegdec_host in.jpg out.bgr   (prints "w h hs vs dri intervals", exit 2 = not a file this decoder takes)
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../image-stitching_b200/csrc/stitch_jpegdec.cuh"

namespace jd {
// Host model of the kernels' rounds (one "thread" after the other): decodes interval [m_first, m_first + n_mcus) from
// `data` with `T` chunks, calling the product's run_chunk; the kernels in stitch_jpeg.inl run the same steps, one launch
// per round.  Test infrastructure: it lives here, not in the product header.
static int decode_interval_chunked_host(const Header &H, const HuffTable *tabs, const uint8_t *nat, const uint8_t *data, uint32_t bytes, int T,
                                        int m_first, int n_mcus, int16_t *const coef[3])
{
    const int nb = H.hs * H.vs + 2, max_block = n_mcus * nb;
    const uint32_t data_bits = bytes * 8u;
    uint32_t chunk_bits = ((data_bits + (uint32_t)T - 1) / (uint32_t)T + 7u) & ~7u;
    if (chunk_bits < 64) chunk_bits = 64;
    SyncState *exit_state = new SyncState[T], *next_state = new SyncState[T];
    int *started = new int[T];
    ChunkOut co;
    for (int j = 0; j < T; ++j) {                 // round 0: guesses
        SyncState s;
        s.pos = (uint32_t)j * chunk_bits;
        s.blk = 0;
        s.k = 0;
        exit_state[j] = run_chunk<false>(H, tabs, nat, data, data_bits, s, (uint32_t)(j + 1) * chunk_bits, 0, 0x7fffffff, m_first, coef, co);
    }
    int rounds = 1;
    for (;; ++rounds) {
        bool changed = false;
        for (int j = 0; j < T; ++j) {
            SyncState s;
            if (j == 0) { s.pos = 0; s.blk = 0; s.k = 0; } else s = exit_state[j - 1];
            next_state[j] = run_chunk<false>(H, tabs, nat, data, data_bits, s, (uint32_t)(j + 1) * chunk_bits, 0, 0x7fffffff, m_first, coef, co);
            started[j] = co.started;
            changed = changed || !same(next_state[j], exit_state[j]);
        }
        for (int j = 0; j < T; ++j) exit_state[j] = next_state[j];
        if (!changed) break;
        if (rounds > T + 2) break;
    }
    // prefix sums over the chunks: index of the first block started in a chunk, DC predictors at its entry
    // (the counts and sums of the last, unchanged round are the true ones)
    int first = 0, total_written = 0;
    int dc_off[3] = {0, 0, 0};
    for (int j = 0; j < T; ++j) {
        SyncState s;
        if (j == 0) { s.pos = 0; s.blk = 0; s.k = 0; } else s = exit_state[j - 1];
        co.dc_base[0] = dc_off[0];
        co.dc_base[1] = dc_off[1];
        co.dc_base[2] = dc_off[2];
        run_chunk<true>(H, tabs, nat, data, data_bits, s, (uint32_t)(j + 1) * chunk_bits, first, max_block, m_first, coef, co);
        for (int c = 0; c < 3; ++c) dc_off[c] += co.dc_sum[c];
        first += started[j];
        total_written += co.started;
    }
    first = total_written;
    delete[] exit_state;
    delete[] next_state;
    delete[] started;
    return first == max_block ? rounds : -rounds;
}

}  // namespace jd

int main(int argc, char **argv)
{
    if (argc < 3) return 1;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 1;
    std::vector<uint8_t> d;
    for (int c; (c = fgetc(f)) != EOF;) d.push_back((uint8_t)c);
    fclose(f);
    jd::Header H;
    if (!jd::parse_header(d.data(), d.size(), H)) return 2;
    const long total = (long)H.mcus_x * H.mcus_y;
    const int chunks = argc > 3 && atoi(argv[3]) > 0 ? atoi(argv[3]) : 1;        // chunks per interval (1: the sequential decode)
    const int max_iv = H.dri > 0 ? (int)((total + H.dri - 1) / H.dri) : 1;
    std::vector<uint32_t> offs(max_iv + 1);
    std::vector<int16_t> coef[3];
    int16_t *cp[3];
    for (int c = 0; c < 3; ++c) {
        coef[c].assign((size_t)64 * jd::blocks_w(H, c) * jd::blocks_h(H, c), 0);
        cp[c] = coef[c].data();
    }
    const jd::HuffTable tabs[4] = {H.dc[0], H.dc[1], H.ac[0], H.ac[1]};
    int n_iv = 0, worst_rounds = 0;
    {
        std::vector<uint8_t> clean(H.scan_len + 16);
        n_iv = jd::unstuff_scan(d.data() + H.scan_off, H.scan_len, H, clean.data(), offs.data(), max_iv);
        if (n_iv < 0) return 2;
        const int dri = H.dri > 0 ? H.dri : (int)total;
        for (int r = 0; r < n_iv; ++r) {
            const int m0 = r * dri, n_mcus = (int)std::min<long>(dri, total - m0);
            const int rounds = jd::decode_interval_chunked_host(H, tabs, jd::kNaturalHost, clean.data() + offs[r], offs[r + 1] - offs[r], chunks, m0, n_mcus, cp);
            if (rounds < 0) { fprintf(stderr, "interval %d: block count mismatch after %d rounds\n", r, -rounds); return 3; }
            worst_rounds = std::max(worst_rounds, rounds);
        }
    }
    std::vector<uint8_t> plane[3];
    size_t pitch[3];
    for (int c = 0; c < 3; ++c) {
        const int bw = jd::blocks_w(H, c), bh = jd::blocks_h(H, c);
        pitch[c] = (size_t)bw * 8;
        plane[c].assign(pitch[c] * bh * 8, 0);
        for (int by = 0; by < bh; ++by)
            for (int bx = 0; bx < bw; ++bx)
                jd::idct_islow(cp[c] + 64 * ((size_t)by * bw + bx), H.quant[c], plane[c].data() + (size_t)by * 8 * pitch[c] + bx * 8, pitch[c]);
    }
    std::vector<uint8_t> out((size_t)H.w * H.h * 3);
    const int cw = (H.w + H.hs - 1) / H.hs, ch = (H.h + H.vs - 1) / H.vs;
    for (int y = 0; y < H.h; ++y)
        for (int x = 0; x < H.w; ++x) {
            const int Y = plane[0][(size_t)y * pitch[0] + x];
            int cb, cr;
            if (H.hs == 2) {
                cb = jd::chroma_h2v2(plane[1].data(), pitch[1], cw, ch, x, y);
                cr = jd::chroma_h2v2(plane[2].data(), pitch[2], cw, ch, x, y);
            } else {
                cb = plane[1][(size_t)y * pitch[1] + x];
                cr = plane[2][(size_t)y * pitch[2] + x];
            }
            const uint32_t p = jd::ycc_to_bgr(Y, cb, cr);
            uint8_t *o = &out[((size_t)y * H.w + x) * 3];
            o[0] = (uint8_t)p;
            o[1] = (uint8_t)(p >> 8);
            o[2] = (uint8_t)(p >> 16);
        }
    f = fopen(argv[2], "wb");
    if (!f) return 1;
    fwrite(out.data(), 1, out.size(), f);
    fclose(f);
    printf("%d %d %d %d %d %d %d\n", H.w, H.h, H.hs, H.vs, H.dri, n_iv, worst_rounds);
    return 0;
}
