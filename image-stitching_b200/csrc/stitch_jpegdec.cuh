// stitch_jpegdec.cuh -- a baseline-JPEG decoder whose entropy stage runs on the GPU.
//
// Why: the reference decodes every frame on the host (cv::imread / cv::VideoCapture, tools/stitch/main.cpp:107,114,147),
// and this image's nvJPEG does the same for its Huffman stage (~20 ms per 4K frame and core), which caps
// bs_submit_jpeg far below what the link and the kernels can carry.  Here the Huffman data is decoded on the device,
// in parallel inside a scan (self-synchronising chunks, see "Parallel entropy decoding" below); inverse DCT, chroma
// up-sampling and colour conversion are two more kernels (stitch_jpeg.inl).
//
// Arithmetic: libjpeg's, so that the pixels are what cv::imread yields (libjpeg-turbo, default settings): the
// integer "islow" inverse DCT (jidctint.c), "fancy" h2v2 chroma upsampling (jdsample.c: the triangle filter with its
// 8/7 rounding pair) and the 16-bit fixed-point YCbCr -> RGB tables (jdcolor.c).  tests/test_jpegdec.py pins the
// host build of these same functions against cv2.imdecode bit for bit.
//
// Supported: baseline sequential DCT (SOF0), 8 bit, three components with Y sampled 2x2 or 1x1 and chroma 1x1,
// one interleaved scan, with or without a restart interval.  Everything else is left to nvJPEG (stitch_jpeg.inl).
//
// The functions below compile for the device and for the host (the CPU check of the decoder needs no GPU).
#pragma once

#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define JD_HD __host__ __device__ __forceinline__
#else
#define JD_HD inline
#endif

namespace jd {

// Zig-zag position -> natural (row-major) index.
#define JD_NATURAL_ORDER                                                                                                  \
    {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, \
     35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63}
static const uint8_t kNaturalHost[64] = JD_NATURAL_ORDER;
#ifdef __CUDACC__
__device__ __constant__ uint8_t kNaturalDev[64] = JD_NATURAL_ORDER;
#endif
JD_HD int natural(int k)
{
#ifdef __CUDA_ARCH__
    return kNaturalDev[k];
#else
    return kNaturalHost[k];
#endif
}

struct HuffTable {
    uint8_t vals[256];
    int32_t maxcode[18];     // largest code of each length (-1: none); [17] ends every search
    int32_t valoff[17];      // index of the first value of a length minus its smallest code
    uint16_t look[512];      // 9 bits of look-ahead: length << 8 | value, 0 when the code is longer
    int16_t fast[512];       // AC tables: when code + magnitude bits fit the 9 bits, value << 8 | run << 4 | total length
};

struct Header {
    int w = 0, h = 0;
    int hs = 0, vs = 0;              // luma sampling factors (chroma is 1x1): 2,2 or 1,1
    int dri = 0;                     // MCUs per restart interval
    int mcus_x = 0, mcus_y = 0;      // MCU grid
    uint16_t quant[3][64];           // per component, NATURAL order
    int dc_tab[3], ac_tab[3];        // Huffman table of each component
    HuffTable dc[2], ac[2];
    size_t scan_off = 0, scan_len = 0;   // entropy-coded data: [scan_off, scan_off + scan_len) of the file (up to EOI)
};

// ---- host: headers --------------------------------------------------------------------------------------
inline bool build_huff(const uint8_t *bits /* 16 counts */, const uint8_t *vals, int nvals, HuffTable &t)
{
    memset(&t, 0, sizeof t);
    memcpy(t.vals, vals, (size_t)nvals);
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        t.valoff[l] = k - code;
        if (bits[l - 1]) {
            for (int i = 0; i < bits[l - 1]; ++i, ++k, ++code) {
                if (l <= 9) {
                    const int first = code << (9 - l), count = 1 << (9 - l);
                    if (first + count > 512 || k >= nvals) return false;
                    for (int j = 0; j < count; ++j) t.look[first + j] = (uint16_t)((l << 8) | vals[k]);
                }
            }
            t.maxcode[l] = code - 1;
        } else {
            t.maxcode[l] = -1;
        }
        if (code > (1 << l)) return false;
        code <<= 1;
    }
    t.maxcode[17] = 0x7fffffff;
    // run/size symbols whose code AND magnitude bits fit the look-ahead decode in one step (most AC coefficients)
    for (int i = 0; i < 512; ++i) {
        t.fast[i] = 0;
        const int e = t.look[i];
        if (!e) continue;
        const int len = e >> 8, run = (e >> 4) & 15, mag = e & 15;
        if (mag && len + mag <= 9) {
            int v = ((i << len) & 511) >> (9 - mag);
            if (v < (1 << (mag - 1))) v -= (1 << mag) - 1;
            if (v >= -128 && v <= 127) t.fast[i] = (int16_t)(v * 256 + run * 16 + len + mag);
        }
    }
    return k == nvals;
}

// Parses the marker segments of a JPEG file up to the first scan.  Returns false for anything this decoder does
// not take (the caller then uses nvJPEG).
inline bool parse_header(const uint8_t *d, size_t n, Header &H)
{
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return false;
    H.w = H.h = H.hs = H.vs = H.dri = 0;                   // (a header object is reused from frame to frame)
    uint16_t qt[4][64];
    bool have_q[4] = {false, false, false, false}, have_dc[2] = {false, false}, have_ac[2] = {false, false}, have_sof = false;
    int comp_id[3] = {0, 0, 0}, comp_tq[3] = {0, 0, 0};
    size_t i = 2;
    while (i + 4 <= n) {
        if (d[i] != 0xFF) return false;
        const uint8_t m = d[i + 1];
        if (m == 0xFF) { ++i; continue; }
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD8)) { i += 2; continue; }
        const size_t len = ((size_t)d[i + 2] << 8) | d[i + 3];
        if (len < 2 || i + 2 + len > n) return false;
        const uint8_t *s = d + i + 4;
        const size_t sl = len - 2;
        if (m == 0xDB) {                                   // DQT
            size_t p = 0;
            while (p < sl) {
                const int pq = s[p] >> 4, tq = s[p] & 15;
                if (tq > 3 || pq > 1 || p + 1 + (pq ? 128 : 64) > sl) return false;
                for (int k = 0; k < 64; ++k) qt[tq][natural(k)] = pq ? (uint16_t)((s[p + 1 + 2 * k] << 8) | s[p + 2 + 2 * k]) : s[p + 1 + k];
                have_q[tq] = true;
                p += 1 + (pq ? 128 : 64);
            }
        } else if (m == 0xC4) {                            // DHT
            size_t p = 0;
            while (p + 17 <= sl) {
                const int tc = s[p] >> 4, th = s[p] & 15;
                int nv = 0;
                for (int k = 0; k < 16; ++k) nv += s[p + 1 + k];
                if (tc > 1 || th > 1 || nv > 256 || p + 17 + nv > sl) return false;
                if (!build_huff(s + p + 1, s + p + 17, nv, tc ? H.ac[th] : H.dc[th])) return false;
                (tc ? have_ac : have_dc)[th] = true;
                p += 17 + nv;
            }
        } else if (m == 0xC0) {                            // SOF0: baseline
            if (sl < 15 || s[0] != 8 || s[5] != 3) return false;
            H.h = (s[1] << 8) | s[2];
            H.w = (s[3] << 8) | s[4];
            for (int c = 0; c < 3; ++c) {
                comp_id[c] = s[6 + 3 * c];
                const int hv = s[7 + 3 * c];
                comp_tq[c] = s[8 + 3 * c];
                if (comp_tq[c] > 3) return false;
                if (c == 0) {
                    H.hs = hv >> 4;
                    H.vs = hv & 15;
                    if (!((H.hs == 2 && H.vs == 2) || (H.hs == 1 && H.vs == 1))) return false;
                } else if (hv != 0x11) {
                    return false;
                }
            }
            if (H.w < 1 || H.h < 1) return false;
            have_sof = true;
        } else if (m == 0xDD) {                            // DRI
            if (sl < 2) return false;
            H.dri = (s[0] << 8) | s[1];
        } else if (m == 0xDA) {                            // SOS
            if (!have_sof || sl < 10 || s[0] != 3) return false;
            for (int c = 0; c < 3; ++c) {
                if (s[1 + 2 * c] != comp_id[c]) return false;          // interleaved in frame order
                H.dc_tab[c] = s[2 + 2 * c] >> 4;
                H.ac_tab[c] = s[2 + 2 * c] & 15;
                if (H.dc_tab[c] > 1 || H.ac_tab[c] > 1 || !have_dc[H.dc_tab[c]] || !have_ac[H.ac_tab[c]] || !have_q[comp_tq[c]]) return false;
                memcpy(H.quant[c], qt[comp_tq[c]], sizeof H.quant[c]);
            }
            if (s[7] != 0 || s[8] != 63 || s[9] != 0) return false;   // spectral selection / approximation of a sequential scan
            H.mcus_x = (H.w + 8 * H.hs - 1) / (8 * H.hs);
            H.mcus_y = (H.h + 8 * H.vs - 1) / (8 * H.vs);
            H.scan_off = i + 2 + len;
            size_t e = n;                                  // the scan runs up to EOI (or to the end of what we were given)
            while (e >= H.scan_off + 2 && !(d[e - 2] == 0xFF && d[e - 1] == 0xD9)) --e;
            H.scan_len = (e >= H.scan_off + 2 ? e - 2 : n) - H.scan_off;
            return true;
        } else if (m >= 0xC1 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return false;                                  // any other frame type (progressive, lossless, arithmetic ...)
        }
        i += 2 + len;
    }
    return false;
}

// ---- entropy decoding (host and device) ------------------------------------------------------------------------
JD_HD uint32_t br_byte(const uint8_t *p)
{
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}

// Where component c keeps its coefficient blocks: `blocks_w(c)` blocks per row, padded to whole MCUs.
JD_HD int blocks_w(const Header &H, int c) { return H.mcus_x * (c == 0 ? H.hs : 1); }
JD_HD int blocks_h(const Header &H, int c) { return H.mcus_y * (c == 0 ? H.vs : 1); }

// ============================================================================================================
// Parallel entropy decoding INSIDE a restart interval (or inside a whole scan without restart markers).
//
// The host removes the byte stuffing (0xFF 0x00 -> 0xFF) and the restart markers while it copies the scan data
// (unstuff_scan), so that a position in an interval is a plain bit offset.  An interval is cut into chunks of equal
// size; the state of the decoder at a symbol boundary is (bit position, block within the MCU, zig-zag index).  Every
// thread decodes ITS chunk starting from the exit state of the chunk before it -- a guess in the first round
// (chunk start, block 0, index 0), the previous round's result afterwards -- and records the state in which it
// leaves the chunk.  Chunk 0 starts from the truth, every exit state is a function of the entry state, so after
// round r at least the first r chunks are right; Huffman streams re-synchronise quickly, so in practice every exit
// state is right after two or three rounds, and a round in which nothing changed proves it (the fixed point is the
// sequential decode).  Then: blocks started per chunk -> prefix sum -> every chunk knows the index of its first
// block; one more pass writes the coefficients (DC as differences); DC predictors follow from a prefix sum of the
// differences per component.  (Weissenberger & Schmidt's self-synchronising scheme, restated for this layout.)
// ============================================================================================================
struct SyncState {
    uint32_t pos;        // bit offset in the interval's data
    int16_t blk;         // block within the MCU whose coefficients come next (0 .. blocks per MCU - 1)
    int16_t k;           // zig-zag index that comes next (0: the block's DC coefficient)
};
JD_HD bool same(const SyncState &a, const SyncState &b) { return a.pos == b.pos && a.blk == b.blk && a.k == b.k; }

// Copies entropy-coded data without stuffing and markers; offs[i] = byte offset of interval i in `dst`
// (n_intervals + 1 entries).  Returns the interval count, or -1 when the markers do not fit the header.
// `dst` needs len + 16 bytes (8 bytes of zero padding are written behind the data).
inline int unstuff_scan(const uint8_t *scan, size_t len, const Header &H, uint8_t *dst, uint32_t *offs, int max_intervals)
{
    const long total = (long)H.mcus_x * H.mcus_y;
    const int want = H.dri > 0 ? (int)((total + H.dri - 1) / H.dri) : 1;
    if (want > max_intervals) return -1;
    int n = 0;
    size_t o = 0;
    offs[n++] = 0;
    const uint8_t *p = scan, *end = scan + len;
    while (p < end) {
        const uint8_t *q = static_cast<const uint8_t *>(memchr(p, 0xFF, (size_t)(end - p)));
        if (!q) q = end;
        memcpy(dst + o, p, (size_t)(q - p));
        o += (size_t)(q - p);
        if (q + 1 >= end) break;
        if (q[1] == 0x00) {
            dst[o++] = 0xFF;
            p = q + 2;
        } else if (q[1] >= 0xD0 && q[1] <= 0xD7) {
            if (n >= want || q[1] != 0xD0 + ((n - 1) & 7)) return -1;
            offs[n++] = (uint32_t)o;
            p = q + 2;
        } else if (q[1] == 0xFF) {
            p = q + 1;                                   // fill byte
        } else {
            break;                                       // any other marker ends the scan
        }
    }
    if (n != want) return -1;
    offs[n] = (uint32_t)o;
    memset(dst + o, 0, 8);
    return want;
}

// A bit reader over stuffing-free data: position = a bit offset.
struct CleanReader {
    const uint8_t *base;
    uint32_t pos, end;       // bits
    uint64_t buf;
    int cnt;                 // valid bits in buf (the reader has consumed `pos` bits; buf holds the next cnt)
};
JD_HD void cr_init(CleanReader &b, const uint8_t *base, uint32_t pos, uint32_t end_bits)
{
    b.base = base;
    b.pos = pos;
    b.end = end_bits;
    b.buf = 0;
    b.cnt = 0;
    // the first byte may be used from the middle: load it and drop the bits in front
    const uint32_t byte = pos >> 3, off = pos & 7;
    b.buf = (uint64_t)br_byte(base + byte) << 56;
    b.buf <<= off;
    b.cnt = 8 - (int)off;
}
JD_HD void cr_fill(CleanReader &b)      // data has 8 readable bytes behind its end
{
    uint32_t next = (b.pos + (uint32_t)b.cnt) >> 3;      // first byte not in the buffer
    while (b.cnt <= 56) {
        b.buf |= (uint64_t)br_byte(b.base + next) << (56 - b.cnt);
        b.cnt += 8;
        ++next;
    }
}
JD_HD uint32_t cr_peek(const CleanReader &b, int n) { return (uint32_t)(b.buf >> (64 - n)); }
JD_HD void cr_skip(CleanReader &b, int n)
{
    b.buf <<= n;
    b.cnt -= n;
    b.pos += (uint32_t)n;
}

// What a pass over a chunk does besides following the symbols.
struct ChunkOut {
    int started;             // blocks whose DC symbol lies in the chunk
    int dc_sum[3];           // sum of the DC differences per component of the blocks started here
    int dc_base[3];          // in, write pass: DC predictor of each component at the chunk's entry (a block's DC
                             // coefficient is dc_base + the running dc_sum)
};

// Decodes from state `s` up to the first symbol boundary at or behind bit `limit` (or the end of the data, or block
// `max_block` of the interval).  kWrite: coefficients go to their blocks (DC as a difference); `first_block` = index,
// within the interval, of the first block STARTED in this chunk (the block in progress at entry is first_block - 1).
template <bool kWrite>
JD_HD SyncState run_chunk(const Header &H, const HuffTable *tabs, const uint8_t *nat, const uint8_t *data, uint32_t data_bits, SyncState s,
                          uint32_t limit, int first_block, int max_block, int m_first, int16_t *const coef[3], ChunkOut &out)
{
    const int nb = H.hs * H.vs + 2;
    out.started = 0;
    out.dc_sum[0] = out.dc_sum[1] = out.dc_sum[2] = 0;
    if (s.pos >= limit || s.pos >= data_bits) return s;
    CleanReader b;
    cr_init(b, data, s.pos, data_bits);
    int blk = s.blk, k = s.k;
    int block_index = first_block - 1;          // of the block in progress
    int16_t *cf = nullptr;
    const int bw0 = blocks_w(H, 0);
    auto block_ptr = [&](int bi, int sub) -> int16_t * {
        const int m = m_first + bi / nb, mx = m % H.mcus_x, my = m / H.mcus_x;
        if (sub < nb - 2) {
            const int u = sub % H.hs, v = sub / H.hs;
            return coef[0] + 64 * ((size_t)(my * H.vs + v) * bw0 + (size_t)(mx * H.hs + u));
        }
        return coef[sub - (nb - 2) + 1] + 64 * ((size_t)my * H.mcus_x + (size_t)mx);
    };
    if (kWrite && k > 0 && block_index >= 0 && block_index < max_block) cf = block_ptr(block_index, blk);
    while (b.pos < limit && b.pos < data_bits) {
        const int comp = blk < nb - 2 ? 0 : blk - (nb - 2) + 1;
        if (b.cnt < 32) cr_fill(b);
        if (k == 0) {                                          // DC
            if (first_block + out.started >= max_block) break; // the interval's blocks are all there (padding bits follow)
            const HuffTable &t = tabs[H.dc_tab[comp]];
            const uint32_t e = t.look[cr_peek(b, 9)];
            int sym, len;
            if (e) {
                len = (int)(e >> 8);
                sym = (int)(e & 0xFF);
            } else {
                const uint32_t c16 = cr_peek(b, 16);
                len = 10;
                while (len <= 16 && (int)(c16 >> (16 - len)) > t.maxcode[len]) ++len;
                sym = len <= 16 ? t.vals[((int)(c16 >> (16 - len)) + t.valoff[len]) & 0xFF] : 0;
                if (len > 16) len = 16;
            }
            cr_skip(b, len);
            sym &= 15;
            int diff = 0;
            if (sym) {
                const int v = (int)cr_peek(b, sym);
                cr_skip(b, sym);
                diff = v < (1 << (sym - 1)) ? v - (1 << sym) + 1 : v;
            }
            block_index = first_block + out.started;
            ++out.started;
            out.dc_sum[comp] += diff;
            if (kWrite) {
                cf = block_ptr(block_index, blk);
                cf[0] = (int16_t)(out.dc_base[comp] + out.dc_sum[comp]);     // predictor so far + this difference (or the bare difference)
            }
            k = 1;
            continue;
        }
        const HuffTable &t = tabs[2 + H.ac_tab[comp]];
        const int f = t.fast[cr_peek(b, 9)];
        if (f) {
            k += (f >> 4) & 15;
            cr_skip(b, f & 15);
            if (kWrite && cf && k < 64) cf[nat[k]] = (int16_t)(f >> 8);
            ++k;
        } else {
            const uint32_t e = t.look[cr_peek(b, 9)];
            int sym, len;
            if (e) {
                len = (int)(e >> 8);
                sym = (int)(e & 0xFF);
            } else {
                const uint32_t c16 = cr_peek(b, 16);
                len = 10;
                while (len <= 16 && (int)(c16 >> (16 - len)) > t.maxcode[len]) ++len;
                sym = len <= 16 ? t.vals[((int)(c16 >> (16 - len)) + t.valoff[len]) & 0xFF] : 0;
                if (len > 16) len = 16;
            }
            cr_skip(b, len);
            const int r = sym >> 4, sz = sym & 15;
            if (sz) {
                k += r;
                const int v = (int)cr_peek(b, sz);
                cr_skip(b, sz);
                if (kWrite && cf && k < 64) cf[nat[k]] = (int16_t)(v < (1 << (sz - 1)) ? v - (1 << sz) + 1 : v);
                ++k;
            } else if (r == 15) {
                k += 16;
            } else {
                k = 64;                                        // end of block
            }
        }
        if (k >= 64) {
            k = 0;
            blk = blk + 1 == nb ? 0 : blk + 1;
            cf = nullptr;
        }
    }
    SyncState o;
    o.pos = b.pos;
    o.blk = (int16_t)blk;
    o.k = (int16_t)k;
    return o;
}

// ---- inverse DCT: libjpeg's jpeg_idct_islow (jidctint.c), 13-bit constants, two passes -----------------------
JD_HD uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
JD_HD int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

JD_HD void idct_islow(const int16_t *coef, const uint16_t *q, uint8_t *out, size_t pitch)
{
    constexpr int C = 13, P1 = 2;
    constexpr int F0_298 = 2446, F0_390 = 3196, F0_541 = 4433, F0_765 = 6270, F0_899 = 7373, F1_175 = 9633, F1_501 = 12299,
                  F1_847 = 15137, F1_961 = 16069, F2_053 = 16819, F2_562 = 20995, F3_072 = 25172;
    int ws[64];
    for (int x = 0; x < 8; ++x) {       // columns
        int z2 = coef[16 + x] * q[16 + x], z3 = coef[48 + x] * q[48 + x];
        int z1 = (z2 + z3) * F0_541;
        int tmp2 = z1 + z3 * (-F1_847), tmp3 = z1 + z2 * F0_765;
        z2 = coef[x] * q[x];
        z3 = coef[32 + x] * q[32 + x];
        int tmp0 = (z2 + z3) * (1 << C), tmp1 = (z2 - z3) * (1 << C);      // (a multiplication: the sums may be negative)
        const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
        tmp0 = coef[56 + x] * q[56 + x];
        tmp1 = coef[40 + x] * q[40 + x];
        tmp2 = coef[24 + x] * q[24 + x];
        tmp3 = coef[8 + x] * q[8 + x];
        z1 = tmp0 + tmp3;
        z2 = tmp1 + tmp2;
        z3 = tmp0 + tmp2;
        int z4 = tmp1 + tmp3;
        const int z5 = (z3 + z4) * F1_175;
        tmp0 *= F0_298;
        tmp1 *= F2_053;
        tmp2 *= F3_072;
        tmp3 *= F1_501;
        z1 *= -F0_899;
        z2 *= -F2_562;
        z3 = z3 * (-F1_961) + z5;
        z4 = z4 * (-F0_390) + z5;
        tmp0 += z1 + z3;
        tmp1 += z2 + z4;
        tmp2 += z2 + z3;
        tmp3 += z1 + z4;
        ws[x] = descale(tmp10 + tmp3, C - P1);
        ws[56 + x] = descale(tmp10 - tmp3, C - P1);
        ws[8 + x] = descale(tmp11 + tmp2, C - P1);
        ws[48 + x] = descale(tmp11 - tmp2, C - P1);
        ws[16 + x] = descale(tmp12 + tmp1, C - P1);
        ws[40 + x] = descale(tmp12 - tmp1, C - P1);
        ws[24 + x] = descale(tmp13 + tmp0, C - P1);
        ws[32 + x] = descale(tmp13 - tmp0, C - P1);
    }
    for (int y = 0; y < 8; ++y) {       // rows
        const int *w = ws + 8 * y;
        int z2 = w[2], z3 = w[6];
        int z1 = (z2 + z3) * F0_541;
        int tmp2 = z1 + z3 * (-F1_847), tmp3 = z1 + z2 * F0_765;
        int tmp0 = (w[0] + w[4]) * (1 << C), tmp1 = (w[0] - w[4]) * (1 << C);
        const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
        tmp0 = w[7];
        tmp1 = w[5];
        tmp2 = w[3];
        tmp3 = w[1];
        z1 = tmp0 + tmp3;
        z2 = tmp1 + tmp2;
        z3 = tmp0 + tmp2;
        int z4 = tmp1 + tmp3;
        const int z5 = (z3 + z4) * F1_175;
        tmp0 *= F0_298;
        tmp1 *= F2_053;
        tmp2 *= F3_072;
        tmp3 *= F1_501;
        z1 *= -F0_899;
        z2 *= -F2_562;
        z3 = z3 * (-F1_961) + z5;
        z4 = z4 * (-F0_390) + z5;
        tmp0 += z1 + z3;
        tmp1 += z2 + z4;
        tmp2 += z2 + z3;
        tmp3 += z1 + z4;
        uint8_t *o = out + (size_t)y * pitch;
        constexpr int S = C + P1 + 3;
        o[0] = clamp_u8(descale(tmp10 + tmp3, S) + 128);
        o[7] = clamp_u8(descale(tmp10 - tmp3, S) + 128);
        o[1] = clamp_u8(descale(tmp11 + tmp2, S) + 128);
        o[6] = clamp_u8(descale(tmp11 - tmp2, S) + 128);
        o[2] = clamp_u8(descale(tmp12 + tmp1, S) + 128);
        o[5] = clamp_u8(descale(tmp12 - tmp1, S) + 128);
        o[3] = clamp_u8(descale(tmp13 + tmp0, S) + 128);
        o[4] = clamp_u8(descale(tmp13 - tmp0, S) + 128);
    }
}

// ---- chroma upsampling (jdsample.c h2v2_fancy_upsample) and colour conversion (jdcolor.c) for one pixel ---------
// `plane` = a chroma plane of `cw` x `ch` REAL samples (ceil(w/2) x ceil(h/2)), row pitch `pitch`.
JD_HD int chroma_h2v2(const uint8_t *plane, size_t pitch, int cw, int ch, int x, int y)
{
    const int cy = y >> 1, cx = x >> 1;
    int ny = (y & 1) ? cy + 1 : cy - 1;                 // the nearer neighbour row (image edges repeat the edge row)
    ny = ny < 0 ? 0 : (ny > ch - 1 ? ch - 1 : ny);
    const uint8_t *r0 = plane + (size_t)cy * pitch, *r1 = plane + (size_t)ny * pitch;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) {
        if (cx == cw - 1) return (cur * 4 + 7) >> 4;
        return (cur * 3 + (3 * r0[cx + 1] + r1[cx + 1]) + 7) >> 4;
    }
    if (cx == 0) return (cur * 4 + 8) >> 4;
    return (cur * 3 + (3 * r0[cx - 1] + r1[cx - 1]) + 8) >> 4;
}
// B | G << 8 | R << 16
JD_HD uint32_t ycc_to_bgr(int y, int cb, int cr)
{
    const int xb = cb - 128, xr = cr - 128;
    const int r = y + ((91881 * xr + 32768) >> 16);
    const int g = y + ((-22554 * xb + 32768 - 46802 * xr) >> 16);
    const int b = y + ((116130 * xb + 32768) >> 16);
    return (uint32_t)clamp_u8(b) | ((uint32_t)clamp_u8(g) << 8) | ((uint32_t)clamp_u8(r) << 16);
}

}  // namespace jd
