// stitch_undistort.cuh -- cv::undistort, the step the reference runs on every frame right before the
// warp in video mode (/root/reference/tools/stitch/main.cpp:148-149).  SURVEY.md row f1, Appendix A.6.
//
// OpenCV builds a fixed-point map with initUndistortRectifyMap (in row stripes of
// min(max(1, 4096 / w), h) rows, the principal point shifted per stripe) and resamples with the same
// 1/32-pixel bilinear kernel as the warp, BORDER_CONSTANT 0.  The normalised x coordinate is
// advanced along a row by repeated addition of 1/fx, so its value at column j carries j rounding
// steps; that sequence is identical for every row when the camera matrix has the usual form
// [[fx 0 cx] [0 fy cy] [0 0 1]], so the host tabulates it once (xs[j]) together with the per-row
// normalised y (ys[i]); the distortion model is then evaluated per pixel in fp64 with individually
// rounded operations, in the order of the scalar reference (oracle/warp_oracle.c).
//
// The reference rebuilds that map for every frame, but it depends on the calibration only.  A handle
// therefore evaluates it once (undistort_map_kernel) into one packed word per pixel and resamples
// every frame with undistort_remap_kernel: one coalesced map word, six aligned tap words, the frame
// kernel's dp2a blend.  undistort_kernel (map and resampling in one pass, fp64 per pixel) remains for
// the one-shot entry point, for unaligned buffers and for lenses whose displacement does not fit
// the packed word.
#pragma once

#include "stitch_fast.cuh"

namespace bs {

struct UndistortArgs {
    const uint8_t *src;
    uint8_t *dst;
    long long spitch, dpitch;
    int w, h;
    const double *xs;      // w entries: _x at column j (accumulated sequence)
    const double *ys;      // h entries: _y at row i (its stripe's inverse applied)
    double winv;           // 1 / _w
    double fx, fy, u0, v0;
    double k1, k2, p1, p2, k3, k4, k5, k6;
};

// Source position of destination pixel (j, i) in 1/32 pixel, before rounding (OpenCV's operation order).
__device__ __forceinline__ void undistort_real(const UndistortArgs &A, int j, int i, double &fu, double &fv)
{
    const double x = __dmul_rn(A.xs[j], A.winv), y = __dmul_rn(A.ys[i], A.winv);
    const double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
    const double r2 = __dadd_rn(x2, y2), _2xy = __dmul_rn(__dmul_rn(2.0, x), y);
    const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(A.k3, r2), A.k2), r2), A.k1), r2));
    const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(A.k6, r2), A.k5), r2), A.k4), r2));
    const double kr = __ddiv_rn(num, den);
    const double xd = __dadd_rn(__dadd_rn(__dmul_rn(x, kr), __dmul_rn(A.p1, _2xy)),
                                __dmul_rn(A.p2, __dadd_rn(r2, __dmul_rn(2.0, x2))));
    const double yd = __dadd_rn(__dadd_rn(__dmul_rn(y, kr), __dmul_rn(A.p1, __dadd_rn(r2, __dmul_rn(2.0, y2)))),
                                __dmul_rn(A.p2, _2xy));
    const double u = __dadd_rn(__dmul_rn(A.fx, xd), A.u0), v = __dadd_rn(__dmul_rn(A.fy, yd), A.v0);
    fu = __dmul_rn(u, (double)kFracOne);
    fv = __dmul_rn(v, (double)kFracOne);
}

// Fixed-point source position of destination pixel (j, i): texel (sx, sy) as OpenCV's int16 map holds
// it, fractions a, b in 1/32; (iu, iv) = the rounded 1/32-pixel coordinates before the int16 cut.
__device__ __forceinline__ void undistort_coord(const UndistortArgs &A, int j, int i, int &sx, int &sy, int &a, int &b, int &iu, int &iv)
{
    double fu, fv;
    undistort_real(A, j, i, fu, fv);
    iu = (fu == fu) ? __double2int_rn(fu) : INT_MIN;
    iv = (fv == fv) ? __double2int_rn(fv) : INT_MIN;
    sx = (int)(short)(iu >> kFracBits);   // OpenCV stores int16
    sy = (int)(short)(iv >> kFracBits);
    a = iu & (kFracOne - 1);
    b = iv & (kFracOne - 1);
}
__device__ __forceinline__ void undistort_coord(const UndistortArgs &A, int j, int i, int &sx, int &sy, int &a, int &b)
{
    int iu, iv;
    undistort_coord(A, j, i, sx, sy, a, b, iu, iv);
}

// Bilinear sample with BORDER_CONSTANT 0, tap by tap (any position, any alignment).
__device__ __forceinline__ void undistort_sample_checked(const uint8_t *__restrict__ src, long long spitch, int w, int h,
                                                         int sx, int sy, int a, int b, uint8_t *o)
{
    const int wgt[4] = {(kFracOne - a) * (kFracOne - b), a * (kFracOne - b), (kFracOne - a) * b, a * b};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        int acc = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int tx = sx + (t & 1), ty = sy + (t >> 1);
            if (tx >= 0 && tx < w && ty >= 0 && ty < h)
                acc += wgt[t] * __ldg(src + (long long)ty * spitch + (long long)tx * 3 + c);   // outside taps are 0
        }
        o[c] = (uint8_t)((acc + 512) >> 10);   // == (32 acc + 2^14) >> 15
    }
}

__global__ void __launch_bounds__(256) undistort_kernel(const __grid_constant__ UndistortArgs A)
{
    const int j = blockIdx.x * 64 + (threadIdx.x & 63);
    const int i = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (j >= A.w || i >= A.h) return;
    int sx, sy, a, b;
    undistort_coord(A, j, i, sx, sy, a, b);
    undistort_sample_checked(A.src, A.spitch, A.w, A.h, sx, sy, a, b, A.dst + (long long)i * A.dpitch + (long long)j * 3);
}

// ---- block plan of a lens model -------------------------------------------------------------------
// The lens field is as smooth as a homography's, so cv::undistort gets the warp's treatment: the frame is
// cut into the same 32x16 blocks, and a block all of whose pixels are interior samples (four taps inside the
// frame) becomes a FAST item of the compact frame kernel (stitch_fast2.cuh) -- source box by TMA, coordinates
// from a second-order fp32 expansion that this kernel checks pixel by pixel against undistort_coord().
// Blocks with a pixel that has a tap outside the frame (BORDER_CONSTANT 0 applies there), or that stick out
// of the frame, are EXACT (lens_quad below); blocks whose pixels all fall outside are EMPTY (zeros).
// One warp per block, lane = column, as in plan_kernel.
struct LensPlanArgs {
    UndistortArgs U;
    int blocks_x, blocks_y;
    BlockRec *out;
    CubicCoef *cubic;      // one per block, row-major like `out`
};

__global__ void __launch_bounds__(32 * kPlanWarps) lens_plan_kernel(const __grid_constant__ LensPlanArgs P)
{
    const UndistortArgs &A = P.U;
    const int lane = threadIdx.x & 31;
    const int bx = blockIdx.x * kPlanWarps + (threadIdx.x >> 5), by = blockIdx.y;
    if (bx >= P.blocks_x) return;
    const unsigned FULLMASK = 0xffffffffu;
    const int x = bx * kBlockW + lane, y0 = by * kBlockH;
    const bool whole = bx * kBlockW + kBlockW <= A.w && y0 + kBlockH <= A.h;

    BlockRec rec;
    memset(&rec, 0, sizeof rec);
    rec.bx = (uint16_t)bx;
    rec.by = (uint16_t)by;
    BlockRec *dst = P.out + (size_t)by * P.blocks_x + bx;

    // the expansion around the block centre: derivatives by central differences of the exact (unrounded) field
    BlockCoef bc;
    CubicCoef cq;
    memset(&bc, 0, sizeof bc);
    memset(&cq, 0, sizeof cq);
    if (whole) {
        const int xc = bx * kBlockW + kBlockW / 2, yc = y0 + kBlockH / 2;      // xc +- 1, yc +- 1 lie inside the frame
        double g[3][3][2];
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) undistort_real(A, xc + dx, yc + dy, g[dy + 1][dx + 1][0], g[dy + 1][dx + 1][1]);
        double d[2][6];     // per coordinate: value, d/du, d/dv, half d2/du2, d2/dudv, half d2/dv2
        for (int c = 0; c < 2; ++c) {
            d[c][0] = fmin(fmax(g[1][1][c], -1.0e9), 1.0e9);
            d[c][1] = 0.5 * (g[1][2][c] - g[1][0][c]);
            d[c][2] = 0.5 * (g[2][1][c] - g[0][1][c]);
            d[c][3] = 0.5 * (g[1][2][c] - 2.0 * g[1][1][c] + g[1][0][c]);
            d[c][4] = 0.25 * (g[2][2][c] - g[2][0][c] - g[0][2][c] + g[0][0][c]);
            d[c][5] = 0.5 * (g[2][1][c] - 2.0 * g[1][1][c] + g[0][1][c]);
        }
        bc = make_block_coef(d[0][0], d[1][0], d[0][1], d[0][2], d[1][1], d[1][2], d[0][3], d[0][4], d[0][5], d[1][3], d[1][4], d[1][5]);
        // third order with a step of 4 pixels (the samples stay inside the block)
        const int s4 = 4;
        auto F = [&](int dx, int dy, double &fu, double &fv) { undistort_real(A, xc + dx * s4, yc + dy * s4, fu, fv); };
        double p2[2], p1[2], m1[2], m2[2], c0[2], t[2];
        F(2, 0, p2[0], p2[1]); F(1, 0, p1[0], p1[1]); F(-1, 0, m1[0], m1[1]); F(-2, 0, m2[0], m2[1]); F(0, 0, c0[0], c0[1]);
        const double h3 = (double)(s4 * s4 * s4);
        double uuu[2], uuv[2], uvv[2];
        for (int c = 0; c < 2; ++c) uuu[c] = (p2[c] - 2.0 * p1[c] + 2.0 * m1[c] - m2[c]) / (2.0 * h3) / 6.0;      // f_uuu / 6
        double a11[2], a01[2], am1[2], b11[2], b01[2], bm1[2];
        F(1, 1, a11[0], a11[1]); F(0, 1, a01[0], a01[1]); F(-1, 1, am1[0], am1[1]);
        F(1, -1, b11[0], b11[1]); F(0, -1, b01[0], b01[1]); F(-1, -1, bm1[0], bm1[1]);
        for (int c = 0; c < 2; ++c) {
            uuv[c] = ((a11[c] - 2.0 * a01[c] + am1[c]) - (b11[c] - 2.0 * b01[c] + bm1[c])) / (2.0 * h3) / 2.0;    // f_uuv / 2
            uvv[c] = ((a11[c] - 2.0 * p1[c] + b11[c]) - (am1[c] - 2.0 * m1[c] + bm1[c])) / (2.0 * h3) / 2.0;      // f_uvv / 2
            t[c] = 0.0;
        }
        (void)t;
        cq.xuuu = (float)uuu[0]; cq.yuuu = (float)uuu[1];
        cq.xuuv = (float)uuv[0]; cq.yuuv = (float)uuv[1];
        cq.xuvv = (float)uvv[0]; cq.yuvv = (float)uvv[1];
    }
    const LaneCoef lc = lane_coef_cubic(bc, cq, lane - kBlockW / 2);

    bool all_interior = whole, all_outside = true, zf = false;
    unsigned mism = 0u;
    int bx0 = INT_MAX, by0 = INT_MAX, bx1 = INT_MIN, by1 = INT_MIN;
    if (x < A.w) {
        for (int k = 0; k < kRun; ++k) {
            const int y = y0 + k;
            if (y >= A.h) break;
            int sx, sy, a, b, X, Y;
            undistort_coord(A, x, y, sx, sy, a, b, X, Y);
            const bool interior = sx >= 0 && sy >= 0 && sx < A.w - 1 && sy < A.h - 1 && (X >> kFracBits) == sx && (Y >> kFracBits) == sy;
            all_interior = all_interior && interior;
            all_outside = all_outside && (sx < -1 || sx >= A.w || sy < -1 || sy >= A.h);
            if (interior) {
                bx0 = min(bx0, sx);
                by0 = min(by0, sy);
                bx1 = max(bx1, sx + 1);
                by1 = max(by1, sy + 1);
                zf = zf || ((a | b) == 0);
            }
            if (whole) {
                const int v = k - kRun / 2;
                int xbits, ybits;
                lane_coord_bits(lc, (float)v, xbits, ybits);
                if (X != xbits + lc.bx + 32 * lc.nxv * v || Y != ybits + lc.by + 32 * lc.nyv * v) mism |= 1u << (k >> 2);
            }
        }
    }
    all_interior = __all_sync(FULLMASK, all_interior);
    all_outside = __all_sync(FULLMASK, all_outside);
    if (all_outside) {
        rec.type = kBlockEmpty;
    } else if (!all_interior) {
        rec.type = kBlockExact;
    } else {
        rec.type = kBlockFast;
        rec.n_layers = 1;
        rec.layer[0].x0 = (int16_t)__reduce_min_sync(FULLMASK, bx0);
        rec.layer[0].y0 = (int16_t)__reduce_min_sync(FULLMASK, by0);
        rec.layer[0].x1 = (int16_t)__reduce_max_sync(FULLMASK, bx1);
        rec.layer[0].y1 = (int16_t)__reduce_max_sync(FULLMASK, by1);
        rec.layer[0].cam = 0;
        rec.layer[0].full = 1;
        rec.layer[0].groups = (uint8_t)__reduce_or_sync(FULLMASK, mism);
        rec.layer[0].zero_frac = __any_sync(FULLMASK, zf) ? 1 : 0;
        rec.coef[0] = bc;
    }
    if (lane == 0) {
        *dst = rec;
        P.cubic[(size_t)by * P.blocks_x + bx] = cq;
    }
}

// Four horizontally consecutive destination pixels starting at (x0, y), the exact way: fp64 lens model,
// tap by tap with BORDER_CONSTANT 0 (the EXACT blocks of a lens plan).
__device__ __forceinline__ void lens_quad(const UndistortArgs &A, int x0, int y)
{
    if (y >= A.h) return;
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x >= A.w) break;
        int sx, sy, a, b;
        undistort_coord(A, x, y, sx, sy, a, b);
        undistort_sample_checked(A.src, A.spitch, A.w, A.h, sx, sy, a, b, A.dst + (long long)y * A.dpitch + (long long)x * 3);
    }
}

// ---- cached map -------------------------------------------------------------------------------
// One word per pixel: a | b << 5 | (sx - j) << 10 | (sy - i) << 21, displacements as 11-bit two's
// complement; the row displacement -1024 is reserved for "every tap lies outside the frame" (the pixel
// is 0 whatever the frame holds).  A lens that moves a visible pixel by more than 1023 pixels sets
// *overflow and the handle keeps the one-pass kernel.
constexpr int kMapDispBits = 11;
constexpr int kMapOutside = -(1 << (kMapDispBits - 1));

__global__ void __launch_bounds__(256) undistort_map_kernel(const __grid_constant__ UndistortArgs A, uint32_t *__restrict__ map,
                                                            int *__restrict__ overflow)
{
    const int j = blockIdx.x * 64 + (threadIdx.x & 63);
    const int i = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (j >= A.w || i >= A.h) return;
    int sx, sy, a, b;
    undistort_coord(A, j, i, sx, sy, a, b);
    int dx = sx - j, dy = sy - i;
    const int lim = 1 << (kMapDispBits - 1);
    if (sx < -1 || sx >= A.w || sy < -1 || sy >= A.h) {
        dx = 0;
        dy = kMapOutside;
    } else if (dx < -lim || dx >= lim || dy <= -lim || dy >= lim) {
        atomicExch(overflow, 1);
    }
    map[(size_t)i * A.w + j] = (uint32_t)a | ((uint32_t)b << 5) | (((uint32_t)dx & 0x7FFu) << 10) | ((uint32_t)dy << 21);
}

struct RemapArgs {
    const uint8_t *src;    // 4-byte aligned, pitch a multiple of 4
    uint8_t *dst;          // 4-byte aligned, pitch a multiple of 4
    long long spitch, dpitch;
    int w, h;
    const uint32_t *map;   // w * h words, row-major without padding
};

// Lane = destination column within a 32-pixel segment; a warp owns kRemapRows consecutive rows of it
// (one strip); warps are persistent and walk the strips of the frame with a grid stride; the map words
// of the next strip are requested before the current strip is resampled, and all 24 tap words of a
// strip before any arithmetic (one pixel per thread was bound by the latency of the map -> tap load
// chain).  Interior pixels (all four taps inside, with a margin
// so that the three aligned tap words stay inside the row) take the frame kernel's path -- 2 x 3
// aligned words through L1, funnel-shift alignment, dp2a blend; a strip with any other pixel goes
// tap by tap.  Full segments are stored as 24 words per row (three 32-byte sectors).
#ifndef BS_REMAP_ROWS
#define BS_REMAP_ROWS 4
#endif
constexpr int kRemapRows = BS_REMAP_ROWS;

__device__ __forceinline__ void remap_store_row(uint8_t *row, uint32_t px, int lane, bool full, bool live)
{
    if (full) {
        const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
        const uint32_t nxt = __shfl_down_sync(0xffffffffu, px, 1);
        const uint32_t word = __byte_perm(px, nxt, sel);
        if ((lane & 3) != 3) *reinterpret_cast<uint32_t *>(row + 4 * (3 * (lane >> 2) + (lane & 3))) = word;
    } else if (live) {
        row[3 * lane] = (uint8_t)px;
        row[3 * lane + 1] = (uint8_t)(px >> 8);
        row[3 * lane + 2] = (uint8_t)(px >> 16);
    }
}

__global__ void __launch_bounds__(256) undistort_remap_kernel(const __grid_constant__ RemapArgs A)
{
    const int lane = threadIdx.x & 31;
    const int segs = (A.w + 31) >> 5;                      // 32-pixel segments per row
    const int strips_y = (A.h + kRemapRows - 1) / kRemapRows;
    const int stride = (int)gridDim.x * 8;                 // persistent warps: strip s, s + stride, ...
    const int dq = stride / segs, dr = stride - dq * segs; // the same step in (segment, strip row) form
    const int s0 = (int)blockIdx.x * 8 + (threadIdx.x >> 5);
    int by = s0 / segs, bx = s0 - by * segs;
    if (by >= strips_y) return;
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const uint32_t word_off = 4u * (uint32_t)(3 * (lane >> 2) + (lane & 3));

    // map words of a strip; rows past the frame and dead lanes shadow the last row / column
    auto load_map = [&](int sx_, int sy_, uint32_t (&m)[kRemapRows]) {
        const int jc = min(sx_ * 32 + lane, A.w - 1);
#pragma unroll
        for (int r = 0; r < kRemapRows; ++r) m[r] = __ldg(A.map + (size_t)min(sy_ * kRemapRows + r, A.h - 1) * A.w + jc);
    };
    // (Deeper pipelines -- tap words one strip ahead in a second register set -- need 128 registers and
    // lose more to occupancy than they hide: 37 us against 33 us per 4K frame.)
    uint32_t m[kRemapRows], mn[kRemapRows];
    load_map(bx, by, m);
    while (true) {
        int nbx = bx + dr, nby = by + dq;
        if (nbx >= segs) {
            nbx -= segs;
            ++nby;
        }
        const bool more = nby < strips_y;                  // warp-uniform
        if (more) load_map(nbx, nby, mn);                  // in flight while this strip is resampled
        const int i0 = by * kRemapRows;
        const int j = bx * 32 + lane;
        const bool whole = bx * 32 + 32 <= A.w && i0 + kRemapRows <= A.h;   // no dead lane, no row past the frame
        uint8_t *const out = A.dst + (long long)i0 * A.dpitch + (long long)(bx * 32) * 3;
        uint32_t off[kRemapRows];                          // byte offset of texel (sx, sy) in the frame (< 4 GB)
        bool interior = whole;
#pragma unroll
        for (int r = 0; r < kRemapRows; ++r) {
            const int sx = j + (((int)(m[r] << 11)) >> 21), sy = i0 + r + ((int)m[r] >> 21);
            // all four taps inside, with a margin for the third aligned word; "outside" words fail on dy
            interior = interior && (m[r] >> 21) != (uint32_t)(kMapOutside & 0x7FF) && (uint32_t)sx < (uint32_t)(A.w - 3) &&
                       (uint32_t)sy < (uint32_t)(A.h - 1);
            off[r] = (uint32_t)sy * (uint32_t)A.spitch + (uint32_t)sx * 3u;
        }
        if (__all_sync(0xffffffffu, interior)) {
            Taps t[kRemapRows];
#pragma unroll
            for (int r = 0; r < kRemapRows; ++r) {
                const uint32_t *r0 = reinterpret_cast<const uint32_t *>(A.src + (off[r] & ~3u));
                const uint32_t *r1 = reinterpret_cast<const uint32_t *>(A.src + (off[r] & ~3u) + (uint32_t)A.spitch);
                t[r].w0 = __ldg(r0); t[r].w1 = __ldg(r0 + 1); t[r].w2 = __ldg(r0 + 2);
                t[r].v0 = __ldg(r1); t[r].v1 = __ldg(r1 + 1); t[r].v2 = __ldg(r1 + 2);
                t[r].sh = off[r] << 3;                     // the frame base is 4-byte aligned
                t[r].wt = tap_weights(m[r] & 31u, (m[r] >> 5) & 31u);
            }
            uint8_t *wp = out + word_off;
#pragma unroll
            for (int r = 0; r < kRemapRows; ++r) {
                const uint32_t px = blend_taps(t[r]);
                const uint32_t nxt = __shfl_down_sync(0xffffffffu, px, 1);
                const uint32_t word = __byte_perm(px, nxt, sel);
                if ((lane & 3) != 3) *reinterpret_cast<uint32_t *>(wp) = word;
                wp += A.dpitch;
            }
        } else {
            const bool live = j < A.w;
            const bool full = bx * 32 + 32 <= A.w;
            const int jc = live ? j : A.w - 1;
            for (int r = 0; r < kRemapRows && i0 + r < A.h; ++r) {
                const int dy = (int)m[r] >> 21;
                const int sx = jc + (((int)(m[r] << 11)) >> 21), sy = i0 + r + dy;
                uint32_t px = 0;
                if (dy != kMapOutside) {
                    uint8_t o[3];
                    undistort_sample_checked(A.src, A.spitch, A.w, A.h, sx, sy, (int)(m[r] & 31u), (int)((m[r] >> 5) & 31u), o);
                    px = (uint32_t)o[0] | ((uint32_t)o[1] << 8) | ((uint32_t)o[2] << 16);
                }
                remap_store_row(out + (long long)r * A.dpitch, px, lane, full, live);
            }
        }
        if (!more) break;
        bx = nbx;
        by = nby;
#pragma unroll
        for (int r = 0; r < kRemapRows; ++r) m[r] = mn[r];
    }
}

}  // namespace bs
