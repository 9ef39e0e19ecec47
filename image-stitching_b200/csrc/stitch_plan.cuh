// stitch_plan.cuh -- the block plan: static per calibration, built once on the device.
//
// The reference recomputes everything per frame (cv::warpPerspective regenerates the coordinate
// maps on every call, /root/reference/tools/stitch/main.cpp:150-151), but H and result_size are
// fixed for the whole run (:53-73).  The plan classifies every 32x16 canvas block once:
//   EMPTY  no camera writes any pixel                      -> zero fill
//   FAST   every written pixel is an interior (4-tap) pixel of one of <= 2 cameras whose source
//          bounding boxes fit a shared-memory stage        -> staged kernel, one pass per camera
//   EXACT  anything else (last-row/column pixels, three or more cameras, huge boxes)
//                                                            -> per-pixel fp64 path
// and, for FAST blocks, proves by exhaustive comparison which 32x4 pixel groups may use the cheap
// fp32 coordinate recurrence (run_coord_bits) instead of the exact fp64 form.
#pragma once

#include "stitch_common.cuh"

namespace bs {

constexpr int kBlockW = 32;        // one warp: lane = canvas column
constexpr int kRun = 16;           // vertical run: consecutive canvas rows per lane = block height
constexpr int kBlockH = kRun;
constexpr int kMaxLayers = 2;      // cameras that may share one FAST block
constexpr int kMaxPlanCams = 64;
constexpr int kPlanWarps = 8;      // blocks per CTA of the plan kernel

enum BlockType : uint8_t { kBlockEmpty = 0, kBlockFast = 1, kBlockExact = 2 };

// Second-order expansion of one camera's source coordinate field around a block's centre, in the
// form the frame kernel evaluates it (see block_coef / lane_coef / lane_coord_bits below).
struct BlockCoef {                 // 64 bytes; x and y members adjacent, so a kernel can hold them as packed pairs
    int ix, iy;                    // integer bases of the centre coordinate (multiples of 32)
    int8_t nxu, nxv, nyu, nyv;     // whole source texels advanced per canvas column (u) and row (v)
    uint32_t aux;                  // set by the host: first converted texel column | (its byte offset in the staged box) << 16
    float fx, fy;                  // centre coordinate minus its base, in [-0.5, 32.5)
    float xu, yu;                  // first derivatives minus their whole-texel parts (1/32 px per pixel)
    float xv, yv;
    float hxuu, hyuu;              // second derivatives (the pure ones halved)
    float xuv, yuv;
    float hxvv, hyvv;
};
static_assert(sizeof(BlockCoef) == 64, "BlockCoef layout");

// Raw output of the plan kernel, one per block (read back by the host).
struct BlockRec {                  // 160 bytes
    uint16_t bx, by;               // block coordinates (units of 32 x 16 pixels)
    uint8_t type;                  // BlockType
    uint8_t n_layers;
    uint8_t has_uncovered;         // some in-canvas pixel of the block is written by no camera
    uint8_t reserved;
    struct Layer {                 // ascending camera index = compositing order
        int16_t x0, y0, x1, y1;    // source bounding box (texels, taps included) of the pixels it writes
        uint8_t cam;
        uint8_t full;              // writes every pixel of the block and the block lies inside the canvas
        uint8_t groups;            // bit g: rows 4g..4g+3 of the block must use exact coordinates
        uint8_t zero_frac;         // some pixel it writes has both fractions 0 (its top-left weight is 2^15: see tap_weights)
    } layer[kMaxLayers];
    BlockCoef coef[kMaxLayers];
};
static_assert(sizeof(BlockRec) == 160, "BlockRec layout");

// What the frame kernel consumes: one item per (block, camera layer), 80 bytes (five 16-byte
// words), a block's items adjacent and in compositing order.  EXACT and EMPTY blocks are single items.
struct WorkItem {
    uint16_t bx, by;
    uint8_t type;                  // BlockType
    uint8_t flags;                 // kItem* bits
    uint8_t cam;
    uint8_t groups;                // bit g: rows 4g..4g+3 must use exact coordinates
    int16_t c0, y0;                // TMA coordinates of the source box: first 4-byte element, first row
    uint32_t origin_off;           // y0 * stage pitch + 4 * c0: offset of the box origin from texel (0, 0)
    BlockCoef coef;
};
static_assert(sizeof(WorkItem) == 80, "WorkItem layout");
constexpr uint8_t kItemFirst = 1, kItemLast = 2, kItemFull = 4;
// compact kernel only (stitch_fast2.cuh): kItemDirect = single full layer without flagged groups (its rows may go
// straight to the canvas); kItemPlain = kItemDirect, no pixel with both fractions 0, and one source row per
// canvas row (the hot loop's leanest variant); bits 4-5 = (byte offset of the first converted texel in the
// staged row) / 4
constexpr uint8_t kItemDirect = 8, kItemPlain = 0x40;
// BS_FLAG_FEATHER: the item is one of the two layers of a block whose pixels are blended by weight (run_feather2)
constexpr uint8_t kItemFeather = 0x80;
constexpr int kItemRawOffShift = 4;

// ---- cheap coordinates ---------------------------------------------------------------------------
// For camera map M the 1/32-pixel source coordinate of canvas pixel (x, y) is g(x, y) = 32 N / W
// with N, W affine in (x, y).  Around a block's centre (xc, yc), with u = x - xc in [-16, 15] and
// v = y - yc in [-8, 7],
//     g(u, v) ~ g + g_u u + g_v v + g_uu u^2 / 2 + g_uv u v + g_vv v^2 / 2.
// The derivatives follow from g W = 32 N:  g_u = (32 M0 - M6 g) / W,  g_v = (32 M1 - M7 g) / W,
// g_uu = -2 M6 g_u / W,  g_vv = -2 M7 g_v / W,  g_uv = -(M6 g_v + M7 g_u) / W.
// The plan computes them once per block and camera in fp64 (block_coef) and stores them in the work
// item; the whole-texel parts of g and of its first derivatives are split off as integers, so the
// frame kernel evaluates only a small remainder in fp32 -- per lane (lane_coef), then two FFMA per
// coordinate and pixel (lane_coord_bits).  The expansion is only used where the plan has checked,
// pixel by pixel, that it reproduces coord_exact(); every floating-point operation on the way is an
// explicit intrinsic so plan kernel and frame kernel execute the same instruction sequence.
constexpr int kMagicBias = 0x4B400000;   // bits of 12582912.0f = 1.5 * 2^23 (a multiple of 32)

// The expansion in the form the frame kernel evaluates it, from the centre value g = (gx, gy) and its first and
// second derivatives with respect to the canvas column u and row v (h*uu, h*vv: the pure ones already halved).
__device__ __forceinline__ BlockCoef make_block_coef(double gx, double gy, double gxu, double gxv, double gyu, double gyv, double hxuu,
                                                     double xuv, double hxvv, double hyuu, double yuv, double hyvv)
{
    BlockCoef c;
    c.ix = (int)rint(gx) & ~31;
    c.iy = (int)rint(gy) & ~31;
    c.fx = (float)(gx - (double)c.ix);
    c.fy = (float)(gy - (double)c.iy);
    auto whole = [](double d) { return (int)fmin(fmax(rint(d * 0.03125), -127.0), 127.0); };
    c.nxu = (int8_t)whole(gxu);
    c.nxv = (int8_t)whole(gxv);
    c.nyu = (int8_t)whole(gyu);
    c.nyv = (int8_t)whole(gyv);
    c.aux = 0;
    c.xu = (float)(gxu - 32.0 * c.nxu);
    c.xv = (float)(gxv - 32.0 * c.nxv);
    c.yu = (float)(gyu - 32.0 * c.nyu);
    c.yv = (float)(gyv - 32.0 * c.nyv);
    c.hxuu = (float)hxuu;
    c.hxvv = (float)hxvv;
    c.xuv = (float)xuv;
    c.hyuu = (float)hyuu;
    c.hyvv = (float)hyvv;
    c.yuv = (float)yuv;
    return c;
}

__device__ __forceinline__ BlockCoef block_coef(const double *__restrict__ M, int xc, int yc)
{
    const double dx = (double)xc, dy = (double)yc;
    const double Xn = fma(M[1], dy, fma(M[0], dx, M[2]));
    const double Yn = fma(M[4], dy, fma(M[3], dx, M[5]));
    const double Wn = fma(M[7], dy, fma(M[6], dx, M[8]));
    const double inv = 1.0 / Wn;
    const double gx = fmin(fmax(32.0 * Xn * inv, -1.0e9), 1.0e9), gy = fmin(fmax(32.0 * Yn * inv, -1.0e9), 1.0e9);
    const double gxu = (32.0 * M[0] - M[6] * gx) * inv, gxv = (32.0 * M[1] - M[7] * gx) * inv;
    const double gyu = (32.0 * M[3] - M[6] * gy) * inv, gyv = (32.0 * M[4] - M[7] * gy) * inv;
    return make_block_coef(gx, gy, gxu, gxv, gyu, gyv, -M[6] * gxu * inv, -(M[6] * gxv + M[7] * gxu) * inv, -M[7] * gxv * inv,
                           -M[6] * gyu * inv, -(M[6] * gyv + M[7] * gyu) * inv, -M[7] * gyv * inv);
}

// Per-lane form: everything that depends on the column u = lane - 16 only.
struct LaneCoef {
    float a0x, a1x, a2x;   // s_x(v) = a0x + v (a1x + v a2x)
    float a0y, a1y, a2y;
    int bx, by;            // X = bits_x + bx + 32 nxv v   (float-magic bias folded in)
    int nxv, nyv;
};

__device__ __forceinline__ LaneCoef lane_coef(const BlockCoef &c, int u)
{
    const float uf = (float)u;
    LaneCoef l;
    l.a0x = __fmaf_rn(uf, __fmaf_rn(uf, c.hxuu, c.xu), c.fx);
    l.a1x = __fmaf_rn(uf, c.xuv, c.xv);
    l.a2x = c.hxvv;
    l.a0y = __fmaf_rn(uf, __fmaf_rn(uf, c.hyuu, c.yu), c.fy);
    l.a1y = __fmaf_rn(uf, c.yuv, c.yv);
    l.a2y = c.hyvv;
    l.bx = c.ix - kMagicBias + 32 * (int)c.nxu * u;
    l.by = c.iy - kMagicBias + 32 * (int)c.nyu * u;
    l.nxv = c.nxv;
    l.nyv = c.nyv;
    return l;
}

// Lens plans (stitch_undistort.cuh) carry the expansion to third order in the column offset u: a radial lens
// field has a constant third derivative, which the second-order form misses by up to 1/400 of a 1/32-pixel
// step at a block's corners -- enough to flag a third of all blocks.  The u^3, u^2 v and u v^2 terms only
// change the lane coefficients (no cost per pixel); v^3 is left out (|v| <= 8).
struct CubicCoef {                 // 32 bytes per work item, kept beside the item list
    float xuuu, yuuu;              // third derivatives / 6, / 2, / 2
    float xuuv, yuuv;
    float xuvv, yuvv;
    float pad0, pad1;
};
__device__ __forceinline__ LaneCoef lane_coef_cubic(const BlockCoef &c, const CubicCoef &q, int u)
{
    const float uf = (float)u;
    LaneCoef l;
    l.a0x = __fmaf_rn(uf, __fmaf_rn(uf, __fmaf_rn(uf, q.xuuu, c.hxuu), c.xu), c.fx);
    l.a1x = __fmaf_rn(uf, __fmaf_rn(uf, q.xuuv, c.xuv), c.xv);
    l.a2x = __fmaf_rn(uf, q.xuvv, c.hxvv);
    l.a0y = __fmaf_rn(uf, __fmaf_rn(uf, __fmaf_rn(uf, q.yuuu, c.hyuu), c.yu), c.fy);
    l.a1y = __fmaf_rn(uf, __fmaf_rn(uf, q.yuuv, c.yuv), c.yv);
    l.a2y = __fmaf_rn(uf, q.yuvv, c.hyvv);
    l.bx = c.ix - kMagicBias + 32 * (int)c.nxu * u;
    l.by = c.iy - kMagicBias + 32 * (int)c.nyu * u;
    l.nxv = c.nxv;
    l.nyv = c.nyv;
    return l;
}

// Row v = k - kRun/2 of the lane's run.  Returns the raw bits of (s + 1.5*2^23): their low bits are
// rint(s) (round half to even, done by the fp32 adder), so
//   X = bits + l.bx + 32 l.nxv v,   X & 31 = bits & 31,   X >> 5 = (bits >> 5) + (l.bx >> 5) + l.nxv v.
__device__ __forceinline__ void lane_coord_bits(const LaneCoef &l, float vf, int &xbits, int &ybits)
{
    const float sX = __fmaf_rn(vf, __fmaf_rn(vf, l.a2x, l.a1x), l.a0x);
    const float sY = __fmaf_rn(vf, __fmaf_rn(vf, l.a2y, l.a1y), l.a0y);
    xbits = __float_as_int(__fadd_rn(sX, 12582912.0f));
    ybits = __float_as_int(__fadd_rn(sY, 12582912.0f));
}

struct PlanArgs {
    double M[kMaxPlanCams][9];
    int w[kMaxPlanCams], h[kMaxPlanCams];
    int n_cams;
    int W, H;
    int block_w;                   // OpenCV's coordinate block width (not ours)
    int legacy_border;
    int keep_canvas;               // blocks with uncovered pixels then need the read-modify-write path
    int feather;                   // BS_FLAG_FEATHER: a block's layers are all cameras writing any of its pixels (at most two, else per-pixel path)
    int blocks_x, blocks_y;
    BlockRec *out;                 // blocks_y * blocks_x records, row-major
    int *exact_bbox;               // may be null; per camera x0, y0, x1, y1 (init INT_MAX, INT_MAX, -1, -1): the source
                                   // texels the per-pixel blocks can read, so that uploads can still be trimmed to a rectangle
};

// 0 = not written, 1 = interior (Appendix A.3), 2 = written by the last-row/column rule (A.4)
__device__ __forceinline__ int cover_class(int X, int Y, int w, int h, int legacy)
{
    const int sx = sat_s16(X >> kFracBits), sy = sat_s16(Y >> kFracBits);
    if (sx < 0 || sy < 0 || sx > w - 1 || sy > h - 1) return 0;
    if (sx < w - 1 && sy < h - 1) return 1;
    return legacy ? 0 : 2;
}

// One warp per block: lane = column, kRun rows per lane, exactly the frame kernel's mapping.
__global__ void __launch_bounds__(32 * kPlanWarps) plan_kernel(const __grid_constant__ PlanArgs P)
{
    const int lane = threadIdx.x & 31;
    const int bx = blockIdx.x * kPlanWarps + (threadIdx.x >> 5), by = blockIdx.y;
    if (bx >= P.blocks_x) return;
    const unsigned FULLMASK = 0xffffffffu;
    const int x = bx * kBlockW + lane;
    const int y0 = by * kBlockH;
    const int xb = (x / P.block_w) * P.block_w;

    // ---- phase 1: who wins each pixel (highest camera index that writes it, Appendix A.5) -----
    unsigned win_lo = 0u, win_hi = 0u;
    bool edge = false, unc = false;
    if (x < P.W) {
        for (int k = 0; k < kRun; ++k) {
            const int y = y0 + k;
            if (y >= P.H) break;
            if (P.feather) {
                // BS_FLAG_FEATHER: every camera that writes a pixel contributes, so the block's layers are all
                // cameras that write any of its pixels; border-rule pixels and more than kMaxLayers cameras on one
                // pixel go to the per-pixel path
                int n_cov = 0;
                for (int i = 0; i < P.n_cams; ++i) {
                    int X, Y;
                    coord_exact(P.M[i], xb, x - xb, y, X, Y);
                    const int c = cover_class(X, Y, P.w[i], P.h[i], P.legacy_border);
                    if (!c) continue;
                    ++n_cov;
                    if (i < 32) win_lo |= 1u << i;
                    else win_hi |= 1u << (i - 32);
                    if (c == 2) edge = true;
                }
                if (n_cov == 0) unc = true;
                if (n_cov > kMaxLayers) edge = true;
                continue;
            }
            int cls = 0, i = P.n_cams - 1;
            for (; i >= 0; --i) {
                int X, Y;
                coord_exact(P.M[i], xb, x - xb, y, X, Y);
                cls = cover_class(X, Y, P.w[i], P.h[i], P.legacy_border);
                if (cls) break;
            }
            if (cls == 0) unc = true;
            else {
                if (i < 32) win_lo |= 1u << i;
                else win_hi |= 1u << (i - 32);
                if (cls == 2) edge = true;
            }
        }
    }
    win_lo = __reduce_or_sync(FULLMASK, win_lo);
    win_hi = __reduce_or_sync(FULLMASK, win_hi);
    edge = __any_sync(FULLMASK, edge);
    unc = __any_sync(FULLMASK, unc);

    BlockRec rec;
    memset(&rec, 0, sizeof rec);
    rec.bx = (uint16_t)bx;
    rec.by = (uint16_t)by;
    rec.has_uncovered = unc ? 1 : 0;
    BlockRec *dst = P.out + (size_t)by * P.blocks_x + bx;
    const int n_layers = __popc(win_lo) + __popc(win_hi);
    if (n_layers == 0) {
        rec.type = kBlockEmpty;
        if (lane == 0) *dst = rec;
        return;
    }
    if (n_layers > kMaxLayers || edge || (P.keep_canvas && unc)) {
        rec.type = kBlockExact;
        if (lane == 0) *dst = rec;
        if (P.exact_bbox) {
            // What the per-pixel path can read here: the taps of every camera in the mask (the winners; with
            // BS_FLAG_FEATHER all cameras that write) at the pixels that camera covers -- a superset, which is fine.
            unsigned long long mm = ((unsigned long long)win_hi << 32) | win_lo;
            while (mm) {
                const int c = __ffsll((long long)mm) - 1;
                mm &= mm - 1;
                int bx0 = INT_MAX, by0 = INT_MAX, bx1 = -1, by1 = -1;
                if (x < P.W)
                    for (int k = 0; k < kRun && y0 + k < P.H; ++k) {
                        int X, Y;
                        coord_exact(P.M[c], xb, x - xb, y0 + k, X, Y);
                        if (!cover_class(X, Y, P.w[c], P.h[c], P.legacy_border)) continue;
                        const int sx = X >> kFracBits, sy = Y >> kFracBits;
                        bx0 = min(bx0, sx);
                        by0 = min(by0, sy);
                        bx1 = max(bx1, min(sx + 1, P.w[c] - 1));
                        by1 = max(by1, min(sy + 1, P.h[c] - 1));
                    }
                bx0 = __reduce_min_sync(FULLMASK, bx0);
                by0 = __reduce_min_sync(FULLMASK, by0);
                bx1 = __reduce_max_sync(FULLMASK, bx1);
                by1 = __reduce_max_sync(FULLMASK, by1);
                if (lane == 0 && bx1 >= 0) {
                    atomicMin(P.exact_bbox + 4 * c + 0, bx0);
                    atomicMin(P.exact_bbox + 4 * c + 1, by0);
                    atomicMax(P.exact_bbox + 4 * c + 2, bx1);
                    atomicMax(P.exact_bbox + 4 * c + 3, by1);
                }
            }
        }
        return;
    }
    // ---- phase 2: per layer, source bounding box and validity of the cheap coordinates -----------
    unsigned long long m = ((unsigned long long)win_hi << 32) | win_lo;
    rec.type = kBlockFast;
    rec.n_layers = (uint8_t)n_layers;
    for (int l = 0; l < n_layers; ++l) {
        const int c = __ffsll((long long)m) - 1;
        m &= m - 1;
        int bx0 = INT_MAX, by0 = INT_MAX, bx1 = INT_MIN, by1 = INT_MIN;
        bool all_in = (x < P.W) && (y0 + kRun <= P.H);
        bool zf = false;
        unsigned mism = 0u;
        const BlockCoef bc = block_coef(P.M[c], bx * kBlockW + kBlockW / 2, y0 + kRun / 2);
        rec.coef[l] = bc;
        if (x < P.W) {
            const LaneCoef lc = lane_coef(bc, lane - kBlockW / 2);
            for (int k = 0; k < kRun; ++k) {
                const int y = y0 + k;
                if (y >= P.H) break;
                int X, Y, xbits, ybits;
                coord_exact(P.M[c], xb, x - xb, y, X, Y);
                const int v = k - kRun / 2;
                lane_coord_bits(lc, (float)v, xbits, ybits);
                if (X != xbits + lc.bx + 32 * lc.nxv * v || Y != ybits + lc.by + 32 * lc.nyv * v) mism |= 1u << (k >> 2);
                if (cover_class(X, Y, P.w[c], P.h[c], 1) == 1) {   // interior only
                    const int sx = X >> kFracBits, sy = Y >> kFracBits;
                    zf = zf || ((X | Y) & (kFracOne - 1)) == 0;
                    bx0 = min(bx0, sx);
                    by0 = min(by0, sy);
                    bx1 = max(bx1, sx + 1);
                    by1 = max(by1, sy + 1);
                } else {
                    all_in = false;
                }
            }
        }
        bx0 = __reduce_min_sync(FULLMASK, bx0);
        by0 = __reduce_min_sync(FULLMASK, by0);
        bx1 = __reduce_max_sync(FULLMASK, bx1);
        by1 = __reduce_max_sync(FULLMASK, by1);
        mism = __reduce_or_sync(FULLMASK, mism);
        all_in = __all_sync(FULLMASK, all_in);
        rec.layer[l].x0 = (int16_t)bx0;     // every layer wins, hence interior-covers, at least one pixel
        rec.layer[l].y0 = (int16_t)by0;
        rec.layer[l].x1 = (int16_t)bx1;
        rec.layer[l].y1 = (int16_t)by1;
        rec.layer[l].cam = (uint8_t)c;
        rec.layer[l].full = all_in ? 1 : 0;
        rec.layer[l].groups = (uint8_t)mism;
        rec.layer[l].zero_frac = __any_sync(FULLMASK, zf) ? 1 : 0;
    }
    if (lane == 0) *dst = rec;
}

}  // namespace bs
