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

// Raw output of the plan kernel, one per block (read back by the host).
struct BlockRec {                  // 32 bytes
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
        uint8_t pad;
    } layer[kMaxLayers];
};
static_assert(sizeof(BlockRec) == 32, "BlockRec layout");

// What the frame kernel consumes: one item per (block, camera layer), 16 bytes, a block's items
// adjacent and in compositing order.  EXACT and EMPTY blocks are single items.
struct WorkItem {
    uint16_t bx, by;
    uint8_t type;                  // BlockType
    uint8_t flags;                 // kItem* bits
    uint8_t cam;
    uint8_t groups;                // bit g: rows 4g..4g+3 must use exact coordinates
    int16_t c0, y0;                // TMA coordinates of the source box: first 4-byte element, first row
    uint32_t origin_off;           // y0 * stage pitch + 4 * c0: offset of the box origin from texel (0, 0)
};
static_assert(sizeof(WorkItem) == 16, "WorkItem layout");
constexpr uint8_t kItemFirst = 1, kItemLast = 2, kItemFull = 4;

// ---- cheap coordinates for a vertical run of kRun pixels (one canvas column) -----------------
// Along a column the source coordinate is fX(yc + j) = fXc + j (32 M1 - M7 fXc) / (Wc + j M7),
// j = k - kRun/2.  The run is anchored at its 9th pixel in fp64; the increment is expanded to
// second order in j.  Its whole-texel part, 32 n per step with n = round(dX / 32), is exact integer
// arithmetic and goes straight into the address; only the small remainder is evaluated in fp32:
//     sX(j) = fx + j (Dx + j Ex),   X(j) = rint(sX(j)) + base_x + 32 n_x j.
// The result is only used where the plan has checked, pixel by pixel, that it equals
// coord_exact(); every floating-point operation is an explicit intrinsic (or PTX) so the plan
// kernel and the frame kernel execute the same instruction sequence.
struct RunAnchor {
    int bx, by;        // integer bases (multiples of 32) minus the float-magic bias
    int nx, ny;        // whole source texels advanced per canvas row
    float fx, fy;      // fXc - base, in [-0.5, 32.5)
    float Dx, Ex, Dy, Ey;
};

constexpr int kMagicBias = 0x4B400000;   // bits of 12582912.0f = 1.5 * 2^23 (a multiple of 32)

// 1/w to a few ulp: hardware seed (about 20 bits) and two Newton steps, no IEEE special cases.
// A bad value (w = 0, denormal, ...) only makes the plan reject the run.
__device__ __forceinline__ double fast_rcp(double w)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(w));
    r = fma(r, fma(-w, r, 1.0), r);
    r = fma(r, fma(-w, r, 1.0), r);
    return r;
}

__device__ __forceinline__ RunAnchor run_anchor(const double *__restrict__ M, int x, int yc)
{
    const double dx = (double)x, dy = (double)yc;
    const double Xn = fma(M[1], dy, fma(M[0], dx, M[2]));
    const double Yn = fma(M[4], dy, fma(M[3], dx, M[5]));
    const double Wn = fma(M[7], dy, fma(M[6], dx, M[8]));
    const double inv = fast_rcp(Wn);
    const double inv32 = inv * 32.0;
    const double fX = Xn * inv32, fY = Yn * inv32;
    const double dX = fma(-M[7], fX, 32.0 * M[1]) * inv;   // d fX / dj at j = 0
    const double dY = fma(-M[7], fY, 32.0 * M[4]) * inv;
    const float e = (float)(-M[7] * inv);
    const float dXf = (float)dX, dYf = (float)dY;
    RunAnchor a;
    a.nx = __float2int_rn(dXf * 0.03125f);
    a.ny = __float2int_rn(dYf * 0.03125f);
    const int ix = __double2int_rn(fX) & ~31, iy = __double2int_rn(fY) & ~31;
    a.bx = ix - kMagicBias;
    a.by = iy - kMagicBias;
    a.fx = (float)(fX - (double)ix);
    a.fy = (float)(fY - (double)iy);
    a.Dx = (float)(dX - (double)(32 * a.nx));
    a.Dy = (float)(dY - (double)(32 * a.ny));
    a.Ex = __fmul_rn(dXf, e);
    a.Ey = __fmul_rn(dYf, e);
    return a;
}

// k in [0, kRun): pixel row yc - kRun/2 + k.  Returns the raw bits of (s + 1.5*2^23): their low
// bits are rint(s) (round half to even, done by the fp32 adder), so with j = k - kRun/2
//   X = bits + a.bx + 32 a.nx j,   X & 31 = bits & 31,   X >> 5 = (bits >> 5) + (a.bx >> 5) + a.nx j.
__device__ __forceinline__ void run_coord_bits(const RunAnchor &a, float jf, int &xbits, int &ybits)
{
    const float sX = __fmaf_rn(jf, __fmaf_rn(jf, a.Ex, a.Dx), a.fx);
    const float sY = __fmaf_rn(jf, __fmaf_rn(jf, a.Ey, a.Dy), a.fy);
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
    int blocks_x, blocks_y;
    BlockRec *out;                 // blocks_y * blocks_x records, row-major
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
        unsigned mism = 0u;
        if (x < P.W) {
            const RunAnchor a = run_anchor(P.M[c], x, y0 + kRun / 2);
            for (int k = 0; k < kRun; ++k) {
                const int y = y0 + k;
                if (y >= P.H) break;
                int X, Y, xbits, ybits;
                coord_exact(P.M[c], xb, x - xb, y, X, Y);
                const int j = k - kRun / 2;
                run_coord_bits(a, (float)j, xbits, ybits);
                if (X != xbits + a.bx + 32 * a.nx * j || Y != ybits + a.by + 32 * a.ny * j) mism |= 1u << (k >> 2);
                if (cover_class(X, Y, P.w[c], P.h[c], 1) == 1) {   // interior only
                    const int sx = X >> kFracBits, sy = Y >> kFracBits;
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
    }
    if (lane == 0) *dst = rec;
}

}  // namespace bs
