// stitch_plan.cuh -- the tile plan: static per calibration, built once on the device.
//
// The reference recomputes everything per frame (cv::warpPerspective regenerates the coordinate
// maps on every call, /root/reference/tools/stitch/main.cpp:150-151), but H and result_size are
// fixed for the whole run (:53-73).  The plan classifies every 128x32 canvas tile once:
//   EMPTY  no camera writes any pixel                      -> zero fill
//   FAST   every written pixel is an interior (4-tap) pixel of one of <= 4 cameras whose source
//          bounding boxes fit a shared-memory stage        -> staged kernel, one pass per camera
//   EXACT  anything else (last-row/column pixels, too many cameras, huge boxes) -> per-pixel kernel
// and, for FAST tiles, proves by exhaustive comparison which 8-pixel runs may use the cheap
// fp32 coordinate recurrence (fast_run_coords) instead of the exact fp64 form.
#pragma once

#include "stitch_common.cuh"

namespace bs {

constexpr int kTileW = 128;
#ifndef BS_BLOCKS_PER_TILE
#define BS_BLOCKS_PER_TILE 8
#endif
constexpr int kBlocksPerTile = BS_BLOCKS_PER_TILE;   // 32x16 blocks per tile = consumer warps per CTA (4: one row of blocks)
constexpr int kThreads = 32 * kBlocksPerTile;
constexpr int kRun = 16;           // vertical run: consecutive rows per thread
constexpr int kTileH = kRun * (kBlocksPerTile / 4);
constexpr int kMaxLayers = 4;      // cameras that may share one FAST tile
constexpr int kMaxPlanCams = 64;

enum TileType : uint8_t { kTileEmpty = 0, kTileFast = 1, kTileExact = 2 };

struct TileRec {                   // 80 bytes, read by the producer as five 16-byte loads
    uint16_t tx, ty;               // tile coordinates
    uint8_t type;                  // TileType
    uint8_t n_layers;
    uint8_t has_uncovered;         // some in-canvas pixel of the tile is written by no layer
    uint8_t reserved;
    uint8_t cam[kMaxLayers];       // ascending camera index = compositing order
    // per layer, bit w = 32x16 block w of the tile (block w belongs to warp w of the kernels):
    uint8_t all_mask[kMaxLayers];  //   the block lies inside the canvas and the layer writes every pixel of it
    uint8_t none_mask[kMaxLayers]; //   the layer writes no pixel of the block
    int16_t bx0[kMaxLayers], by0[kMaxLayers], bx1[kMaxLayers], by1[kMaxLayers];  // texel bbox incl. taps
    uint32_t exact_mask[kMaxLayers];   // bit (w*4 + row group): those 32x4 pixels must use exact coords
    uint32_t pad[3];
};
static_assert(sizeof(TileRec) == 80, "TileRec layout");

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

// thread -> pixels of a tile, shared by the plan and the staged kernel: thread t owns column
// (t & 127) and the kRun rows starting at kRun * (t >> 7).  A warp is 32 adjacent columns, so its
// shared-memory gathers touch one contiguous span per source row (no bank conflicts).
__device__ __forceinline__ int tile_col(int tid) { return tid & (kTileW - 1); }
__device__ __forceinline__ int tile_row0(int tid) { return (tid >> 7) * kRun; }
// bit of TileRec::exact_mask covering row k of thread tid's run: (warp, group of 4 rows)
__device__ __forceinline__ int exact_bit(int tid, int k) { return (tid >> 5) * 4 + (k >> 2); }
// (warp w of the plan kernel covers the same 32x16 block as consumer warp w of the frame kernel)

struct PlanArgs {
    double M[kMaxPlanCams][9];
    int w[kMaxPlanCams], h[kMaxPlanCams];
    int n_cams;
    int W, H;
    int block_w;
    int legacy_border;
    int tiles_x, tiles_y;
    TileRec *out;
};

// 0 = not written, 1 = interior (Appendix A.3), 2 = written by the last-row/column rule (A.4)
__device__ __forceinline__ int cover_class(int X, int Y, int w, int h, int legacy)
{
    const int sx = sat_s16(X >> kFracBits), sy = sat_s16(Y >> kFracBits);
    if (sx < 0 || sy < 0 || sx > w - 1 || sy > h - 1) return 0;
    if (sx < w - 1 && sy < h - 1) return 1;
    return legacy ? 0 : 2;
}

// One CTA per tile.
__global__ void __launch_bounds__(kThreads) plan_kernel(const __grid_constant__ PlanArgs P)
{
    __shared__ unsigned long long s_win;       // cameras that win at least one pixel
    __shared__ int s_edge_win, s_uncovered;
    __shared__ int s_bx0[kMaxLayers], s_by0[kMaxLayers], s_bx1[kMaxLayers], s_by1[kMaxLayers];
    __shared__ unsigned s_all[kMaxLayers], s_none[kMaxLayers];
    __shared__ unsigned s_exact[kMaxLayers];

    const int tid = threadIdx.x;
    const int tx = blockIdx.x, ty = blockIdx.y;
    if (tid == 0) {
        s_win = 0ull;
        s_edge_win = 0;
        s_uncovered = 0;
    }
    if (tid < kMaxLayers) {
        s_bx0[tid] = s_by0[tid] = INT_MAX;
        s_bx1[tid] = s_by1[tid] = INT_MIN;
        s_all[tid] = 0u;
        s_none[tid] = 0u;
        s_exact[tid] = 0u;
    }
    __syncthreads();

    const int x = tx * kTileW + tile_col(tid);
    const int y0 = ty * kTileH + tile_row0(tid);
    const int xb = (x / P.block_w) * P.block_w;
    // ---- phase 1: who wins each pixel -----------------------------------------------------
    unsigned long long my_win = 0ull;
    int my_edge = 0, my_unc = 0;
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
            if (cls == 0) my_unc = 1;
            else {
                my_win |= 1ull << i;
                if (cls == 2) my_edge = 1;
            }
        }
    }
    if (my_win) atomicOr(&s_win, my_win);
    if (my_edge) atomicOr(&s_edge_win, 1);
    if (my_unc) atomicOr(&s_uncovered, 1);
    __syncthreads();

    TileRec rec;
    memset(&rec, 0, sizeof rec);
    rec.tx = (uint16_t)tx;
    rec.ty = (uint16_t)ty;
    const unsigned long long win = s_win;
    const int n_layers = __popcll(win);
    rec.has_uncovered = (uint8_t)s_uncovered;
    TileRec *dst = P.out + (size_t)ty * P.tiles_x + tx;
    if (n_layers == 0) {
        rec.type = kTileEmpty;
        if (tid == 0) *dst = rec;
        return;
    }
    if (n_layers > kMaxLayers || s_edge_win) {
        rec.type = kTileExact;
        if (tid == 0) *dst = rec;
        return;
    }
    int cams[kMaxLayers];
    {
        unsigned long long m = win;
        for (int l = 0; l < n_layers; ++l) {
            cams[l] = __ffsll((long long)m) - 1;
            m &= m - 1;
        }
    }
    // ---- phase 2: per layer, source bounding box and validity of the cheap coordinates -----
    for (int l = 0; l < n_layers; ++l) {
        const int c = cams[l];
        int bx0 = INT_MAX, by0 = INT_MAX, bx1 = INT_MIN, by1 = INT_MIN;
        bool all_in = (x < P.W) && (y0 + kRun <= P.H), any_in = false;
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
                if (X != xbits + a.bx + 32 * a.nx * j || Y != ybits + a.by + 32 * a.ny * j)
                    mism |= 1u << exact_bit(tid, k);
                if (cover_class(X, Y, P.w[c], P.h[c], 1) == 1) {   // interior only
                    const int sx = X >> kFracBits, sy = Y >> kFracBits;
                    bx0 = min(bx0, sx);
                    by0 = min(by0, sy);
                    bx1 = max(bx1, sx + 1);
                    by1 = max(by1, sy + 1);
                    any_in = true;
                } else {
                    all_in = false;
                }
            }
        }
        if (mism) atomicOr(&s_exact[l], mism);
        if (bx0 != INT_MAX) {
            atomicMin(&s_bx0[l], bx0);
            atomicMin(&s_by0[l], by0);
            atomicMax(&s_bx1[l], bx1);
            atomicMax(&s_by1[l], by1);
        }
        // block-level coverage (a warp is exactly one 32x16 block)
        const bool blk_all = __all_sync(0xffffffffu, all_in), blk_any = __any_sync(0xffffffffu, any_in);
        if ((tid & 31) == 0) {
            if (blk_all) atomicOr(&s_all[l], 1u << (tid >> 5));
            if (!blk_any) atomicOr(&s_none[l], 1u << (tid >> 5));
        }
    }
    __syncthreads();
    if (tid == 0) {
        rec.type = kTileFast;
        rec.n_layers = (uint8_t)n_layers;
        for (int l = 0; l < n_layers; ++l) {
            rec.cam[l] = (uint8_t)cams[l];
            rec.bx0[l] = (int16_t)s_bx0[l];
            rec.by0[l] = (int16_t)s_by0[l];
            rec.bx1[l] = (int16_t)s_bx1[l];
            rec.by1[l] = (int16_t)s_by1[l];
            rec.exact_mask[l] = s_exact[l];
            rec.all_mask[l] = (uint8_t)s_all[l];
            rec.none_mask[l] = (uint8_t)s_none[l];
        }
        *dst = rec;
    }
}

}  // namespace bs
