// stitch_fast.cuh -- the per-frame kernel: one persistent CTA loop over the planned tiles.
//
// For a FAST tile and each of its cameras (lowest index first, so later cameras overwrite
// earlier ones exactly as the reference's loop does, /root/reference/tools/stitch/main.cpp:145-152):
//   1. the camera's source bounding box is prefetched into shared memory by one TMA tensor copy
//      (cp.async.bulk.tensor.2d, mbarrier-signalled, double buffered across tiles);
//   2. each thread computes two runs of 8 canvas pixels: coordinates from the plan-verified fp32
//      recurrence (or exact fp64 where the plan says so), 2x3 aligned words per pixel from the
//      staged box, OpenCV's fixed-point bilinear weights from a 1024-entry table, three
//      channel dot products (dp4a) -- bit-identical to (sum w*S + 2^14) >> 15;
//   3. the finished 128x32 tile goes from shared memory to the canvas as 16-byte vector stores.
// EXACT tiles run the per-pixel fp64 path of stitch_exact.cuh; EMPTY tiles are zero-filled.
#pragma once

#include <cuda.h>   // CUtensorMap (type only; the encoder is fetched at run time)

#include "stitch_exact.cuh"
#include "stitch_plan.cuh"

namespace bs {

constexpr int kOutTileBytes = kTileW * kTileH * 3;   // 12288
constexpr int kLutBytes = 1024 * 8;
constexpr int kStagePad = 128;                       // slack behind a stage for the 3rd word of the last texel

struct FastArgs {
    FrameArgs F;
    CUtensorMap tmap[kMaxCamsPerLaunch];   // source frame of camera i as a [h][pitch/4] u32 tensor
    const TileRec *tiles;
    int n_tiles;
    int box_we[kMaxCamsPerLaunch];         // TMA box of camera i: width in 4-byte elements, rows
    int box_h[kMaxCamsPerLaunch];
    int stage_bytes;                       // shared memory per stage (multiple of 128)
};

// First 4-byte element of the staged box for a bounding box whose first texel column is bx0.
// TMA requires the box to start on a 16-byte boundary in global memory (an unaligned start traps
// with "illegal instruction"), so the start is rounded down to a multiple of 4 elements.
__host__ __device__ __forceinline__ int box_elem0(int bx0) { return ((3 * bx0) >> 4) << 2; }

// ---- PTX helpers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tmap, int c0, int c1, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(dst)),
        "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// Weight table entry for fractional position (a, b), a = X & 31, b = Y & 31.
// w' = w / 32 are OpenCV's 15-bit weights without their common factor 32; they sum to 1024.
// Entry = { bytes of w' >> 2, bytes of (w' & 3) << 6 } in tap order 00, 01, 10, 11, so that
//   256 * dp4a(S, hi) + dp4a(S, lo) = 64 * sum(w' S),  and byte 2 of (that + 2^15) is the pixel.
// (a, b) = (0, 0) would need w'00 = 1024 (9 bits after >> 2); 1023 gives the same pixel for
// every S in 0..255: (1023 S + 512) >> 10 == S.
__device__ __forceinline__ uint2 lut_entry(int idx)
{
    const int a = idx & 31, b = idx >> 5;
    int w[4] = {(32 - a) * (32 - b), a * (32 - b), (32 - a) * b, a * b};
    if (w[0] == 1024) w[0] = 1023;
    uint2 e = make_uint2(0u, 0u);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        e.x |= (uint32_t)(w[t] >> 2) << (8 * t);
        e.y |= (uint32_t)((w[t] & 3) << 6) << (8 * t);
    }
    return e;
}

// One interior pixel from a staged box.  `o` = shared byte address of texel (sx, sy) (any
// alignment), `pitch` = stage row pitch in bytes, `lut_off` = shared address of the weight-table
// entry.  Returns B | G<<8 | R<<16.
__device__ __forceinline__ uint32_t sample_staged(uint32_t o, int pitch, uint32_t lut_off)
{
    const uint32_t wa = o & ~3u;
    const uint32_t sh = o << 3;          // the funnel shift uses the low 5 bits: 8 * (o & 3)
    const uint32_t w0 = lds32(wa), w1 = lds32(wa + 4), w2 = lds32(wa + 8);
    const uint32_t v0 = lds32(wa + pitch), v1 = lds32(wa + pitch + 4), v2 = lds32(wa + pitch + 8);
    const uint2 wt = lds64(lut_off);
    // six aligned bytes of each row: r?a = [B0 G0 R0 B1], r?b = [G1 R1 . .]
    const uint32_t r0a = __funnelshift_r(w0, w1, sh), r0b = __funnelshift_r(w1, w2, sh);
    const uint32_t r1a = __funnelshift_r(v0, v1, sh), r1b = __funnelshift_r(v1, v2, sh);
    // transpose to one word per channel, taps in order 00 01 10 11
    const uint32_t qb = __byte_perm(r0a, r1a, 0x7430);
    const uint32_t t0 = __byte_perm(r0a, r0b, 0x5241);   // [G00 G01 R00 R01]
    const uint32_t t1 = __byte_perm(r1a, r1b, 0x5241);   // [G10 G11 R10 R11]
    const uint32_t qg = __byte_perm(t0, t1, 0x5410);
    const uint32_t qr = __byte_perm(t0, t1, 0x7632);
    const uint32_t vb = __dp4a(qb, wt.x, 128u) * 256u + __dp4a(qb, wt.y, 0u);
    const uint32_t vg = __dp4a(qg, wt.x, 128u) * 256u + __dp4a(qg, wt.y, 0u);
    const uint32_t vr = __dp4a(qr, wt.x, 128u) * 256u + __dp4a(qr, wt.y, 0u);
    // results sit in byte 2 of vb, vg, vr; byte 3 is always 0
    const uint32_t bg = __byte_perm(vb, vg, 0x3362);
    return __byte_perm(bg, vr, 0x7610);
}

// Store one row's pixel of every lane into the output tile: lanes 4q..4q+3 hold 4 adjacent
// pixels = 12 bytes = 3 words; lane 4q+i (i < 3) builds word i from its own pixel and its right
// neighbour's and stores it.  `row_addr` = shared address of the tile row, `keep` = this lane's
// word must not be stored (used by partially covered layers, which fall back to byte stores).
__device__ __forceinline__ void store_row_pixel(uint32_t px, uint32_t row_addr, int lane, uint32_t sel)
{
    const uint32_t nxt = __shfl_down_sync(0xffffffffu, px, 1);
    const uint32_t word = __byte_perm(px, nxt, sel);
    if ((lane & 3) != 3)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(row_addr + 4u * (uint32_t)(3 * (lane >> 2) + (lane & 3))), "r"(word)
                     : "memory");
}

// Shared address of the first output-tile row of thread tid's run, at the start of its warp's
// 96-byte (32-pixel) span.
__device__ __forceinline__ uint32_t out_span(uint32_t out_tile, int tid)
{
    return out_tile + (uint32_t)(tile_row0(tid) * (kTileW * 3) + ((tid & (kTileW - 1)) >> 5) * 96);
}

// Hot path: the layer covers every pixel of the tile with interior samples, the tile lies inside
// the canvas and the plan verified the cheap coordinates for this warp's whole 32x16 block.
// No coverage tests, no fp64 beyond the run's anchor; fully unrolled so k is a constant.
__device__ __forceinline__ void run_layer_full(const CamDev &cam, int pitch, uint32_t stage_origin, uint32_t lut_base,
                                               uint32_t out_tile, int x, int y0, int tid)
{
    const int lane = tid & 31;
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const uint32_t out_col = out_span(out_tile, tid);
    const RunAnchor a = run_anchor(cam.M, x, y0 + kRun / 2);
    const uint32_t run_base = stage_origin + (uint32_t)((a.by >> kFracBits) * pitch + (a.bx >> kFracBits) * 3);
#pragma unroll
    for (int k = 0; k < kRun; ++k) {
        int xbits, ybits;
        run_coord_bits(a, k, xbits, ybits);
        const uint32_t o = run_base + (uint32_t)((ybits >> kFracBits) * pitch + (xbits >> kFracBits) * 3);
        const uint32_t lut_off = lut_base + (uint32_t)(((ybits & 31) << 8) + ((xbits & 31) << 3));
        const uint32_t px = sample_staged(o, pitch, lut_off);
        store_row_pixel(px, out_col + (uint32_t)(k * (kTileW * 3)), lane, sel);
    }
}

// General path: partially covered layers, tiles that stick out of the canvas, and warps for which
// the plan found at least one pixel where the cheap coordinates differ (those 32x4 groups use
// the exact fp64 form).  Same arithmetic per pixel, with coverage tests and byte stores.
__device__ __noinline__ void run_layer_general(const FastArgs &A, const CamDev &cam, int pitch, uint32_t stage_origin,
                                               uint32_t lut_base, uint32_t out_tile, uint32_t exact_mask, bool full,
                                               bool first_layer, int x, int y0, int tid)
{
    const int lane = tid & 31;
    const uint32_t out_col = out_span(out_tile, tid);
    const RunAnchor a = run_anchor(cam.M, x, y0 + kRun / 2);
    const int xb = (x / A.F.block_w) * A.F.block_w;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        int X, Y;
        if ((exact_mask >> exact_bit(tid, k)) & 1) {     // warp-uniform
            coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        } else {
            // same instruction sequence as run_coord_bits, with k at run time
            const float kf = (float)(k - kRun / 2);
            const float sX = __fmaf_rn(kf, __fmaf_rn(kf, a.Ex, a.Dx), a.fx);
            const float sY = __fmaf_rn(kf, __fmaf_rn(kf, a.Ey, a.Dy), a.fy);
            X = __float_as_int(__fadd_rn(sX, 12582912.0f)) + a.bx;
            Y = __float_as_int(__fadd_rn(sY, 12582912.0f)) + a.by;
        }
        const int sx = X >> kFracBits, sy = Y >> kFracBits;
        bool in = (x < A.F.W) && (y0 + k < A.F.H);
        if (!full) in = in && sx >= 0 && sy >= 0 && sx < cam.w - 1 && sy < cam.h - 1;
        const uint32_t o = in ? stage_origin + (uint32_t)(sy * pitch + sx * 3) : out_tile;   // any valid address
        const uint32_t lut_off = lut_base + (uint32_t)((((Y & 31) << 5) | (X & 31)) << 3);
        uint32_t px = sample_staged(o, pitch, lut_off);
        // the first layer also defines the background: pixels it does not cover start at 0
        if (in || first_layer) {
            if (!in) px = 0;
            const uint32_t p8 = out_col + (uint32_t)(k * (kTileW * 3)) + 3u * (uint32_t)lane;
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8), "r"(px) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 1), "r"(px >> 8) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 2), "r"(px >> 16) : "memory");
        }
    }
}

__global__ void __launch_bounds__(kThreads, 3) stitch_fast_kernel(const __grid_constant__ FastArgs A)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *const out_base = smem + 2 * A.stage_bytes;              // two output tiles
    uint2 *const lut = reinterpret_cast<uint2 *>(out_base + 2 * kOutTileBytes);
    uint64_t *const bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(lut) + kLutBytes);

    const int tid = threadIdx.x;
    for (int i = tid; i < 1024; i += kThreads) lut[i] = lut_entry(i);
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t lut_base = smem_u32(lut);

    // ---- prefetch cursor (thread 0 only): walks the (tile, layer) items of this CTA ------------
    int pf_k = blockIdx.x, pf_layer = 0;     // next item to issue
    uint32_t n_issued = 0;
    auto issue_next = [&]() {                // thread 0: issue the TMA of the next FAST item, if any
        while (pf_k < A.n_tiles) {
            const TileRec *r = &A.tiles[pf_k];
            const bool fast = (r->type == kTileFast) && !(A.F.keep_canvas && r->has_uncovered);
            if (fast && pf_layer < r->n_layers) {
                const int c = r->cam[pf_layer];
                const int s = n_issued & 1;
                const uint32_t bytes = (uint32_t)(A.box_we[c] * 4 * A.box_h[c]);
                mbar_expect_tx(&bars[s], bytes);
                tma_load_2d(smem + s * A.stage_bytes, &A.tmap[c], box_elem0((int)r->bx0[pf_layer]),
                            (int)r->by0[pf_layer], &bars[s]);
                ++n_issued;
                ++pf_layer;
                return;
            }
            pf_k += gridDim.x;
            pf_layer = 0;
        }
    };
    if (tid == 0) {
        issue_next();
        issue_next();
    }

    uint32_t n_consumed = 0;
    int out_sel = 0;
    for (int k = blockIdx.x; k < A.n_tiles; k += gridDim.x) {
        const TileRec *const rec = &A.tiles[k];
        const int tx0 = rec->tx * kTileW, ty0 = rec->ty * kTileH;
        const int n_layers = rec->n_layers;
        const int rec_type = rec->type;
        const bool fast = (rec_type == kTileFast) && !(A.F.keep_canvas && rec->has_uncovered);
        // tiles that stick out of the canvas: their outside pixels must not be sampled
        const bool edge_tile = (tx0 + kTileW > A.F.W) || (ty0 + kTileH > A.F.H);
        if (!fast) {
            if (rec_type == kTileEmpty) {
                if (!A.F.keep_canvas) {
                    // zero fill: rows of the tile, 16-byte chunks where aligned
                    const int wbytes = min(kTileW, A.F.W - tx0) * 3;
                    for (int q = tid; q < kTileH * 24; q += kThreads) {
                        const int row = q / 24, cb = (q % 24) * 16;
                        if (ty0 + row >= A.F.H || cb >= wbytes) continue;
                        uint8_t *p = A.F.canvas + (long long)(ty0 + row) * A.F.cpitch + (long long)tx0 * 3 + cb;
                        if (cb + 16 <= wbytes) *reinterpret_cast<uint4 *>(p) = make_uint4(0, 0, 0, 0);
                        else
                            for (int j = 0; j < wbytes - cb; ++j) p[j] = 0;
                    }
                }
            } else {
                // exact tile: thread t takes the 4-pixel group (t & 31) of rows (t >> 5) + 8 j
                for (int j = 0; j < kTileH / 8; ++j) exact_quad(A.F, tx0 + (tid & 31) * 4, ty0 + (tid >> 5) + 8 * j);
            }
            continue;
        }

        const uint32_t outt = smem_u32(out_base + out_sel * kOutTileBytes);
        const int x = tx0 + tile_col(tid);
        const int y0 = ty0 + tile_row0(tid);
        for (int l = 0; l < n_layers; ++l) {
            const int c = rec->cam[l];
            const CamDev &cam = A.F.cam[c];
            const int s = n_consumed & 1;
            const int pitch = A.box_we[c] * 4;
            const bool full = ((rec->full_mask >> l) & 1) && !edge_tile;
            const uint32_t exact_mask = rec->exact_mask[l];
            // shared address texel (0, 0) of this camera would have, were the whole frame staged
            const uint32_t stage_origin = smem_u32(smem + s * A.stage_bytes) -
                                          (uint32_t)((int)rec->by0[l] * pitch + 4 * box_elem0((int)rec->bx0[l]));
            mbar_wait(&bars[s], (n_consumed >> 1) & 1);
            const uint32_t warp_mask = (exact_mask >> ((tid >> 5) * 4)) & 15u;    // this warp's 4 row groups
            if (full && warp_mask == 0u) run_layer_full(cam, pitch, stage_origin, lut_base, outt, x, y0, tid);
            else run_layer_general(A, cam, pitch, stage_origin, lut_base, outt, exact_mask, full, l == 0, x, y0, tid);
            ++n_consumed;
            __syncthreads();                 // stage s fully read; this layer's pixels are in the tile
            if (tid == 0) issue_next();      // refill stage s while the next item computes
        }
        // ---- flush the tile: 32 rows x 384 bytes as 16-byte vector stores -----------------------
        {
            const uint8_t *const ot = out_base + out_sel * kOutTileBytes;
            const int wbytes = min(kTileW, A.F.W - tx0) * 3;
#pragma unroll
            for (int q = tid; q < kTileH * 24; q += kThreads) {
                const int row = q / 24, cb = (q % 24) * 16;
                if (ty0 + row >= A.F.H || cb >= wbytes) continue;
                const uint4 v = *reinterpret_cast<const uint4 *>(ot + row * (kTileW * 3) + cb);
                uint8_t *p = A.F.canvas + (long long)(ty0 + row) * A.F.cpitch + (long long)tx0 * 3 + cb;
                if (cb + 16 <= wbytes) {
                    *reinterpret_cast<uint4 *>(p) = v;
                } else {
                    const uint8_t *sv = ot + row * (kTileW * 3) + cb;
                    for (int j = 0; j < wbytes - cb; ++j) p[j] = sv[j];
                }
            }
        }
        out_sel ^= 1;
    }
}

}  // namespace bs
