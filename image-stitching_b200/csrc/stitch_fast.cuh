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
    unsigned *sched;                       // [0] next tile to hand out, [1] CTAs that ran dry (self-resetting)
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
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(20000u)   // suspend-time hint (ns): sleep in hardware, woken by the completing arrive
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

// The twelve tap bytes of one interior pixel as they come out of shared memory, plus its weights.
struct Taps {
    uint32_t w0, w1, w2;   // three aligned words of source row sy starting at the word that holds texel sx
    uint32_t v0, v1, v2;   // same for row sy + 1
    uint2 wt;              // weight-table entry
    uint32_t sh;           // 8 * (byte offset of texel sx in w0); the funnel shifter uses the low 5 bits
};

// `o` = shared byte address of texel (sx, sy) (any alignment), `pitch` = stage row pitch in bytes,
// `lut_off` = shared address of the weight-table entry.
__device__ __forceinline__ Taps fetch_taps(uint32_t o, int pitch, uint32_t lut_off)
{
    Taps t;
    const uint32_t wa = o & ~3u;
    t.sh = o << 3;
    t.w0 = lds32(wa);
    t.w1 = lds32(wa + 4);
    t.w2 = lds32(wa + 8);
    t.v0 = lds32(wa + pitch);
    t.v1 = lds32(wa + pitch + 4);
    t.v2 = lds32(wa + pitch + 8);
    t.wt = lds64(lut_off);
    return t;
}

// Fixed-point bilinear blend of one pixel, bit-identical to (sum w*S + 2^14) >> 15 per channel.
// Returns B | G<<8 | R<<16.
__device__ __forceinline__ uint32_t blend_taps(const Taps &t)
{
    // six aligned bytes of each row: r?a = [B0 G0 R0 B1], r?b = [G1 R1 . .]
    const uint32_t r0a = __funnelshift_r(t.w0, t.w1, t.sh), r0b = __funnelshift_r(t.w1, t.w2, t.sh);
    const uint32_t r1a = __funnelshift_r(t.v0, t.v1, t.sh), r1b = __funnelshift_r(t.v1, t.v2, t.sh);
    // transpose to one word per channel, taps in order 00 01 10 11
    const uint32_t qb = __byte_perm(r0a, r1a, 0x7430);
    const uint32_t t0 = __byte_perm(r0a, r0b, 0x5241);   // [G00 G01 R00 R01]
    const uint32_t t1 = __byte_perm(r1a, r1b, 0x5241);   // [G10 G11 R10 R11]
    const uint32_t qg = __byte_perm(t0, t1, 0x5410);
    const uint32_t qr = __byte_perm(t0, t1, 0x7632);
    const uint32_t vb = __dp4a(qb, t.wt.x, 128u) * 256u + __dp4a(qb, t.wt.y, 0u);
    const uint32_t vg = __dp4a(qg, t.wt.x, 128u) * 256u + __dp4a(qg, t.wt.y, 0u);
    const uint32_t vr = __dp4a(qr, t.wt.x, 128u) * 256u + __dp4a(qr, t.wt.y, 0u);
    // results sit in byte 2 of vb, vg, vr; byte 3 is always 0
    const uint32_t bg = __byte_perm(vb, vg, 0x3362);
    return __byte_perm(bg, vr, 0x7610);
}

__device__ __forceinline__ uint32_t sample_staged(uint32_t o, int pitch, uint32_t lut_off)
{
    return blend_taps(fetch_taps(o, pitch, lut_off));
}

// Store one row's pixel of every lane into the warp's output block: lanes 4q..4q+3 hold 4 adjacent
// pixels = 12 bytes = 3 words; lane 4q+i (i < 3) builds word i from its own pixel and its right
// neighbour's and stores it.  `row_addr` = shared address of the block row (96 bytes).
__device__ __forceinline__ void store_row_pixel(uint32_t px, uint32_t row_addr, int lane, uint32_t sel)
{
    const uint32_t nxt = __shfl_down_sync(0xffffffffu, px, 1);
    const uint32_t word = __byte_perm(px, nxt, sel);
    if ((lane & 3) != 3)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(row_addr + 4u * (uint32_t)(3 * (lane >> 2) + (lane & 3))), "r"(word)
                     : "memory");
}

// ---- work decomposition ------------------------------------------------------------------------
// A CTA is 8 consumer warps + 1 producer warp.  A 128x32 tile is 8 blocks of 32 columns x 16 rows;
// consumer warp w owns block (w & 3, w >> 2) of every tile its CTA processes: lane = column, one
// vertical run of kRun pixels per lane.  The warps never meet at a CTA barrier: the producer feeds
// source boxes through a ring of kStages stages guarded by full/empty mbarriers, and every warp
// composites into, and flushes, its own 16 x 96-byte output block.
constexpr int kConsumerWarps = kBlocksPerTile;
constexpr int kFastThreads = 32 * (kConsumerWarps + 1);
#ifndef BS_MIN_CTAS
#define BS_MIN_CTAS 2
#endif
#ifndef BS_STAGES
#define BS_STAGES 3
#endif
constexpr int kStages = BS_STAGES;
constexpr int kBlockRowBytes = 96;                     // 32 pixels
constexpr int kWarpOutBytes = kRun * kBlockRowBytes;   // 1536

// Hot path: the layer covers every pixel of the block with interior samples, the tile lies inside
// the canvas and the plan verified the cheap coordinates for the whole block.  No coverage tests,
// no fp64 beyond the run's anchor; fully unrolled so k is a constant.
__device__ __forceinline__ void run_layer_full(const CamDev &cam, int pitch, uint32_t stage_origin, uint32_t lut_base,
                                               uint32_t out_blk, int x, int yc, int lane)
{
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const RunAnchor a = run_anchor(cam.M, x, yc);
    const uint32_t run_base = stage_origin + (uint32_t)((a.by >> kFracBits) * pitch + (a.bx >> kFracBits) * 3);
    const int step = a.ny * pitch + a.nx * 3;      // whole texels per canvas row, as a byte offset
    // software pipeline: the shared-memory loads of pixel k+1 are issued before the arithmetic of
    // pixel k, so a warp has work to do while its loads are in flight
    auto fetch = [&](int k) {
        const int j = k - kRun / 2;
        int xbits, ybits;
        run_coord_bits(a, (float)j, xbits, ybits);
        const uint32_t o = run_base + (uint32_t)(j * step) +
                           (uint32_t)((ybits >> kFracBits) * pitch + (xbits >> kFracBits) * 3);
        // the table sits on an 8 KB boundary, so its base can be OR-ed in after masking the index
        const uint32_t lut_off = lut_base | ((uint32_t)((ybits << 8) + ((xbits << 3) & 0xF8)) & 0x1FF8u);
        return fetch_taps(o, pitch, lut_off);
    };
    Taps cur = fetch(0);
#pragma unroll
    for (int k = 0; k < kRun; ++k) {
        Taps nxt = cur;
        if (k + 1 < kRun) nxt = fetch(k + 1);
        const uint32_t px = blend_taps(cur);
        store_row_pixel(px, out_blk + (uint32_t)(k * kBlockRowBytes), lane, sel);
        cur = nxt;
    }
}

// General path: partially covered layers, tiles that stick out of the canvas, and blocks for which
// the plan found at least one pixel where the cheap coordinates differ (those 32x4 groups use
// the exact fp64 form).  Same arithmetic per pixel, with coverage tests and byte stores.
__device__ __noinline__ void run_layer_general(const FastArgs &A, const CamDev &cam, int pitch, uint32_t stage_origin,
                                               uint32_t lut_base, uint32_t out_blk, uint32_t group_mask, bool full,
                                               bool first_layer, int x, int y0, int lane)
{
    const RunAnchor a = run_anchor(cam.M, x, y0 + kRun / 2);
    const int xb = (x / A.F.block_w) * A.F.block_w;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        int X, Y;
        if ((group_mask >> (k >> 2)) & 1) {     // warp-uniform
            coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        } else {
            const int j = k - kRun / 2;
            int xbits, ybits;
            run_coord_bits(a, (float)j, xbits, ybits);
            X = xbits + a.bx + 32 * a.nx * j;
            Y = ybits + a.by + 32 * a.ny * j;
        }
        const int sx = X >> kFracBits, sy = Y >> kFracBits;
        bool in = (x < A.F.W) && (y0 + k < A.F.H);
        if (!full) in = in && sx >= 0 && sy >= 0 && sx < cam.w - 1 && sy < cam.h - 1;
        const uint32_t o = in ? stage_origin + (uint32_t)(sy * pitch + sx * 3) : out_blk;   // any valid address
        const uint32_t lut_off = lut_base + (uint32_t)((((Y & 31) << 5) | (X & 31)) << 3);
        uint32_t px = sample_staged(o, pitch, lut_off);
        // the first layer also defines the background: pixels it does not cover start at 0
        if (in || first_layer) {
            if (!in) px = 0;
            const uint32_t p8 = out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane;
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8), "r"(px) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 1), "r"(px >> 8) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 2), "r"(px >> 16) : "memory");
        }
    }
}

// What the producer publishes with every ring slot: everything a consumer needs to know about the
// item, so consumers never wait on plan records in global memory.
//   word 0  tx | ty << 16
//   word 1  type | last_layer << 8 | all_mask << 16 | skip_mask << 24
//   word 2  camera
//   word 3  by0 * pitch + 4 * box_elem0(bx0): stage offset of the box origin texel
//   word 4  exact_mask
//   word 5  stage row pitch in bytes
constexpr int kMetaBytes = 32;

__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(bar),
        "r"(parity), "r"(20000u)   // suspend-time hint (ns): sleep in hardware, woken by the completing arrive
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// Rows x 96 bytes of one block, with bounds checks: zero fill (src == 0) or copy from shared.
__device__ __noinline__ void store_block_checked(const FrameArgs &F, uint32_t src, int tx0, int ty0, int lane)
{
    const int wbytes = min(32, F.W - tx0) * 3;
    for (int c = lane; c < kRun * 6; c += 32) {
        const int row = c / 6, cb = (c % 6) * 16;
        if (ty0 + row >= F.H || cb >= wbytes) continue;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (src) v = lds128(src + (uint32_t)(row * kBlockRowBytes + cb));
        uint8_t *p = F.canvas + (long long)(ty0 + row) * F.cpitch + (long long)tx0 * 3 + cb;
        if (cb + 16 <= wbytes) {
            *reinterpret_cast<uint4 *>(p) = v;
        } else {
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            for (int j = 0; j < wbytes - cb; ++j) p[j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
        }
    }
}

__global__ void __launch_bounds__(kFastThreads, BS_MIN_CTAS) stitch_fast_kernel(const __grid_constant__ FastArgs A)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    // layout: [pad to an 8 KB boundary][weight table 8 KB][kStages stages][8 output blocks][meta][barriers]
    const uint32_t raw_base = smem_u32(smem_raw);
    const uint32_t lut_base = (raw_base + 8191u) & ~8191u;
    const uint32_t stage0 = lut_base + kLutBytes;                                // 8 KB aligned
    const uint32_t out0 = stage0 + (uint32_t)(kStages * A.stage_bytes);          // 8 private output blocks
    const uint32_t meta0 = out0 + kConsumerWarps * kWarpOutBytes;
    const uint32_t full0 = meta0 + kStages * kMetaBytes;
    const uint32_t empty0 = full0 + kStages * 8;
    uint8_t *const gen0 = smem_raw + (lut_base - raw_base);                      // generic pointer to the table

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    // Programmatic dependent launch: let the next frame's kernel start filling SMs as this one's CTAs
    // retire.  It may prefetch and composite its first tile, but touches the canvas and the tile
    // scheduler only after griddepcontrol.wait, i.e. after this grid has completed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    {
        uint2 *lut = reinterpret_cast<uint2 *>(gen0);
        for (int i = tid; i < 1024; i += kFastThreads) lut[i] = lut_entry(i);
    }
    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full0 + 8 * s), "r"(1));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty0 + 8 * s), "r"(kConsumerWarps));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == kConsumerWarps) {
        // ===== producer: one lane walks the tiles of this CTA and keeps the ring full ===============
        // Every tile becomes one ring item per layer (FAST) or a single descriptor-only item.
        if (lane == 0) {
            uint32_t slot = 0, phase = 1;    // a fresh mbarrier passes a wait on parity 1
            // The first tile of a CTA is static (its index is the CTA's), the rest are handed out
            // dynamically, heaviest first in the plan's order, so there is no tail imbalance.  The
            // counter may still belong to the previous launch until that grid has completed.
            bool prior_grid_done = false;
            for (int k = (int)blockIdx.x; k < A.n_tiles;) {
                const uint4 *rp = reinterpret_cast<const uint4 *>(&A.tiles[k]);
                TileRec r;
                {
                    uint4 *dst = reinterpret_cast<uint4 *>(&r);
#pragma unroll
                    for (int q = 0; q < (int)(sizeof(TileRec) / 16); ++q) dst[q] = __ldg(rp + q);
                }
                const bool fast = (r.type == kTileFast) && !(A.F.keep_canvas && r.has_uncovered);
                const int n_items = fast ? r.n_layers : 1;
                uint32_t higher_all = 0;     // blocks fully written by a later (higher) layer
                uint32_t skip[kMaxLayers];
                if (fast)
                    for (int l = n_items - 1; l >= 0; --l) {
                        skip[l] = (uint32_t)r.none_mask[l] | higher_all;
                        higher_all |= r.all_mask[l];
                    }
                for (int l = 0; l < n_items; ++l) {
                    mbar_wait_addr(empty0 + 8 * slot, phase);   // all 8 warps released the slot
                    uint32_t kind, cam = 0, origin_off = 0, exact_mask = 0, pitch = 0;
                    if (fast) {
                        cam = r.cam[l];
                        pitch = (uint32_t)(A.box_we[cam] * 4);
                        kind = (uint32_t)kTileFast | ((l == n_items - 1) ? 1u << 8 : 0u) | ((uint32_t)r.all_mask[l] << 16) |
                               (skip[l] << 24);
                        origin_off = (uint32_t)((int)r.by0[l] * (int)pitch + 4 * box_elem0((int)r.bx0[l]));
                        exact_mask = r.exact_mask[l];
                    } else {
                        kind = (uint32_t)(r.type == kTileFast ? kTileExact : r.type) | (1u << 8);
                    }
                    const uint32_t ma = meta0 + kMetaBytes * slot;
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(ma),
                                 "r"((uint32_t)r.tx | ((uint32_t)r.ty << 16)), "r"(kind), "r"(cam), "r"(origin_off)
                                 : "memory");
                    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(ma + 16), "r"(exact_mask), "r"(pitch) : "memory");
                    if (fast) {
                        // arrive has release semantics: the meta words are visible to whoever sees the phase flip
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full0 + 8 * slot),
                                     "r"((uint32_t)(A.box_we[cam] * 4 * A.box_h[cam]))
                                     : "memory");
                        asm volatile(
                            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                                stage0 + slot * (uint32_t)A.stage_bytes),
                            "l"(&A.tmap[cam]), "r"(box_elem0((int)r.bx0[l])), "r"((int)r.by0[l]), "r"(full0 + 8 * slot)
                            : "memory");
                    } else {
                        mbar_arrive_addr(full0 + 8 * slot);
                    }
                    if (++slot == kStages) {
                        slot = 0;
                        phase ^= 1u;
                    }
                }
                if (!prior_grid_done) {
                    asm volatile("griddepcontrol.wait;" ::: "memory");
                    prior_grid_done = true;
                }
                k = (int)atomicAdd(&A.sched[0], 1u);
            }
            if (!prior_grid_done) asm volatile("griddepcontrol.wait;" ::: "memory");
            // out of tiles: tell the consumers to stop, and let the last CTA re-arm the scheduler
            mbar_wait_addr(empty0 + 8 * slot, phase);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(meta0 + kMetaBytes * slot), "r"(0u), "r"(0xFFu),
                         "r"(0u), "r"(0u)
                         : "memory");
            mbar_arrive_addr(full0 + 8 * slot);
            if (atomicInc(&A.sched[1], gridDim.x - 1) == gridDim.x - 1) atomicExch(&A.sched[0], gridDim.x);
        }
        return;
    }

    // ===== consumers ====================================================================================
    const uint32_t out_blk = out0 + (uint32_t)(warp * kWarpOutBytes);
    const int bx = (warp & 3) * 32, by = (warp >> 2) * kRun;      // block origin inside a tile
    uint32_t slot = 0, phase = 0;
    bool prior_grid_done = false;     // griddepcontrol.wait executed (needed before any canvas access)
    for (;;) {
        bool wrote = false, last;
        do {
            mbar_wait_addr(full0 + 8 * slot, phase);
            const uint4 m0 = lds128(meta0 + kMetaBytes * slot);
            const int tx0 = (int)(m0.x & 0xFFFFu) * kTileW + bx, ty0 = (int)(m0.x >> 16) * kTileH + by;
            const int type = (int)(m0.y & 0xFFu);
            if (type == 0xFF) return;          // the producer ran out of tiles
            last = (m0.y >> 8) & 1u;
            if (type == kTileFast) {
                if (!((m0.y >> (24 + warp)) & 1u)) {          // this layer writes into this block
                    const uint2 m1 = lds64(meta0 + kMetaBytes * slot + 16);
                    const CamDev &cam = A.F.cam[m0.z];
                    const int pitch = (int)m1.y;
                    const bool full = (m0.y >> (16 + warp)) & 1u;
                    const uint32_t group_mask = (m1.x >> (warp * 4)) & 15u;   // this block's 4 row groups
                    // shared address texel (0, 0) of this camera would have, were the whole frame staged
                    const uint32_t stage_origin = stage0 + slot * (uint32_t)A.stage_bytes - m0.w;
                    if (full && group_mask == 0u)
                        run_layer_full(cam, pitch, stage_origin, lut_base, out_blk, tx0 + lane, ty0 + kRun / 2, lane);
                    else
                        run_layer_general(A, cam, pitch, stage_origin, lut_base, out_blk, group_mask, full, !wrote,
                                          tx0 + lane, ty0, lane);
                    wrote = true;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(empty0 + 8 * slot);    // this warp is done with the slot
                if (last) {
                    if (!prior_grid_done) {
                        asm volatile("griddepcontrol.wait;" ::: "memory");
                        prior_grid_done = true;
                    }
                    if (!wrote) {
                        if (!A.F.keep_canvas) store_block_checked(A.F, 0u, tx0, ty0, lane);   // nobody wrote here
                    } else if (tx0 + 32 <= A.F.W && ty0 + kRun <= A.F.H) {
                        // flush: lane pair (2r, 2r+1) writes one full 32-byte sector of row r per instruction
                        const uint32_t so = out_blk + (uint32_t)((lane >> 1) * kBlockRowBytes + (lane & 1) * 16);
                        uint8_t *const gp = A.F.canvas + (long long)(ty0 + (lane >> 1)) * A.F.cpitch +
                                            (long long)tx0 * 3 + (lane & 1) * 16;
#pragma unroll
                        for (int j = 0; j < 3; ++j) *reinterpret_cast<uint4 *>(gp + 32 * j) = lds128(so + 32u * j);
                    } else {
                        store_block_checked(A.F, out_blk, tx0, ty0, lane);
                    }
                    __syncwarp();   // the block is free again
                }
            } else {
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(empty0 + 8 * slot);
                if (!prior_grid_done) {
                    asm volatile("griddepcontrol.wait;" ::: "memory");
                    prior_grid_done = true;
                }
                if (type == kTileEmpty) {
                    if (!A.F.keep_canvas) store_block_checked(A.F, 0u, tx0, ty0, lane);
                } else {
                    // exact block: lane takes the 4-pixel group (lane & 7) of rows (lane >> 3) + 4 j
                    for (int j = 0; j < kRun / 4; ++j) exact_quad(A.F, tx0 + (lane & 7) * 4, ty0 + (lane >> 3) + 4 * j);
                }
            }
            if (++slot == kStages) {
                slot = 0;
                phase ^= 1u;
            }
        } while (!last);
    }
}

}  // namespace bs
