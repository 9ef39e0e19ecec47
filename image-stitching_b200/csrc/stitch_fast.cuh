// stitch_fast.cuh -- the per-frame kernel: every warp is an independent worker on 32x16 blocks.
//
// Each warp owns a contiguous range of the plan's work items (one item = one canvas block and one
// camera layer; a block's layers are adjacent, lowest camera first, so later cameras overwrite
// earlier ones exactly as the reference's loop does, /root/reference/tools/stitch/main.cpp:145-152).
// For every item the warp
//   1. prefetches, one item ahead, the camera's source bounding box into one of its two shared-memory
//      stages with one TMA tensor copy (cp.async.bulk.tensor.2d) signalled on its own mbarrier --
//      lane 0 is the producer, nothing is shared with other warps, so no warp ever waits for one;
//   2. computes one vertical run of 16 canvas pixels per lane: coordinates from the plan-verified
//      fp32 expansion (flagged rows are redone with exact fp64 coordinates), 2x3 aligned words per
//      pixel from the staged box, OpenCV's fixed-point bilinear weights computed from the fractions,
//      three channel dot products (dp2a) -- bit-identical to (sum w*S + 2^14) >> 15;
//   3. stores each finished row straight to the canvas when the block has a single layer; blocks
//      with two layers, partially covered layers and a warp's first block are composited in a
//      16x96-byte block in shared memory and written out as 16-byte vector stores.
// EXACT blocks run the per-pixel fp64 path of stitch_exact.cuh (four rows per item); EMPTY blocks are
// zero-filled.
#pragma once

#include <cuda.h>   // CUtensorMap (type only; the encoder is fetched at run time)

#include "stitch_exact.cuh"
#include "stitch_plan.cuh"

namespace bs {

constexpr int kStagePad = 128;                       // slack behind a stage for the 3rd word of the last texel

struct FastArgs {
    FrameArgs F;
    CUtensorMap tmap[kMaxCamsPerLaunch];   // source frame of camera i as a [h][pitch/4] u32 tensor
    const WorkItem *items;                 // the plan, flattened and grouped by warp: warp w owns items
    const int *warp_start;                 //   [warp_start[w], warp_start[w+1]); a block's layers are adjacent
    int pitch[kMaxCamsPerLaunch];          // TMA box of camera i: row pitch in shared memory (bytes) ...
    int box_bytes[kMaxCamsPerLaunch];      // ... and total bytes (= what the mbarrier expects)
    int stage_bytes;                       // shared memory per stage (multiple of 128)
    int rec_slots;                         // record ring slots per warp: 16, or 8 when that lets one more CTA fit an SM
    int wait_sources;                      // 1: the source frames may come from the kernel just before this one in the
                                           // stream: order every source read behind it (griddepcontrol.wait first)
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

// Bilinear weights for fractional position (a, b), a = X & 31, b = Y & 31, as dp2a wants them.
// w' = w / 32 are OpenCV's 15-bit weights without their common factor 32; they sum to 1024.
// Result = { 64 w'00 | 64 w'01 << 16, 64 w'10 | 64 w'11 << 16 }, so that
//   sum(64 w' S) + 2^15 = 2 (sum(w S) + 2^14)  has the pixel in byte 2.
// Computed, not looked up: a 1024-entry table in shared memory cost 4-5 wavefronts per warp and
// pixel row (bank conflicts) on a shared-memory pipe that was 80 % busy.  (tests/test_fixed_point_identity.py
// mirrors this function, wrap-around included.)
// (a, b) = (0, 0) would need 64 w'00 = 65536; 65535 gives the same pixel for every S in 0..255:
// (65535 S + 2^15) >> 16 == S  because  0 < 2^15 - S < 2^16.
__device__ __forceinline__ uint2 tap_weights(uint32_t a, uint32_t b)
{
    const uint32_t a2 = a * (0xFFFFu * 64u) + 2048u;   // 64 (32 - a) | 64 a << 16; no product below carries between halves ...
    const uint32_t w23 = b * a2;
    uint32_t w01 = a2 * 32u - w23;
    if (w01 == 0x10000u) w01 = 0xFFFFu;                // ... except 64 * 32 * 32 at (0, 0)
    return make_uint2(w01, w23);
}

// The twelve tap bytes of one interior pixel as they come out of shared memory, plus its weights.
struct Taps {
    uint32_t w0, w1, w2;   // three aligned words of source row sy starting at the word that holds texel sx
    uint32_t v0, v1, v2;   // same for row sy + 1
    uint2 wt;              // tap_weights()
    uint32_t sh;           // 8 * (byte offset of texel sx in w0); the funnel shifter uses the low 5 bits
};

// `o` = shared byte address of texel (sx, sy) (any alignment), `pitch` = stage row pitch in bytes.
__device__ __forceinline__ Taps fetch_taps(uint32_t o, int pitch, uint2 wt)
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
    t.wt = wt;
    return t;
}

// Fixed-point bilinear blend of one pixel, bit-identical to (sum w*S + 2^14) >> 15 per channel.
// Returns B | G<<8 | R<<16.
__device__ __forceinline__ uint32_t blend_taps(const Taps &t)
{
    // six aligned bytes of each row: r?a = [B0 G0 R0 B1], r?b = [G1 R1 . .]
    const uint32_t r0a = __funnelshift_r(t.w0, t.w1, t.sh), r0b = __funnelshift_r(t.w1, t.w2, t.sh);
    const uint32_t r1a = __funnelshift_r(t.v0, t.v1, t.sh), r1b = __funnelshift_r(t.v1, t.v2, t.sh);
    // pair the two taps of a row per channel, as dp2a wants them
    const uint32_t p0 = __byte_perm(r0a, r0b, 0x4130), q0 = __byte_perm(r0a, r0b, 0x5252);   // [B00 B01 G00 G01], [R00 R01 ..]
    const uint32_t p1 = __byte_perm(r1a, r1b, 0x4130), q1 = __byte_perm(r1a, r1b, 0x5252);   // same for row 1
    const uint32_t vb = __dp2a_lo(t.wt.y, p1, __dp2a_lo(t.wt.x, p0, 32768u));
    const uint32_t vg = __dp2a_hi(t.wt.y, p1, __dp2a_hi(t.wt.x, p0, 32768u));
    const uint32_t vr = __dp2a_lo(t.wt.y, q1, __dp2a_lo(t.wt.x, q0, 32768u));
    // results sit in byte 2 of vb, vg, vr; byte 3 is always 0
    const uint32_t bg = __byte_perm(vb, vg, 0x3362);
    return __byte_perm(bg, vr, 0x7610);
}

__device__ __forceinline__ uint32_t sample_staged(uint32_t o, int pitch, uint2 wt)
{
    return blend_taps(fetch_taps(o, pitch, wt));
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
// Same, straight to the canvas: (`hi`:`lo`) = address of this lane's word of the row, `storing` != 0
// for the 24 lanes that own a word (lanes 4q+3 do not).  They cover the row's 96 bytes = three full
// 32-byte sectors.
__device__ __forceinline__ void store_row_pixel_global(uint32_t px, uint32_t lo, uint32_t hi, uint32_t storing, uint32_t sel)
{
    const uint32_t nxt = __shfl_down_sync(0xffffffffu, px, 1);
    const uint32_t word = __byte_perm(px, nxt, sel);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 a;\n"
        "setp.ne.u32 p, %3, 0;\n"
        "mov.b64 a, {%0, %1};\n"
        "@p st.global.u32 [a], %2;\n"
        "}" ::"r"(lo),
        "r"(hi), "r"(word), "r"(storing)
        : "memory");
}

// ---- work decomposition ------------------------------------------------------------------------
#ifndef BS_WARPS_PER_CTA
#define BS_WARPS_PER_CTA 8
#endif
constexpr int kWarpsPerCta = BS_WARPS_PER_CTA;
constexpr int kFastThreads = 32 * kWarpsPerCta;
#ifndef BS_MIN_CTAS
#define BS_MIN_CTAS 2
#endif
#ifndef BS_STAGES
#define BS_STAGES 2
#endif
constexpr int kStages = BS_STAGES;                     // per warp: source boxes of items i .. i + kStages - 1
constexpr int kPrefetch = kStages - 1;                 // boxes in flight ahead of the one being computed
constexpr int kRecSlotsMax = 16;                       // per warp: record ring of 16 or 8 slots (FastArgs::rec_slots),
                                                       // refilled in halves by one bulk copy each
constexpr int kBlockRowBytes = 96;                     // 32 pixels
constexpr int kWarpOutBytes = kRun * kBlockRowBytes;   // 1536
constexpr int kRecBytes = 80;                          // sizeof(WorkItem)

// records + mbarriers, rounded so that every warp's region stays 128-byte aligned
__host__ __device__ constexpr int warp_ctl_bytes(int rec_slots)
{
    return (rec_slots * kRecBytes + (kStages + 2) * 8 + 127) / 128 * 128;
}

__host__ __device__ constexpr int warp_smem_bytes(int stage_bytes, int rec_slots)
{
    return kStages * stage_bytes + kWarpOutBytes + warp_ctl_bytes(rec_slots);
}

// The block's coordinate expansion as it sits in a record slot (words 4..19 of the work item).
__device__ __forceinline__ BlockCoef load_block_coef(uint32_t rec_addr)
{
    uint4 q[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(q[i].x), "=r"(q[i].y), "=r"(q[i].z), "=r"(q[i].w)
                     : "r"(rec_addr + 16u + 16u * i));
    BlockCoef c;
    memcpy(&c, q, sizeof c);
    return c;
}

// Hot path: the layer covers every pixel of the block with interior samples, the block lies inside
// the canvas and the plan verified the cheap coordinates for the whole block.  No coverage tests,
// no fp64; fully unrolled so the row offset v is a constant.
// kDirect: the block has no other layer and the previous frame's kernel is known to have finished,
// so every row goes straight to the canvas (`grow` = this lane's word of the block's first row);
// otherwise rows are collected in the warp's output block in shared memory.
template <bool kDirect>
__device__ __forceinline__ void run_layer_full(const LaneCoef &lc, int pitch, uint32_t stage_origin,
                                               uint32_t out_blk, int lane, uint8_t *grow, long long cpitch,
                                               uint32_t sel, uint32_t storing)
{
    const uint32_t glo = (uint32_t)reinterpret_cast<uintptr_t>(grow), ghi = (uint32_t)(reinterpret_cast<uintptr_t>(grow) >> 32);
    const uint32_t run_base = stage_origin + (uint32_t)((lc.by >> kFracBits) * pitch + (lc.bx >> kFracBits) * 3);
    const int step = lc.nyv * pitch + lc.nxv * 3;      // whole texels per canvas row, as a byte offset
    auto fetch = [&](int k) {
        const int v = k - kRun / 2;
        int xbits, ybits;
        lane_coord_bits(lc, (float)v, xbits, ybits);
        const uint32_t o = run_base + (uint32_t)(v * step) +
                           (uint32_t)((ybits >> kFracBits) * pitch + (xbits >> kFracBits) * 3);
        return fetch_taps(o, pitch, tap_weights((uint32_t)xbits & 31u, (uint32_t)ybits & 31u));
    };
    // software pipeline: the shared-memory loads of pixel k+1 are issued before the arithmetic of
    // pixel k, so a warp has work to do while its loads are in flight
    Taps cur = fetch(0);
#pragma unroll
    for (int k = 0; k < kRun; ++k) {
        Taps nxt = cur;
        if (k + 1 < kRun) nxt = fetch(k + 1);
        const uint32_t px = blend_taps(cur);
        if (kDirect) {
            // the canvas does not straddle a 4 GB boundary (the host checks), so only the low half
            // of the row address advances -- one add per row instead of a 64-bit one
            store_row_pixel_global(px, glo + (uint32_t)k * (uint32_t)cpitch, ghi, storing, sel);
        } else
            store_row_pixel(px, out_blk + (uint32_t)(k * kBlockRowBytes), lane, sel);
        cur = nxt;
    }
}

// Partially covered layer (a seam, the canvas border) whose cheap coordinates hold everywhere: the
// hot path's arithmetic and software pipeline, but every pixel tests whether the camera covers it and
// stores its three bytes on its own.  `first_layer`: pixels this layer does not cover start at 0.
__device__ __noinline__ void run_layer_masked(const LaneCoef &lc, int pitch, uint32_t stage_origin, uint32_t out_blk, int lane,
                                              int src_w, int src_h, bool col_in_canvas, int rows_in_canvas, bool first_layer)
{
    const uint32_t run_base = stage_origin + (uint32_t)((lc.by >> kFracBits) * pitch + (lc.bx >> kFracBits) * 3);
    const int step = lc.nyv * pitch + lc.nxv * 3;
    const uint32_t safe = out_blk;                      // any valid shared address for pixels that are not covered
    auto fetch = [&](int k, bool &in) {
        const int v = k - kRun / 2;
        int xbits, ybits;
        lane_coord_bits(lc, (float)v, xbits, ybits);
        const int sx = (xbits + lc.bx + 32 * lc.nxv * v) >> kFracBits, sy = (ybits + lc.by + 32 * lc.nyv * v) >> kFracBits;
        in = col_in_canvas && k < rows_in_canvas && (uint32_t)sx < (uint32_t)(src_w - 1) && (uint32_t)sy < (uint32_t)(src_h - 1);
        const uint32_t o = run_base + (uint32_t)(v * step) +
                           (uint32_t)((ybits >> kFracBits) * pitch + (xbits >> kFracBits) * 3);
        return fetch_taps(in ? o : safe, pitch, tap_weights((uint32_t)xbits & 31u, (uint32_t)ybits & 31u));
    };
    bool in_cur, in_nxt = false;
    Taps cur = fetch(0, in_cur);
#pragma unroll 4
    for (int k = 0; k < kRun; ++k) {
        Taps nxt = cur;
        if (k + 1 < kRun) nxt = fetch(k + 1, in_nxt);
        uint32_t px = blend_taps(cur);
        if (in_cur || first_layer) {
            if (!in_cur) px = 0;
            const uint32_t p8 = out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane;
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8), "r"(px) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 1), "r"(px >> 8) : "memory");
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 2), "r"(px >> 16) : "memory");
        }
        cur = nxt;
        in_cur = in_nxt;
    }
}

// General path: partially covered layers, blocks that stick out of the canvas, and blocks for which
// the plan found at least one pixel where the cheap coordinates differ (those 32x4 groups use
// the exact fp64 form).  Same arithmetic per pixel, with coverage tests and byte stores.
__device__ __noinline__ void run_layer_general(const FastArgs &A, const CamDev &cam, const LaneCoef &lc, int pitch,
                                               uint32_t stage_origin, uint32_t out_blk,
                                               uint32_t group_mask, bool full, bool first_layer, int x, int y0, int lane)
{
    const int xb = (x / A.F.block_w) * A.F.block_w;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        int X, Y;
        if ((group_mask >> (k >> 2)) & 1) {     // warp-uniform
            coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        } else {
            const int v = k - kRun / 2;
            int xbits, ybits;
            lane_coord_bits(lc, (float)v, xbits, ybits);
            X = xbits + lc.bx + 32 * lc.nxv * v;
            Y = ybits + lc.by + 32 * lc.nyv * v;
        }
        const int sx = X >> kFracBits, sy = Y >> kFracBits;
        bool in = (x < A.F.W) && (y0 + k < A.F.H);
        if (!full) in = in && sx >= 0 && sy >= 0 && sx < cam.w - 1 && sy < cam.h - 1;
        const uint32_t o = in ? stage_origin + (uint32_t)(sy * pitch + sx * 3) : out_blk;   // any valid address
        uint32_t px = sample_staged(o, pitch, tap_weights((uint32_t)X & 31u, (uint32_t)Y & 31u));
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

// Fully covered layer with flagged groups: the hot path has filled the output block with the cheap
// coordinates; redo the flagged 32x4 groups with the exact fp64 coordinates (same taps, same blend).
__device__ __noinline__ void patch_exact_groups(const FastArgs &A, const CamDev &cam, int pitch, uint32_t stage_origin,
                                                uint32_t out_blk, uint32_t group_mask, int x, int y0, int lane)
{
    const int xb = (x / A.F.block_w) * A.F.block_w;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        if (!((group_mask >> (k >> 2)) & 1)) continue;     // warp-uniform
        int X, Y;
        coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        const uint32_t o = stage_origin + (uint32_t)((Y >> kFracBits) * pitch + (X >> kFracBits) * 3);
        const uint32_t px = sample_staged(o, pitch, tap_weights((uint32_t)X & 31u, (uint32_t)Y & 31u));
        const uint32_t p8 = out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane;
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8), "r"(px) : "memory");
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 1), "r"(px >> 8) : "memory");
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 2), "r"(px >> 16) : "memory");
    }
}

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
// One probe of the barrier; lets independent work be issued before the result is needed.
__device__ __forceinline__ bool mbar_try_wait_addr(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
        "selp.u32 %0, 1, 0, P1;\n"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
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
__device__ __noinline__ void store_block_checked(uint8_t *canvas, long long cpitch, int W, int H, uint32_t src, int tx0, int ty0, int lane)
{
    const int wbytes = min(32, W - tx0) * 3;
    for (int c = lane; c < kRun * 6; c += 32) {
        const int row = c / 6, cb = (c % 6) * 16;
        if (ty0 + row >= H || cb >= wbytes) continue;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (src) v = lds128(src + (uint32_t)(row * kBlockRowBytes + cb));
        uint8_t *p = canvas + (long long)(ty0 + row) * cpitch + (long long)tx0 * 3 + cb;
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
    // layout: [pad to a 128-byte boundary][per warp: kStages stages | output block | records | mbarriers]
    const uint32_t raw_base = (smem_u32(smem_raw) + 127u) & ~127u;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const uint32_t wbase = raw_base + (uint32_t)(warp * warp_smem_bytes(A.stage_bytes, A.rec_slots));
    const uint32_t stage0 = wbase;                                             // 128-byte aligned
    const uint32_t out_blk = stage0 + (uint32_t)(kStages * A.stage_bytes);
    const uint32_t rec0 = out_blk + kWarpOutBytes;
    const uint32_t rec_slots = (uint32_t)A.rec_slots, rec_half = rec_slots >> 1;   // powers of two
    const uint32_t dbar0 = rec0 + rec_slots * kRecBytes;                       // data barriers, one per stage
    const uint32_t rbar0 = dbar0 + kStages * 8;                                // record barriers, one per half of the ring

    // Programmatic dependent launch: let the next frame's kernel start filling SMs as this one's CTAs
    // retire.  It may prefetch and composite, but touches the canvas only after griddepcontrol.wait,
    // i.e. after this grid has completed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (lane == 0) {
        for (int s = 0; s < kStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dbar0 + 8 * s), "r"(1));
        for (int s = 0; s < 2; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rbar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();     // the barriers are private to the warp: no CTA-wide synchronisation anywhere in this kernel

    // Work split: the host deals the plan's blocks round-robin to the warps of the grid (block u to
    // warp u mod W, heaviest blocks first, so they spread evenly and at any moment the grid works on W
    // neighbouring blocks) and stores every warp's items contiguously.
    const int gw = (int)blockIdx.x * kWarpsPerCta + warp;
    const int i0 = A.warp_start[gw];
    const int cnt = A.warp_start[gw + 1] - i0;      // this warp's items, numbered n = 0 .. cnt-1
    if (cnt <= 0) return;

    // lane 0 is this warp's producer ------------------------------------------------------------------
    // records of items [first, first + 8) -> their half of the record ring, one bulk copy
    auto request_records = [&](int first) {
        const uint32_t bytes = (uint32_t)(min((int)rec_half, cnt - first) * kRecBytes);
        const uint32_t half = ((uint32_t)first / rec_half) & 1u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rbar0 + 8 * half), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         rec0 + kRecBytes * rec_half * half),
                     "l"(A.items + i0 + first), "r"(bytes), "r"(rbar0 + 8 * half)
                     : "memory");
    };
    // source box of the item whose record words are `r` -> stage; items without a box just arrive
    auto issue_box = [&](const uint4 &r, uint32_t stage) {
        if ((r.y & 0xFFu) == (uint32_t)kBlockFast) {
            const uint32_t cam = (r.y >> 16) & 0xFFu;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dbar0 + 8 * stage),
                         "r"((uint32_t)A.box_bytes[cam])
                         : "memory");
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                    stage0 + stage * (uint32_t)A.stage_bytes),
                "l"(&A.tmap[cam]), "r"((int)(int16_t)(r.z & 0xFFFFu)), "r"((int)(int16_t)(r.z >> 16)), "r"(dbar0 + 8 * stage)
                : "memory");
        } else {
            mbar_arrive_addr(dbar0 + 8 * stage);
        }
    };
    // record words / lane coefficients of item m from the record ring (all lanes)
    auto rec_addr = [&](int m) { return rec0 + kRecBytes * ((uint32_t)m & (rec_slots - 1)); };
    auto wait_half_of = [&](int m) {      // first item of a half: every lane observes the copy's completion itself
        if (((uint32_t)m & (rec_half - 1)) == 0) mbar_wait_addr(rbar0 + 8 * (((uint32_t)m / rec_half) & 1u), ((uint32_t)m / rec_slots) & 1u);
    };

    if (lane == 0) {
        request_records(0);
        if (cnt > (int)rec_half) request_records((int)rec_half);
    }
    __syncwarp();
    wait_half_of(0);
    // The per-item set-up is software-pipelined across items: while item n is computed, the record of
    // item n+2 has been read and the box of item n+1 is in flight.  (The lane coefficients of item n+1
    // are formed after item n: carrying a second set through the hot loop cost ten registers and ten
    // moves per item for nothing.)
    uint4 rec = lds128(rec_addr(0));
    uint4 recn = cnt > 1 ? lds128(rec_addr(1)) : rec;
    LaneCoef lc = lane_coef(load_block_coef(rec_addr(0)), lane - kBlockW / 2);
    bool prior_grid_done = false;     // griddepcontrol.wait executed (needed before any canvas access)
    if (A.wait_sources) {             // only the plan (written long ago) has been read so far
        asm volatile("griddepcontrol.wait;" ::: "memory");
        prior_grid_done = true;
    }
    if (lane == 0) issue_box(rec, 0u);
    __syncwarp();

    // lane constants of the output word merge (store_row_pixel): byte selector, and whether this lane owns a word
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const uint32_t storing = (uint32_t)((lane & 3) != 3);
    bool wrote = false;               // some layer has written into the current block
    uint32_t stage = 0, dpar = 0;     // stage and data-barrier parity of the current item (kStages == 2)
    static_assert(kStages == 2, "the item pipeline below is written for two stages");
    for (int n = 0; n < cnt; ++n) {
        if (lane == 0) {
            // the half of the record ring that items n-8 .. n-1 used is free: refill it with n+8 .. n+15
            if (((uint32_t)n & (rec_half - 1)) == 0 && n >= (int)rec_half && n + (int)rec_half < cnt) request_records(n + (int)rec_half);
            // prefetch: the box of item n+1 goes into the stage item n-1 is done with
            if (n + 1 < cnt) issue_box(recn, stage ^ 1u);
        }
        __syncwarp();
        const bool landed = mbar_try_wait_addr(dbar0 + 8 * stage, dpar);    // probe now, look at the answer later
        // look-ahead read: the record of the item after next
        uint4 recn2 = recn;
        if (n + 2 < cnt) {
            wait_half_of(n + 2);
            recn2 = lds128(rec_addr(n + 2));
        }
        const uint4 r = rec;
        bool direct = false;
        const int tx0 = (int)(r.x & 0xFFFFu) * kBlockW, ty0 = (int)(r.x >> 16) * kBlockH;
        const uint32_t type = r.y & 0xFFu, flags = (r.y >> 8) & 0xFFu;
        if (type == (uint32_t)kBlockFast) {
            const uint32_t c = (r.y >> 16) & 0xFFu;
            const CamDev &cam = A.F.cam[c];
            const int pitch = A.pitch[c];
            // shared address texel (0, 0) of this camera would have, were the whole frame staged
            const uint32_t stage_origin = stage0 + stage * (uint32_t)A.stage_bytes - r.w;
            const uint32_t groups = (r.y >> 24) & 0xFFu;
            const bool full = (flags & kItemFull) != 0;
            if (!landed) mbar_wait_addr(dbar0 + 8 * stage, dpar);     // the box has landed
            if (full) {
                // single-layer block after the first flush: no staging in shared memory, no flush
                direct = groups == 0u && prior_grid_done && A.F.canvas_in_4g &&
                         (flags & (kItemFirst | kItemLast)) == (kItemFirst | kItemLast);
                uint8_t *const grow = A.F.canvas + (long long)ty0 * A.F.cpitch + (long long)tx0 * 3 + 4 * (3 * (lane >> 2) + (lane & 3));
                if (direct)
                    run_layer_full<true>(lc, pitch, stage_origin, out_blk, lane, grow, A.F.cpitch, sel, storing);
                else
                    run_layer_full<false>(lc, pitch, stage_origin, out_blk, lane, grow, A.F.cpitch, sel, storing);
                if (groups != 0u) {
                    __syncwarp();
                    patch_exact_groups(A, cam, pitch, stage_origin, out_blk, groups, tx0 + lane, ty0, lane);
                }
            } else {
                if (groups == 0u)
                    run_layer_masked(lc, pitch, stage_origin, out_blk, lane, cam.w, cam.h, tx0 + lane < A.F.W, A.F.H - ty0, !wrote);
                else
                    run_layer_general(A, cam, lc, pitch, stage_origin, out_blk, groups, full, !wrote, tx0 + lane, ty0, lane);
            }
            wrote = !direct;
        }
        __syncwarp();                 // every lane is done with this item's stage
        if ((flags & kItemLast) && !direct) {
            if (!prior_grid_done) {
                asm volatile("griddepcontrol.wait;" ::: "memory");
                prior_grid_done = true;
            }
            if (type == (uint32_t)kBlockExact) {
                // lane takes the 4-pixel group (lane & 7) of rows (lane >> 3) + 4 j; the item names its row groups j
                const uint32_t rows = (r.y >> 24) & 0xFFu;
                for (int j = 0; j < kRun / 4; ++j)
                    if ((rows >> j) & 1u) exact_quad(A.F, tx0 + (lane & 7) * 4, ty0 + (lane >> 3) + 4 * j);
            } else if (!wrote && A.F.keep_canvas) {
                // nobody writes here and the canvas keeps its contents
            } else if (tx0 + kBlockW <= A.F.W && ty0 + kRun <= A.F.H) {
                // flush (or zero fill of an empty block): lane pair (2r, 2r+1) writes one full 32-byte
                // sector of row r per instruction
                const uint32_t so = out_blk + (uint32_t)((lane >> 1) * kBlockRowBytes + (lane & 1) * 16);
                uint8_t *const gp = A.F.canvas + (long long)(ty0 + (lane >> 1)) * A.F.cpitch + (long long)tx0 * 3 + (lane & 1) * 16;
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    *reinterpret_cast<uint4 *>(gp + 32 * j) = wrote ? lds128(so + 32u * j) : make_uint4(0u, 0u, 0u, 0u);
            } else if (!wrote) {
                store_block_checked(A.F.canvas, A.F.cpitch, A.F.W, A.F.H, 0u, tx0, ty0, lane);
            } else {
                store_block_checked(A.F.canvas, A.F.cpitch, A.F.W, A.F.H, out_blk, tx0, ty0, lane);
            }
            wrote = false;
            __syncwarp();             // the output block is free again
        }
        rec = recn;
        recn = recn2;
        if (n + 1 < cnt) lc = lane_coef(load_block_coef(rec_addr(n + 1)), lane - kBlockW / 2);
        dpar ^= stage;                // the parity flips every second item
        stage ^= 1u;
    }
}

}  // namespace bs
