// stitch_fast.cuh -- the per-frame kernel: every warp is an independent worker on 32x16 blocks.
//
// Each warp owns a contiguous range of the plan's work items (one item = one canvas block and one
// camera layer; a block's layers are adjacent, lowest camera first, so later cameras overwrite
// earlier ones exactly as the reference's loop does, /root/reference/tools/stitch/main.cpp:145-152).
// For every item the warp
//   1. prefetches, two items ahead, the camera's source bounding box into its own shared-memory
//      stage with one TMA tensor copy (cp.async.bulk.tensor.2d) signalled on its own mbarrier --
//      lane 0 is the producer, nothing is shared with other warps, so no warp ever waits for one;
//   2. computes one vertical run of 16 canvas pixels per lane: coordinates from the plan-verified
//      fp32 recurrence (or exact fp64 where the plan says so), 2x3 aligned words per pixel from the
//      staged box, OpenCV's fixed-point bilinear weights from a 1024-entry table, three channel dot
//      products (dp4a) -- bit-identical to (sum w*S + 2^14) >> 15;
//   3. after the block's last layer, writes the finished 16x96-byte block to the canvas as 16-byte
//      vector stores.
// EXACT blocks run the per-pixel fp64 path of stitch_exact.cuh; EMPTY blocks are zero-filled.
#pragma once

#include <cuda.h>   // CUtensorMap (type only; the encoder is fetched at run time)

#include "stitch_exact.cuh"
#include "stitch_plan.cuh"

namespace bs {

constexpr int kLutBytes = 1024 * 8;
constexpr int kStagePad = 128;                       // slack behind a stage for the 3rd word of the last texel

struct FastArgs {
    FrameArgs F;
    CUtensorMap tmap[kMaxCamsPerLaunch];   // source frame of camera i as a [h][pitch/4] u32 tensor
    const WorkItem *items;                 // the plan, flattened: blocks with two layers first (two adjacent items each)
    int n_units;                           // blocks in the plan
    int n_two;                             // of which the first n_two have two items
    int pitch[kMaxCamsPerLaunch];          // TMA box of camera i: row pitch in shared memory (bytes) ...
    int box_bytes[kMaxCamsPerLaunch];      // ... and total bytes (= what the mbarrier expects)
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
constexpr int kWarpsPerCta = 8;
constexpr int kFastThreads = 32 * kWarpsPerCta;
#ifndef BS_MIN_CTAS
#define BS_MIN_CTAS 2
#endif
constexpr int kStages = 2;                             // per warp: source boxes of items i and i+1
constexpr int kRecSlots = 4;                           // per warp: records of items i .. i+3
constexpr int kBlockRowBytes = 96;                     // 32 pixels
constexpr int kWarpOutBytes = kRun * kBlockRowBytes;   // 1536
constexpr int kRecBytes = 80;                          // sizeof(WorkItem)
constexpr int kWarpCtlBytes = (kRecSlots * kRecBytes + (kStages + kRecSlots) * 8 + 127) / 128 * 128;   // records + mbarriers, keeps every warp's region 128-byte aligned

__host__ __device__ constexpr int warp_smem_bytes(int stage_bytes)
{
    return kStages * stage_bytes + kWarpOutBytes + kWarpCtlBytes;
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
__device__ __forceinline__ void run_layer_full(const LaneCoef &lc, int pitch, uint32_t stage_origin, uint32_t lut_base,
                                               uint32_t out_blk, int lane)
{
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const uint32_t run_base = stage_origin + (uint32_t)((lc.by >> kFracBits) * pitch + (lc.bx >> kFracBits) * 3);
    const int step = lc.nyv * pitch + lc.nxv * 3;      // whole texels per canvas row, as a byte offset
    // software pipeline: the shared-memory loads of pixel k+1 are issued before the arithmetic of
    // pixel k, so a warp has work to do while its loads are in flight
    auto fetch = [&](int k) {
        const int v = k - kRun / 2;
        int xbits, ybits;
        lane_coord_bits(lc, (float)v, xbits, ybits);
        const uint32_t o = run_base + (uint32_t)(v * step) +
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

// General path: partially covered layers, blocks that stick out of the canvas, and blocks for which
// the plan found at least one pixel where the cheap coordinates differ (those 32x4 groups use
// the exact fp64 form).  Same arithmetic per pixel, with coverage tests and byte stores.
__device__ __noinline__ void run_layer_general(const FastArgs &A, const CamDev &cam, const LaneCoef &lc, int pitch,
                                               uint32_t stage_origin, uint32_t lut_base, uint32_t out_blk,
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
    // layout: [pad to an 8 KB boundary][weight table 8 KB][per warp: kStages stages | output block | records | mbarriers]
    const uint32_t raw_base = smem_u32(smem_raw);
    const uint32_t lut_base = (raw_base + 8191u) & ~8191u;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const uint32_t wbase = lut_base + kLutBytes + (uint32_t)(warp * warp_smem_bytes(A.stage_bytes));
    const uint32_t stage0 = wbase;                                             // 128-byte aligned
    const uint32_t out_blk = stage0 + (uint32_t)(kStages * A.stage_bytes);
    const uint32_t rec0 = out_blk + kWarpOutBytes;
    const uint32_t dbar0 = rec0 + kRecSlots * kRecBytes;                       // data barriers, one per stage
    const uint32_t rbar0 = dbar0 + kStages * 8;                                // record barriers, one per slot

    // Programmatic dependent launch: let the next frame's kernel start filling SMs as this one's CTAs
    // retire.  It may prefetch and composite, but touches the canvas only after griddepcontrol.wait,
    // i.e. after this grid has completed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    {
        uint2 *lut = reinterpret_cast<uint2 *>(smem_raw + (lut_base - raw_base));
        for (int i = tid; i < 1024; i += kFastThreads) lut[i] = lut_entry(i);
    }
    if (lane == 0) {
        for (int s = 0; s < kStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dbar0 + 8 * s), "r"(1));
        for (int s = 0; s < kRecSlots; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rbar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // Work split: block u of the plan goes to warp u mod W (W = warps in the grid).  The plan lists the
    // heaviest blocks first, so they spread evenly and no cost model is needed; at any moment the grid
    // works on W consecutive blocks, i.e. on neighbouring pixels, which keeps box overlaps in L2.
    const int W = (int)gridDim.x * kWarpsPerCta;
    const int gw = (int)blockIdx.x * kWarpsPerCta + warp;
    if (gw >= A.n_units) return;
    const int my_units = (A.n_units - 1 - gw) / W + 1;
    const int my_two = gw < A.n_two ? (A.n_two - 1 - gw) / W + 1 : 0;
    const int i0 = 0, i1 = my_units + my_two;       // this warp's items, numbered locally
    // cursor over this warp's items in order, used by lane 0 to fetch records ahead
    int cur_unit = gw, cur_layer = 0;
    auto next_item_index = [&]() {                  // global index of the cursor's item; then advance
        const int idx = cur_unit + min(cur_unit, A.n_two) + cur_layer;
        if (cur_layer == 0 && cur_unit < A.n_two) cur_layer = 1;
        else {
            cur_layer = 0;
            cur_unit += W;
        }
        return idx;
    };

    // lane 0 is this warp's producer ------------------------------------------------------------------
    // record of item i -> record slot, by an 80-byte bulk copy
    auto request_record = [&](uint32_t slot) {
        const int i = next_item_index();
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rbar0 + 8 * slot), "r"((uint32_t)kRecBytes)
                     : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         rec0 + kRecBytes * slot),
                     "l"(A.items + i), "r"((uint32_t)kRecBytes), "r"(rbar0 + 8 * slot)
                     : "memory");
    };
    // source box of the item whose record sits in `slot` -> stage; items without a box just arrive
    auto issue_box = [&](uint32_t slot, uint32_t rec_parity, uint32_t stage) {
        mbar_wait_addr(rbar0 + 8 * slot, rec_parity);
        const uint4 r = lds128(rec0 + kRecBytes * slot);
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
    if (lane == 0) {
        for (int d = 0; d < 3 && i0 + d < i1; ++d) request_record((uint32_t)d);
        issue_box(0u, 0u, 0u);
    }
    __syncwarp();

    bool prior_grid_done = false;     // griddepcontrol.wait executed (needed before any canvas access)
    bool wrote = false;               // some layer has written into the current block
    uint32_t stage = 0, dpar = 0;     // stage and data-barrier parity of the current item
    uint32_t slot = 0, rpar = 0;      // record slot and record-barrier parity of the current item
    uint32_t pstage = 1, pslot = 1, ppar = 0;   // producer cursor: item i+1 (its stage, record slot, record parity)
    for (int i = i0; i < i1; ++i) {
        if (lane == 0) {
            // prefetch: the box of item i+1 goes into the other stage (item i-1 is done with it), the
            // record of item i+3 into the slot item i-1 used
            if (i + 1 < i1) issue_box(pslot, ppar, pstage);
            if (i + 3 < i1) request_record((pslot + 2) & (kRecSlots - 1));
        }
        if (++pstage == kStages) pstage = 0;
        if (++pslot == kRecSlots) {
            pslot = 0;
            ppar ^= 1u;
        }
        __syncwarp();
        mbar_wait_addr(rbar0 + 8 * slot, rpar);
        const uint4 r = lds128(rec0 + kRecBytes * slot);
        const int tx0 = (int)(r.x & 0xFFFFu) * kBlockW, ty0 = (int)(r.x >> 16) * kBlockH;
        const uint32_t type = r.y & 0xFFu, flags = (r.y >> 8) & 0xFFu;
        mbar_wait_addr(dbar0 + 8 * stage, dpar);
        if (type == (uint32_t)kBlockFast) {
            const uint32_t c = (r.y >> 16) & 0xFFu;
            const CamDev &cam = A.F.cam[c];
            const int pitch = A.pitch[c];
            // shared address texel (0, 0) of this camera would have, were the whole frame staged
            const uint32_t stage_origin = stage0 + stage * (uint32_t)A.stage_bytes - r.w;
            const uint32_t groups = (r.y >> 24) & 0xFFu;
            const bool full = (flags & kItemFull) != 0;
            const LaneCoef lc = lane_coef(load_block_coef(rec0 + kRecBytes * slot), lane - kBlockW / 2);
            if (full && groups == 0u)
                run_layer_full(lc, pitch, stage_origin, lut_base, out_blk, lane);
            else
                run_layer_general(A, cam, lc, pitch, stage_origin, lut_base, out_blk, groups, full, !wrote, tx0 + lane, ty0, lane);
            wrote = true;
        }
        __syncwarp();                 // every lane is done with this item's stage and record
        if (flags & kItemLast) {
            if (!prior_grid_done) {
                asm volatile("griddepcontrol.wait;" ::: "memory");
                prior_grid_done = true;
            }
            if (type == (uint32_t)kBlockExact) {
                // lane takes the 4-pixel group (lane & 7) of rows (lane >> 3) + 4 j
                for (int j = 0; j < kRun / 4; ++j) exact_quad(A.F, tx0 + (lane & 7) * 4, ty0 + (lane >> 3) + 4 * j);
            } else if (!wrote) {
                if (!A.F.keep_canvas) store_block_checked(A.F, 0u, tx0, ty0, lane);   // nobody writes here
            } else if (tx0 + kBlockW <= A.F.W && ty0 + kRun <= A.F.H) {
                // flush: lane pair (2r, 2r+1) writes one full 32-byte sector of row r per instruction
                const uint32_t so = out_blk + (uint32_t)((lane >> 1) * kBlockRowBytes + (lane & 1) * 16);
                uint8_t *const gp = A.F.canvas + (long long)(ty0 + (lane >> 1)) * A.F.cpitch + (long long)tx0 * 3 + (lane & 1) * 16;
#pragma unroll
                for (int j = 0; j < 3; ++j) *reinterpret_cast<uint4 *>(gp + 32 * j) = lds128(so + 32u * j);
            } else {
                store_block_checked(A.F, out_blk, tx0, ty0, lane);
            }
            wrote = false;
            __syncwarp();             // the output block is free again
        }
        if (++stage == kStages) {
            stage = 0;
            dpar ^= 1u;
        }
        if (++slot == kRecSlots) {
            slot = 0;
            rpar ^= 1u;
        }
    }
}

}  // namespace bs
