// stitch_fast2.cuh -- the per-frame kernel for calibrations whose source boxes are compact
// (a 32x16 canvas block reads at most 18 source rows of 128 bytes: every near-unit-scale video
// panorama, BASELINE.json configs 2-5).  Same work split, same plan and the same arithmetic as
// stitch_fast.cuh (which remains the kernel for large boxes); what differs is how a pixel gets its taps:
//
//   * after the TMA box of an item has landed, the warp rewrites it once from packed 3-byte texels
//     into one aligned word per texel (B G R x): 3 LDS + 3 PRMT + 1 STS.128 per four texels;
//   * a pixel then loads its four taps as four aligned words at immediate offsets from ONE address
//     (no alignment shifts, no third word), pairs them with four PRMT and blends with six dp2a;
//   * the x and y coordinate polynomials are evaluated together with the packed fp32 instructions
//     of sm_100 (fma.rn.f32x2 / add.rn.f32x2): 3 instructions per pixel instead of 6;
//   * two raw stages + one converted buffer: the boxes of items n+1 and n+2 are in flight while item n
//     is rewritten and blended;
//   * blocks in which no pixel has both fractions 0 and whose source advances one row per canvas row
//     (the plan marks them) run a variant of the hot loop without the 2^16 weight fix-up and with the
//     row offsets as immediates.
// Replaces the camera loop body /root/reference/tools/stitch/main.cpp:145-152.
#pragma once

#include <type_traits>

#include "stitch_undistort.cuh"

namespace bs {

#ifndef BS2_RAW_PITCH
#define BS2_RAW_PITCH 144
#endif
constexpr int kRawPitch = BS2_RAW_PITCH;                 // bytes per staged source row: TMA box of 36 four-byte elements.  120 would do;
                                                         // 144 puts consecutive rows 4 banks apart, so the three rows one conversion
                                                         // step reads collide in two banks instead of in all of them
constexpr int kRawRows = 18;                             // rows a box may have
constexpr int kRawStage = (kRawPitch * kRawRows + 127) / 128 * 128;   // 2688
#ifndef BS2_CONV_PITCH
#define BS2_CONV_PITCH 160
#endif
constexpr int kConvGroups = 10;                          // groups of four texels converted per row
constexpr int kConvTexels = 4 * kConvGroups;             // 40
constexpr int kConvPitch = BS2_CONV_PITCH;               // bytes per converted row: 160 = the 40 texels; 256 keeps every row on the same
                                                         // banks, so lanes whose taps straddle two rows do not collide
static_assert(kConvPitch >= 4 * kConvTexels && kConvPitch % 32 == 0, "converted row pitch");
constexpr int kConvRegion = (kConvPitch * kRawRows + 127) / 128 * 128;   // 2944
#ifndef BS2_RAW_STAGES
#define BS2_RAW_STAGES 2
#endif
#ifndef BS2_REC_SLOTS
#define BS2_REC_SLOTS 16
#endif
constexpr int kRawStages = BS2_RAW_STAGES;               // boxes in flight ahead of the one being converted
constexpr int kRec2Slots = BS2_REC_SLOTS;                // record ring slots per warp (two halves)
constexpr int kWarp2Smem = kRawStages * kRawStage + kConvRegion + kWarpOutBytes + (kRec2Slots * kRecBytes + 32 + 127) / 128 * 128;
static_assert(kWarp2Smem % 128 == 0, "per-warp regions stay 128-byte aligned");
// BS_FLAG_FEATHER launches: one blend weight per pixel of the block, kept between a block's two layers
constexpr int kWarp2SmemFeather = kWarp2Smem + kBlockW * kRun * 4;
static_assert(kWarp2SmemFeather % 128 == 0, "weight block");
// lens launches: a second ring beside the records with every item's third-order terms (32 bytes each)
constexpr int kWarp2SmemLens = kWarp2Smem + kRec2Slots * (int)sizeof(CubicCoef);
static_assert(sizeof(CubicCoef) == 32 && kWarp2SmemLens % 128 == 0, "cubic ring");
static_assert(kRawStages == 2, "the item pipeline is written for two raw stages");

#ifndef BS2_WARPS_PER_CTA
#define BS2_WARPS_PER_CTA 8
#endif
#ifndef BS2_MIN_CTAS
#define BS2_MIN_CTAS 2
#endif
constexpr int kWarps2 = BS2_WARPS_PER_CTA;
constexpr int kFast2Threads = 32 * kWarps2;

struct Fast2Args {
    FrameArgs F;
    CUtensorMap tmap[kMaxCamsPerLaunch];   // source frame of camera i as a [h][pitch/4] u32 tensor, box 32 x box_h[i]
    const WorkItem *items;
    const int *warp_start;
    int box_bytes;                         // 128 * box rows (all cameras use the same box; what the mbarrier expects)
    int wait_sources;                      // 1: the source frames may come from the kernel just before this one in
                                           // the stream: order every source read behind it (griddepcontrol.wait first)
    // Lens launches only (stitch_fast2_kernel<true>: cv::undistort of one frame of each of up to 16 cameras in ONE
    // launch, stitch_undistort.cuh).  An item's camera field names the camera c it belongs to: F.cam[c] is the
    // distorted frame, lens_dst[c] the frame to write (same size), lens_models[c] the model its EXACT blocks and
    // flagged rows evaluate (BORDER_CONSTANT 0), and item k's expansion has the third-order terms lens_cubic[k].
    const UndistortArgs *lens_models;      // device array indexed by camera; the frame pointers in it are not used
    const CubicCoef *lens_cubic;
    uint8_t *lens_dst[kMaxCamsPerLaunch];
    long long lens_dpitch[kMaxCamsPerLaunch];
};

// ---- packed fp32 pairs (x in the low half, y in the high half) -------------------------------------
__device__ __forceinline__ uint64_t pack2(float lo, float hi)
{
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, int &lo, int &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c)
{
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b)
{
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// lane_coef() of stitch_plan.cuh with x and y held as pairs.  Each half executes exactly the fma
// sequence the plan kernel verified (fma.rn.f32x2 is two independent IEEE fma.rn.f32).
struct LaneCoef2 {
    uint64_t a0, a1, a2;   // s(v) = a0 + v (a1 + v a2), (x, y)
};
__device__ __forceinline__ LaneCoef2 lane_coef2(const BlockCoef &c, float uf)
{
    const uint64_t u2 = pack2(uf, uf);
    LaneCoef2 l;
    l.a0 = fma2(u2, fma2(u2, pack2(c.hxuu, c.hyuu), pack2(c.xu, c.yu)), pack2(c.fx, c.fy));
    l.a1 = fma2(u2, pack2(c.xuv, c.yuv), pack2(c.xv, c.yv));
    l.a2 = pack2(c.hxvv, c.hyvv);
    return l;
}
// The lens-plan form (lane_coef_cubic() of stitch_plan.cuh): third-order terms in the column offset.
__device__ __forceinline__ LaneCoef2 lane_coef2_cubic(const BlockCoef &c, const CubicCoef &q, float uf)
{
    const uint64_t u2 = pack2(uf, uf);
    LaneCoef2 l;
    l.a0 = fma2(u2, fma2(u2, fma2(u2, pack2(q.xuuu, q.yuuu), pack2(c.hxuu, c.hyuu)), pack2(c.xu, c.yu)), pack2(c.fx, c.fy));
    l.a1 = fma2(u2, fma2(u2, pack2(q.xuuv, q.yuuv), pack2(c.xuv, c.yuv)), pack2(c.xv, c.yv));
    l.a2 = fma2(u2, pack2(q.xuvv, q.yuvv), pack2(c.hxvv, c.hyvv));
    return l;
}
// What only the partially covered layers need: the whole-texel parts, to test a pixel's source texel
// against the frame (X = bits_x + bx + 32 nxv v, as in lane_coef()).
struct LaneBase2 {
    int bx, by, nxv, nyv;
};
__device__ __forceinline__ LaneBase2 lane_base2(const BlockCoef &c, int u)
{
    LaneBase2 l;
    l.bx = c.ix - kMagicBias + 32 * (int)c.nxu * u;
    l.by = c.iy - kMagicBias + 32 * (int)c.nyu * u;
    l.nxv = c.nxv;
    l.nyv = c.nyv;
    return l;
}
__device__ __forceinline__ void lane_coord_bits2(const LaneCoef2 &l, float vf, int &xbits, int &ybits)
{
    const uint64_t v2 = pack2(vf, vf);
    const uint64_t s = fma2(v2, fma2(v2, l.a2, l.a1), l.a0);
    unpack2(add2(s, pack2(12582912.0f, 12582912.0f)), xbits, ybits);
}

// ---- converted box: one word per texel ---------------------------------------------------------
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// One conversion step: three raw words (four packed texels) -> four texel words.
struct ConvWords {
    uint32_t w0, w1, w2;      // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
};
__device__ __forceinline__ ConvWords conv_load(uint32_t s)
{
    ConvWords c;
    c.w0 = lds32(s);
    c.w1 = lds32(s + 4);
    c.w2 = lds32(s + 8);
    return c;
}
__device__ __forceinline__ void conv_store(uint32_t d, const ConvWords &c)
{
    sts128(d, c.w0, __byte_perm(c.w0, c.w1, 0x6543), __byte_perm(c.w1, c.w2, 0x5432), c.w2 >> 8);
}
constexpr int kConvSteps = kRawRows / 3;
#ifndef BS2_ROWS_PER_PASS
#define BS2_ROWS_PER_PASS 16
#endif
constexpr int kRowsPerPass = BS2_ROWS_PER_PASS;                          // rows of the hot loop's unrolled body
static_assert(kRun % kRowsPerPass == 0, "pass shape");
// `src`: this lane's first raw word (row slot/10 of the stage, texel group slot%10, slot = min(lane, 29):
// lanes 30 and 31 repeat the work of 28 and 29), `dst`: its first converted word.  Six steps of three rows.
__device__ __forceinline__ void convert_box(uint32_t src, uint32_t dst)
{
    ConvWords c[kConvSteps];       // every load is issued before the first store: one round trip, not six
#pragma unroll
    for (int it = 0; it < kConvSteps; ++it) c[it] = conv_load(src + (uint32_t)(it * 3 * kRawPitch));
#pragma unroll
    for (int it = 0; it < kConvSteps; ++it) conv_store(dst + (uint32_t)(it * 3 * kConvPitch), c[it]);
}

struct Taps2 {
    uint32_t t00, t01, t10, t11;   // texels (sx, sy), (sx+1, sy), (sx, sy+1), (sx+1, sy+1) as B G R x
    uint2 wt;
};
template <int kImm>
__device__ __forceinline__ uint32_t lds32_at(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(kImm));
    return v;
}
template <int kOff = 0>
__device__ __forceinline__ Taps2 fetch_taps2(uint32_t o, uint2 wt)
{
    Taps2 t;
    t.t00 = lds32_at<kOff>(o);
    t.t01 = lds32_at<kOff + 4>(o);
    t.t10 = lds32_at<kOff + kConvPitch>(o);
    t.t11 = lds32_at<kOff + kConvPitch + 4>(o);
    t.wt = wt;
    return t;
}
// Same blend as blend_taps(): bit-identical to (sum w*S + 2^14) >> 15 per channel.
__device__ __forceinline__ uint32_t blend_taps2(const Taps2 &t)
{
    const uint32_t p0 = __byte_perm(t.t00, t.t01, 0x5140), q0 = __byte_perm(t.t00, t.t01, 0x6262);   // [B00 B01 G00 G01], [R00 R01 ..]
    const uint32_t p1 = __byte_perm(t.t10, t.t11, 0x5140), q1 = __byte_perm(t.t10, t.t11, 0x6262);
    const uint32_t vb = __dp2a_lo(t.wt.y, p1, __dp2a_lo(t.wt.x, p0, 32768u));
    const uint32_t vg = __dp2a_hi(t.wt.y, p1, __dp2a_hi(t.wt.x, p0, 32768u));
    const uint32_t vr = __dp2a_lo(t.wt.y, q1, __dp2a_lo(t.wt.x, q0, 32768u));
    const uint32_t bg = __byte_perm(vb, vg, 0x3362);
    return __byte_perm(bg, vr, 0x7610);
}
__device__ __forceinline__ uint32_t sample_conv(uint32_t o, uint2 wt) { return blend_taps2(fetch_taps2<0>(o, wt)); }

__device__ __forceinline__ void store_px_bytes(uint32_t p8, uint32_t px)
{
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8), "r"(px) : "memory");
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 1), "r"(px >> 8) : "memory");
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(p8 + 2), "r"(px >> 16) : "memory");
}

// Address of a pixel's first tap in the converted buffer: `base` + whole texels of (xbits, ybits).
// Written so that each coordinate costs a mask and one multiply-add / shift-add.
__device__ __forceinline__ uint32_t tap_addr(uint32_t base, int xbits, int ybits)
{
    const uint32_t ym = (uint32_t)ybits & ~31u, xm = (uint32_t)xbits & ~31u;
    return __umulhi(xm, 1u << 29) + (ym * (uint32_t)(kConvPitch / 32) + base);      // LEA.HI + IMAD
}
// The hot loop's form of the same address.  `s` = the coordinate pair after the magic add, i.e. the floats
// 1.5 * 2^23 + X (X = the 1/32-pixel integers).  fma.rm(s, 1/32, 2^23) = floor(393216 + X / 32 + 2^23) exactly
// (the sum lies in [2^23, 2^24): one ulp is 1, round-down is floor), so its bits are kTexelBias + floor(X / 32) and
// the address is base + 4 qx + pitch qy with the bias folded into `base` (tap_base_bias): FFMA2 + IMAD + LEA
// instead of two masks + IMAD + LEA.HI.
constexpr uint32_t kTexelBias = 0x4B000000u + 393216u;
// tap_addr() sees the texel indices with the bias of the magic add, (0x4B400000 >> 5), and the host folded that
// into the item's run base; tap_addr_q() sees them with kTexelBias: the difference, mod 2^32
constexpr uint32_t kTapBaseBias = (kTexelBias - (0x4B400000u >> 5)) * (4u + (uint32_t)kConvPitch);
__device__ __forceinline__ uint64_t fma2_rm(uint64_t a, uint64_t b, uint64_t c)
{
    uint64_t d;
    asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint32_t tap_addr_q(uint32_t biased_base, uint64_t s)
{
    int qx, qy;
    unpack2(fma2_rm(s, pack2(0.03125f, 0.03125f), pack2(8388608.0f, 8388608.0f)), qx, qy);
    uint32_t a;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a) : "r"((uint32_t)qy), "n"(kConvPitch), "r"(biased_base));
    return a + 4u * (uint32_t)qx;
}
// tap_weights() of stitch_fast.cuh with the subtraction kept as one shift-and-subtract (LEA).
__device__ __forceinline__ uint2 tap_weights2(uint32_t a, uint32_t b)
{
    const uint32_t a2 = a * (0xFFFFu * 64u) + 2048u;
    const uint32_t w23 = b * a2;
    uint32_t w01;
    asm("mad.lo.u32 %0, %1, 32, %2;" : "=r"(w01) : "r"(a2), "r"(0u - w23));
    if (w01 == 0x10000u) w01 = 0xFFFFu;
    return make_uint2(w01, w23);
}
// For blocks in which the plan found no pixel with a = b = 0: the 2^16 case cannot occur.
__device__ __forceinline__ uint2 tap_weights2_nz(uint32_t a, uint32_t b)
{
    const uint32_t a2 = a * (0xFFFFu * 64u) + 2048u;
    const uint32_t w23 = b * a2;
    uint32_t w01;
    asm("mad.lo.u32 %0, %1, 32, %2;" : "=r"(w01) : "r"(a2), "r"(0u - w23));
    return make_uint2(w01, w23);
}

template <int K, int N, typename F>
__device__ __forceinline__ void static_for(F &&f)
{
    if constexpr (K < N) {
        f(std::integral_constant<int, K>{});
        static_for<K + 1, N>(f);
    }
}

// Hot path: the layer covers every pixel of the block with interior samples and the plan verified the
// cheap coordinates for the whole block.  `run0` = converted-buffer address of this lane's texel base
// for row v = -kRun/2 (whole-texel parts only), `vstep` = whole texels per canvas row as a byte offset.
// kDirect: every row goes straight to the canvas, `gp` = address of this lane's word of the block's
// first row; otherwise rows are collected in the warp's output block in shared memory.
// kPlain (with kDirect): no pixel of the block has both fractions 0 and a canvas row advances the source by
// exactly one converted row, so the weight fix-up is dropped and the row offsets become immediates.
template <bool kDirect, bool kPlain>
__device__ __forceinline__ void run_full2(const LaneCoef2 &lc, uint32_t run0, int vstep, uint32_t out_blk, int lane,
                                          uint64_t gp, uint32_t cpitch, uint32_t sel, uint32_t storing)
{
    // The 16 rows run as kRun / kRowsPerPass passes of an unrolled body (one pass by default: two passes of
    // eight rows halve the code, but cost four more instructions per row and were 7 % slower).
    float vf = (float)(-kRun / 2);          // row offset v of the pixel being fetched: -8, -7, ... (exact in fp32)
    uint32_t rbase = run0 - kTapBaseBias;   // converted-buffer base of the current pass (tap_addr_q's form)
    auto fetch = [&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const uint64_t v2 = pack2(vf, vf);
        const uint64_t sxy = fma2(v2, fma2(v2, lc.a2, lc.a1), lc.a0);
        int xbits, ybits;
        const uint64_t sm = add2(sxy, pack2(12582912.0f, 12582912.0f));
        unpack2(sm, xbits, ybits);
        vf += 1.0f;
        if constexpr (kPlain)
            return fetch_taps2<j * kConvPitch>(tap_addr_q(rbase, sm), tap_weights2_nz((uint32_t)xbits & 31u, (uint32_t)ybits & 31u));
        else
            return fetch_taps2<0>(tap_addr_q(rbase + (uint32_t)(j * vstep), sm),
                                  tap_weights2((uint32_t)xbits & 31u, (uint32_t)ybits & 31u));
    };
    Taps2 cur = fetch(std::integral_constant<int, 0>{});
#pragma unroll 1
    for (int pass = 0; pass < kRun / kRowsPerPass; ++pass) {
        const bool last_pass = pass == kRun / kRowsPerPass - 1;
        static_for<0, kRowsPerPass>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            Taps2 nxt = cur;
            if constexpr (j + 1 < kRowsPerPass) {
                nxt = fetch(std::integral_constant<int, j + 1>{});
            } else {
                // first pixel of the next pass (not needed after the last one; fetched anyway from a valid
                // address to keep the body branch-free is not possible here: rows beyond the box may not exist)
                rbase += (uint32_t)(kRowsPerPass * (kPlain ? kConvPitch : vstep));
                if (!last_pass) nxt = fetch(std::integral_constant<int, 0>{});
            }
            const uint32_t px = blend_taps2(cur);
            if constexpr (kDirect) {
                // this lane's word of the row, then on to the next row: the canvas does not straddle a 4 GB
                // boundary (the host checks), so only the low half of the address advances, in place
                const uint32_t nb = __shfl_down_sync(0xffffffffu, px, 1);
                const uint32_t word = __byte_perm(px, nb, sel);
                asm volatile(
                    "{\n"
                    ".reg .pred p;\n"
                    ".reg .b32 lo, hi;\n"
                    "setp.ne.u32 p, %2, 0;\n"
                    "@p st.global.u32 [%0], %1;\n"
                    "mov.b64 {lo, hi}, %0;\n"
                    "add.u32 lo, lo, %3;\n"
                    "mov.b64 %0, {lo, hi};\n"
                    "}"
                    : "+l"(gp)
                    : "r"(word), "r"(storing), "r"(cpitch)
                    : "memory");
            } else {
                store_row_pixel(px, out_blk + (uint32_t)(j * kBlockRowBytes), lane, sel);
            }
            cur = nxt;
        });
        if constexpr (!kDirect) out_blk += kRowsPerPass * kBlockRowBytes;
    }
}

// Partially covered layer whose cheap coordinates hold everywhere: coverage test and byte stores per pixel.
__device__ __noinline__ void run_masked2(const LaneCoef2 lc, const LaneBase2 lb, uint32_t run0, int vstep, uint32_t out_blk, int lane,
                                         int src_w, int src_h, bool col_in_canvas, int rows_in_canvas, bool first_layer)
{
    const uint32_t safe = out_blk - 2 * kConvPitch;     // inside the converted buffer: valid for every tap offset
    auto fetch = [&](int k, bool &in) {
        const int v = k - kRun / 2;
        int xbits, ybits;
        lane_coord_bits2(lc, (float)v, xbits, ybits);
        const int sx = (xbits + lb.bx + 32 * lb.nxv * v) >> kFracBits, sy = (ybits + lb.by + 32 * lb.nyv * v) >> kFracBits;
        in = col_in_canvas && k < rows_in_canvas && (uint32_t)sx < (uint32_t)(src_w - 1) && (uint32_t)sy < (uint32_t)(src_h - 1);
        const uint32_t o = tap_addr(run0 + (uint32_t)(k * vstep), xbits, ybits);
        return fetch_taps2(in ? o : safe, tap_weights2((uint32_t)xbits & 31u, (uint32_t)ybits & 31u));
    };
    bool in_cur, in_nxt = false;
    Taps2 cur = fetch(0, in_cur);
#pragma unroll 4
    for (int k = 0; k < kRun; ++k) {
        Taps2 nxt = cur;
        if (k + 1 < kRun) nxt = fetch(k + 1, in_nxt);
        uint32_t px = blend_taps2(cur);
        if (in_cur || first_layer) {
            if (!in_cur) px = 0;
            store_px_bytes(out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane, px);
        }
        cur = nxt;
        in_cur = in_nxt;
    }
}

// BS_FLAG_FEATHER, a block two cameras write into (include/b200_stitch.h): the first layer leaves its pixels in the
// output block and their weights d = 1 + distance to the nearest source edge (1/32 pixel) in `wgt`; the second layer
// blends where both cover, (d0 P0 + d1 P1 + (d0 + d1) / 2) / (d0 + d1) per channel, writes its own pixel where only it
// covers and leaves the rest.  Same fetch as run_masked2 (coverage = interior samples; border-rule pixels made the
// plan send the block to the per-pixel path).
__device__ __noinline__ void run_feather2(const LaneCoef2 lc, const LaneBase2 lb, uint32_t run0, int vstep, uint32_t out_blk, uint32_t wgt,
                                          int lane, int src_w, int src_h, bool col_in_canvas, int rows_in_canvas, bool first_layer)
{
    const uint32_t safe = out_blk - 2 * kConvPitch;     // inside the converted buffer: valid for every tap offset
    const int xmax = 32 * (src_w - 1), ymax = 32 * (src_h - 1);
#pragma unroll 2
    for (int k = 0; k < kRun; ++k) {
        const int v = k - kRun / 2;
        int xbits, ybits;
        lane_coord_bits2(lc, (float)v, xbits, ybits);
        const int X = xbits + lb.bx + 32 * lb.nxv * v, Y = ybits + lb.by + 32 * lb.nyv * v;
        const int sx = X >> kFracBits, sy = Y >> kFracBits;
        const bool in = col_in_canvas && k < rows_in_canvas && (uint32_t)sx < (uint32_t)(src_w - 1) && (uint32_t)sy < (uint32_t)(src_h - 1);
        const uint32_t o = tap_addr(run0 + (uint32_t)(k * vstep), xbits, ybits);
        const uint32_t px = blend_taps2(fetch_taps2(in ? o : safe, tap_weights2((uint32_t)xbits & 31u, (uint32_t)ybits & 31u)));
        const uint32_t d = in ? (uint32_t)max(1, 1 + min(min(X, xmax - X), min(Y, ymax - Y))) : 0u;
        const uint32_t p8 = out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane;
        const uint32_t wa = wgt + 4u * (uint32_t)(k * kBlockW + lane);
        if (first_layer) {
            store_px_bytes(p8, in ? px : 0u);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(wa), "r"(d) : "memory");
        } else if (in) {
            const uint32_t d0 = lds32(wa);
            uint32_t res = px;
            if (d0 != 0u) {
                uint32_t b0, g0, r0;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(b0) : "r"(p8));
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(g0) : "r"(p8 + 1));
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r0) : "r"(p8 + 2));
                const uint32_t den = d0 + d, half = den >> 1;
                const float rden = __frcp_rn((float)den);           // den < 2^21: exact as a float
                auto quot = [&](uint32_t num) {                     // floor(num / den), num < 2^29: estimate, then one correction
                    uint32_t q = (uint32_t)(__uint2float_rz(num) * rden);
                    int rem = (int)(num - q * den);
                    if (rem < 0) --q;
                    else if (rem >= (int)den) ++q;
                    return q;
                };
                res = quot(d0 * b0 + d * (px & 0xFFu) + half) | (quot(d0 * g0 + d * ((px >> 8) & 0xFFu) + half) << 8) |
                      (quot(d0 * r0 + d * ((px >> 16) & 0xFFu) + half) << 16);
            }
            store_px_bytes(p8, res);
        }
    }
}

// General path: partially covered layers with flagged 32x4 groups (those use the exact fp64 coordinates).
// `conv_origin` = address texel (0, 0) of the camera would have in the converted buffer.
__device__ __noinline__ void run_general2(const FrameArgs &F, const CamDev &cam, const LaneCoef2 lc, const LaneBase2 lb,
                                          uint32_t conv_origin, uint32_t out_blk, uint32_t group_mask, bool first_layer, int x, int y0, int lane)
{
    const int xb = (x / F.block_w) * F.block_w;
    const uint32_t safe = out_blk - 2 * kConvPitch;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        int X, Y;
        if ((group_mask >> (k >> 2)) & 1) {     // warp-uniform
            coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        } else {
            const int v = k - kRun / 2;
            int xbits, ybits;
            lane_coord_bits2(lc, (float)v, xbits, ybits);
            X = xbits + lb.bx + 32 * lb.nxv * v;
            Y = ybits + lb.by + 32 * lb.nyv * v;
        }
        const int sx = X >> kFracBits, sy = Y >> kFracBits;
        const bool in = (x < F.W) && (y0 + k < F.H) && sx >= 0 && sy >= 0 && sx < cam.w - 1 && sy < cam.h - 1;
        const uint32_t o = in ? conv_origin + (uint32_t)(sy * kConvPitch + sx * 4) : safe;
        uint32_t px = sample_conv(o, tap_weights2((uint32_t)X & 31u, (uint32_t)Y & 31u));
        if (in || first_layer) {
            if (!in) px = 0;
            store_px_bytes(out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane, px);
        }
    }
}

// Fully covered layer with flagged groups: redo the flagged rows of the output block with exact coordinates.
__device__ __noinline__ void patch_exact2(const FrameArgs &F, const CamDev &cam, uint32_t conv_origin, uint32_t out_blk,
                                          uint32_t group_mask, int x, int y0, int lane)
{
    const int xb = (x / F.block_w) * F.block_w;
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        if (!((group_mask >> (k >> 2)) & 1)) continue;     // warp-uniform
        int X, Y;
        coord_exact(cam.M, xb, x - xb, y0 + k, X, Y);
        const uint32_t o = conv_origin + (uint32_t)((Y >> kFracBits) * kConvPitch + (X >> kFracBits) * 4);
        const uint32_t px = sample_conv(o, tap_weights2((uint32_t)X & 31u, (uint32_t)Y & 31u));
        store_px_bytes(out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane, px);
    }
}

// The same for a lens plan: the flagged rows are redone with the lens model's exact coordinates.
__device__ __noinline__ void patch_lens2(const UndistortArgs &U, uint32_t conv_origin, uint32_t out_blk, uint32_t group_mask, int x,
                                         int y0, int lane)
{
#pragma unroll 1
    for (int k = 0; k < kRun; ++k) {
        if (!((group_mask >> (k >> 2)) & 1)) continue;     // warp-uniform
        int sx, sy, a, b;
        undistort_coord(U, x, y0 + k, sx, sy, a, b);
        const uint32_t px = sample_conv(conv_origin + (uint32_t)(sy * kConvPitch + sx * 4), tap_weights2((uint32_t)a, (uint32_t)b));
        store_px_bytes(out_blk + (uint32_t)(k * kBlockRowBytes) + 3u * (uint32_t)lane, px);
    }
}

// kLens = false: the panorama frame kernel.  kLens = true: the same loop over a lens plan (stitch_undistort.cuh).
// kFeather (with kLens = false): the BS_FLAG_FEATHER build of the frame kernel (two-layer blocks are blended).
template <bool kLens, bool kFeather = false>
__global__ void __launch_bounds__(kFast2Threads, BS2_MIN_CTAS) stitch_fast2_kernel(const __grid_constant__ Fast2Args A)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    // layout: [pad to a 128-byte boundary][per warp: raw stage 0 | raw stage 1 | converted box | output block | records | mbarriers]
    const uint32_t raw_base = (smem_u32(smem_raw) + 127u) & ~127u;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const uint32_t raw0 = raw_base + (uint32_t)(warp * (kLens ? kWarp2SmemLens : (kFeather ? kWarp2SmemFeather : kWarp2Smem)));   // 128-byte aligned
    const uint32_t conv0 = raw0 + kRawStages * kRawStage;
    const uint32_t out_blk = conv0 + kConvRegion;
    const uint32_t rec0 = out_blk + kWarpOutBytes;
    constexpr uint32_t rec_slots = kRec2Slots, rec_half = kRec2Slots / 2;
    const uint32_t dbar0 = rec0 + rec_slots * kRecBytes;                       // data barriers, one per raw stage
    const uint32_t rbar0 = dbar0 + 16;                                         // record barriers, one per half of the ring
    const uint32_t cub0 = raw0 + kWarp2Smem;                                   // lens launches: ring of third-order terms
    const uint32_t wgt0 = raw0 + kWarp2Smem;                                   // feather launches: the block's weights

    // Programmatic dependent launch, as in stitch_fast_kernel: the next frame's kernel may start at once;
    // it touches the canvas only after griddepcontrol.wait.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (lane == 0) {
        for (int s = 0; s < kRawStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dbar0 + 8 * s), "r"(1));
        for (int s = 0; s < 2; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rbar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const int gw = (int)blockIdx.x * kWarps2 + warp;
    const int i0 = A.warp_start[gw];
    const int cnt = A.warp_start[gw + 1] - i0;
    if (cnt <= 0) return;

    // lane 0 is this warp's producer ------------------------------------------------------------------
    auto request_records = [&](int first) {
        const uint32_t n_rec = (uint32_t)min((int)rec_half, cnt - first);
        const uint32_t bytes = n_rec * kRecBytes, cbytes = kLens ? n_rec * (uint32_t)sizeof(CubicCoef) : 0u;
        const uint32_t half = ((uint32_t)first / rec_half) & 1u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rbar0 + 8 * half), "r"(bytes + cbytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         rec0 + kRecBytes * rec_half * half),
                     "l"(A.items + i0 + first), "r"(bytes), "r"(rbar0 + 8 * half)
                     : "memory");
        if constexpr (kLens)        // the items' third-order terms ride on the same barrier
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             cub0 + (uint32_t)sizeof(CubicCoef) * rec_half * half),
                         "l"(A.lens_cubic + i0 + first), "r"(cbytes), "r"(rbar0 + 8 * half)
                         : "memory");
    };
    // source box of the item whose header words are `r` -> raw stage; items without a box just arrive
    const uint64_t tmap0 = reinterpret_cast<uint64_t>(&A.tmap[0]);
    static_assert(sizeof(CUtensorMap) == 128, "tensor map stride");
    auto issue_box = [&](const uint4 &r, uint32_t stage) {
        if ((r.y & 0xFFu) == (uint32_t)kBlockFast) {
            const uint32_t cam = (r.y >> 16) & 0xFFu;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dbar0 + 8 * stage),
                         "r"((uint32_t)A.box_bytes)
                         : "memory");
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                    raw0 + stage * (uint32_t)kRawStage),
                "l"(tmap0 + 128ull * cam), "r"((int)(int16_t)(r.z & 0xFFFFu)), "r"((int)(int16_t)(r.z >> 16)), "r"(dbar0 + 8 * stage)
                : "memory");
        } else {
            mbar_arrive_addr(dbar0 + 8 * stage);
        }
    };
    auto rec_addr = [&](int m) { return rec0 + kRecBytes * ((uint32_t)m & (rec_slots - 1)); };
    auto wait_half_of = [&](int m) {      // first item of a half: every lane observes the copy's completion itself
        if (((uint32_t)m & (rec_half - 1)) == 0) mbar_wait_addr(rbar0 + 8 * (((uint32_t)m / rec_half) & 1u), ((uint32_t)m / rec_slots) & 1u);
    };
    // item m: its data barrier and parity, its raw stage, and where this lane's conversion slot reads and writes
    const uint32_t slot = (uint32_t)min(lane, 3 * kConvGroups - 1);
    const uint32_t cv_src_lane = (slot / kConvGroups) * kRawPitch + 12 * (slot % kConvGroups);
    const uint32_t cv_dst_lane = (slot / kConvGroups) * kConvPitch + 16 * (slot % kConvGroups);
    auto box_landed_wait = [&](int m) { mbar_wait_addr(dbar0 + 8 * ((uint32_t)m & 1u), ((uint32_t)m >> 1) & 1u); };
    auto cv_src_of = [&](int m, uint32_t hdr_y) {      // flags bits 4-5: where the first converted texel sits in the staged row
        return raw0 + ((uint32_t)m & 1u) * (uint32_t)kRawStage + cv_src_lane + ((hdr_y >> (8 + kItemRawOffShift - 2)) & 12u);
    };
    const uint32_t conv = conv0, cv_dst = conv0 + cv_dst_lane;
    // address texel (0, 0) of an item's camera would have in its converted buffer (rare paths only)
    auto conv_origin_of = [&](uint32_t conv, const uint4 &h) {
        const int c0 = (int)(int16_t)(h.z & 0xFFFFu), y0 = (int)(int16_t)(h.z >> 16);
        const int t0 = (4 * c0 + (int)((h.y >> (8 + kItemRawOffShift - 2)) & 12u)) / 3;
        return conv - (uint32_t)(y0 * kConvPitch + t0 * 4);
    };

    if (lane == 0) {
        request_records(0);
        if (cnt > (int)rec_half) request_records((int)rec_half);
    }
    __syncwarp();
    wait_half_of(0);
    bool prior_grid_done = false;     // griddepcontrol.wait executed (needed before any canvas access)
    if (A.wait_sources) {             // only the plan (written long ago) has been read so far
        asm volatile("griddepcontrol.wait;" ::: "memory");
        prior_grid_done = true;
    }
    if (lane == 0) {
        issue_box(lds128(rec_addr(0)), 0u);
        if (cnt > 1) issue_box(lds128(rec_addr(1)), 1u);
    }
    __syncwarp();
    // lane constants: output word merge (store_row_pixel), column offsets
    const uint32_t sel = (lane & 3) == 0 ? 0x4210u : ((lane & 3) == 1 ? 0x5421u : 0x6542u);
    const uint32_t storing = (uint32_t)((lane & 3) != 3);
    const int u = lane - kBlockW / 2;
    const float uf = (float)u;
    // this lane's word of canvas row 0, block column 0; a block adds by * 16 * pitch + bx * 96 (32-bit: direct
    // stores need the canvas inside one 4 GB window anyway)
    const uint32_t lane_word = 4u * (uint32_t)(3 * (lane >> 2) + (lane & 3));
    const uint64_t gbase = reinterpret_cast<uint64_t>(A.F.canvas) + lane_word;
    const uint32_t cpitch = (uint32_t)A.F.cpitch, brow = cpitch * kBlockH;
    // where an item's pixels go: the canvas, or (lens launches) the undistorted frame of the item's camera
    struct Dst {
        uint8_t *base;
        long long pitch;
        int W, H;
    };
    auto dst_of = [&](uint32_t hdr_y) {
        if constexpr (kLens) {
            const uint32_t c = (hdr_y >> 16) & 0xFFu;
            return Dst{A.lens_dst[c], A.lens_dpitch[c], A.F.cam[c].w, A.F.cam[c].h};
        } else {
            return Dst{A.F.canvas, A.F.cpitch, A.F.W, A.F.H};
        }
    };
    auto lens_of = [&](uint32_t hdr_y) {     // the model of the item's camera with this launch's frames (rare paths)
        const uint32_t c = (hdr_y >> 16) & 0xFFu;
        UndistortArgs U = A.lens_models[c];
        U.src = A.F.cam[c].ptr;
        U.spitch = A.F.cam[c].pitch;
        U.dst = A.lens_dst[c];
        U.dpitch = A.lens_dpitch[c];
        return U;
    };
    bool wrote = false;               // some layer has written into the current block
    bool can_direct = prior_grid_done && A.F.canvas_in_4g;
    for (int n = 0; n < cnt; ++n) {
        if (lane == 0 && ((uint32_t)n & (rec_half - 1)) == 0 && n >= (int)rec_half && n + (int)rec_half < cnt)
            request_records(n + (int)rec_half);      // the half items n-8 .. n-1 used is free: refill it with n+8 .. n+15
        // the record ring is read two items ahead: make the next half visible when item n+2 is its first
        if (n + 2 < cnt) wait_half_of(n + 2);
        // item header (layout_plan, compact form): bx | by << 16, type | flags << 8 | cam << 16 | groups << 24,
        // c0 | y0 << 16 (TMA coordinates of the box), run base of lane u = 0 in the converted buffer
        const uint32_t ra = rec_addr(n);
        const uint4 r = lds128(ra);
        bool direct = false;
        const uint32_t bx = r.x & 0xFFFFu, by = r.x >> 16;
        const uint32_t type = r.y & 0xFFu, flags = (r.y >> 8) & 0xFFu;
        if (type == (uint32_t)kBlockFast) {
            box_landed_wait(n);
            convert_box(cv_src_of(n, r.y), cv_dst);
        }
        __syncwarp();                 // converted box complete; its raw stage takes the box of item n+2
        if (n + 2 < cnt && lane == 0) issue_box(lds128(rec_addr(n + 2)), (uint32_t)n & 1u);
        if (type == (uint32_t)kBlockFast) {
            const BlockCoef bc = load_block_coef(ra);
            LaneCoef2 lc;
            if constexpr (kLens) {
                CubicCoef q;
                const uint32_t qa = cub0 + (uint32_t)sizeof(CubicCoef) * ((uint32_t)n & (rec_slots - 1));
                const uint4 q0 = lds128(qa);
                const uint2 q1 = lds64(qa + 16);
                q.xuuu = __uint_as_float(q0.x); q.yuuu = __uint_as_float(q0.y); q.xuuv = __uint_as_float(q0.z);
                q.yuuv = __uint_as_float(q0.w); q.xuvv = __uint_as_float(q1.x); q.yuvv = __uint_as_float(q1.y);
                lc = lane_coef2_cubic(bc, q, uf);
            } else {
                lc = lane_coef2(bc, uf);
            }
            const int ustep = (int)(int16_t)(bc.aux & 0xFFFFu), vstep = (int)(int16_t)(bc.aux >> 16);
            const uint32_t run0 = conv + r.w + (uint32_t)(u * ustep);
            const uint32_t groups = r.y >> 24;
            if (kFeather && (flags & kItemFeather)) {
                const CamDev &cam = A.F.cam[(r.y >> 16) & 0xFFu];
                const int tx0 = (int)bx * kBlockW, ty0 = (int)by * kBlockH;
                run_feather2(lc, lane_base2(bc, u), run0, vstep, out_blk, wgt0, lane, cam.w, cam.h, tx0 + lane < A.F.W, A.F.H - ty0, !wrote);
            } else if (flags & kItemFull) {
                direct = can_direct && (flags & kItemDirect);
                uint64_t grow;
                uint32_t rpitch = cpitch;
                if constexpr (kLens) {
                    const uint32_t c = (r.y >> 16) & 0xFFu;
                    rpitch = (uint32_t)A.lens_dpitch[c];
                    grow = reinterpret_cast<uint64_t>(A.lens_dst[c]) + lane_word + (by * (rpitch * kBlockH) + bx * (uint32_t)kBlockRowBytes);
                } else {
                    grow = gbase + (by * brow + bx * (uint32_t)kBlockRowBytes);
                }
                if (direct && (flags & kItemPlain))
                    run_full2<true, true>(lc, run0, vstep, out_blk, lane, grow, rpitch, sel, storing);
                else if (direct)
                    run_full2<true, false>(lc, run0, vstep, out_blk, lane, grow, rpitch, sel, storing);
                else
                    run_full2<false, false>(lc, run0, vstep, out_blk, lane, grow, rpitch, sel, storing);
                if (groups != 0u) {
                    __syncwarp();
                    if constexpr (kLens)
                        patch_lens2(lens_of(r.y), conv_origin_of(conv, r), out_blk, groups, (int)bx * kBlockW + lane, (int)by * kBlockH, lane);
                    else
                        patch_exact2(A.F, A.F.cam[(r.y >> 16) & 0xFFu], conv_origin_of(conv, r), out_blk, groups, (int)bx * kBlockW + lane,
                                     (int)by * kBlockH, lane);
                }
            } else {
                const CamDev &cam = A.F.cam[(r.y >> 16) & 0xFFu];
                const LaneBase2 lb = lane_base2(bc, u);
                const int tx0 = (int)bx * kBlockW, ty0 = (int)by * kBlockH;
                const Dst d = dst_of(r.y);
                if (groups == 0u)
                    run_masked2(lc, lb, run0, vstep, out_blk, lane, cam.w, cam.h, tx0 + lane < d.W, d.H - ty0, !wrote);
                else          // (the lens plan has no partially covered layers: its blocks are full, EXACT or EMPTY)
                    run_general2(A.F, cam, lc, lb, conv_origin_of(conv, r), out_blk, groups, !wrote, tx0 + lane, ty0, lane);
            }
            wrote = !direct;
        }
        __syncwarp();                 // every lane is done with the converted buffer
        if ((flags & kItemLast) && !direct) {
            const int tx0 = (int)bx * kBlockW, ty0 = (int)by * kBlockH;
            if (!prior_grid_done) {
                asm volatile("griddepcontrol.wait;" ::: "memory");
                prior_grid_done = true;
                can_direct = A.F.canvas_in_4g != 0;
            }
            const Dst d = dst_of(r.y);
            if (type == (uint32_t)kBlockExact) {
                // lane takes the 4-pixel group (lane & 7) of rows (lane >> 3) + 4 j; the item names its row groups j
                const uint32_t rows = r.y >> 24;
                if constexpr (kLens) {
                    const UndistortArgs U = lens_of(r.y);
                    for (int j = 0; j < kRun / 4; ++j)
                        if ((rows >> j) & 1u) lens_quad(U, tx0 + (lane & 7) * 4, ty0 + (lane >> 3) + 4 * j);
                } else {
                    for (int j = 0; j < kRun / 4; ++j)
                        if ((rows >> j) & 1u) exact_quad(A.F, tx0 + (lane & 7) * 4, ty0 + (lane >> 3) + 4 * j);
                }
            } else if (!wrote && !kLens && A.F.keep_canvas) {
                // nobody writes here and the canvas keeps its contents
            } else if (tx0 + kBlockW <= d.W && ty0 + kRun <= d.H) {
                // flush (or zero fill of an empty block): lane pair (2r, 2r+1) writes one full 32-byte sector of row r per instruction
                const uint32_t so = out_blk + (uint32_t)((lane >> 1) * kBlockRowBytes + (lane & 1) * 16);
                uint8_t *const gp = d.base + (long long)(ty0 + (lane >> 1)) * d.pitch + (long long)tx0 * 3 + (lane & 1) * 16;
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    *reinterpret_cast<uint4 *>(gp + 32 * j) = wrote ? lds128(so + 32u * j) : make_uint4(0u, 0u, 0u, 0u);
            } else {
                store_block_checked(d.base, d.pitch, d.W, d.H, wrote ? out_blk : 0u, tx0, ty0, lane);
            }
            wrote = false;
            __syncwarp();             // the output block is free again
        }
    }
}

}  // namespace bs
