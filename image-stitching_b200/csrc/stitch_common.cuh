// stitch_common.cuh -- shared device-side definitions of the warp + composite path.
//
// Arithmetic reproduced: what cv::warpPerspective(..., INTER_LINEAR, BORDER_TRANSPARENT)
// computes for CV_8UC3 at /root/reference/tools/stitch/main.cpp:116-117 and :150-151
// (OpenCV imgproc, un-vendored; algorithm as specified in SURVEY.md Appendix A).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace bs {

constexpr int kFracBits = 5;            // OpenCV INTER_BITS
constexpr int kFracOne = 1 << kFracBits;
constexpr int kMaxCamsPerLaunch = 16;   // cameras carried in one kernel's parameter block

// One camera as the kernels see it.  M is the INVERSE homography (canvas -> source).
struct CamDev {
    double M[9];
    const uint8_t *ptr;   // packed BGR rows
    long long pitch;      // bytes
    int w, h;
};

struct FrameArgs {
    CamDev cam[kMaxCamsPerLaunch];
    uint8_t *canvas;
    long long cpitch;
    int n_cams;
    int W, H;
    int block_w;          // x-origin granularity of OpenCV's coordinate blocks (64 for canvases >= 64x16)
    int legacy_border;    // Appendix A.4 rule
    int keep_canvas;      // 1: uncovered pixels untouched, 0: uncovered pixels = 0
    int cam_base;         // index of cam[0] among all cameras (multi-launch for > 16 cameras)
    int canvas_in_4g;     // 1: all canvas addresses share their upper 32 bits (staged kernel: rows may be stored with 32-bit address steps)
    int border_constant;  // 1: cv::BORDER_CONSTANT 0 instead of BORDER_TRANSPARENT (exact kernel only)
    int feather;          // 1: BS_FLAG_FEATHER, pixels written by several cameras are a weighted mean (composite_feather)
};

// ---- Appendix A.1, exact form -------------------------------------------------------------
// Same operation order as OpenCV's scalar block loop, every operation individually rounded
// (no FMA contraction), IEEE divide.  The oracle (oracle/warp_oracle.c) is the same sequence.
__device__ __forceinline__ void coord_exact(const double *__restrict__ M, int xb, int x1, int y,
                                            int &X, int &Y)
{
    const double dxb = (double)xb, dy = (double)y, dx1 = (double)x1;
    const double X0 = __dadd_rn(__dadd_rn(__dmul_rn(M[0], dxb), __dmul_rn(M[1], dy)), M[2]);
    const double Y0 = __dadd_rn(__dadd_rn(__dmul_rn(M[3], dxb), __dmul_rn(M[4], dy)), M[5]);
    const double W0 = __dadd_rn(__dadd_rn(__dmul_rn(M[6], dxb), __dmul_rn(M[7], dy)), M[8]);
    double W = __dadd_rn(W0, __dmul_rn(M[6], dx1));
    W = (W != 0.0) ? __ddiv_rn((double)kFracOne, W) : 0.0;
    double fX = __dmul_rn(__dadd_rn(X0, __dmul_rn(M[0], dx1)), W);
    double fY = __dmul_rn(__dadd_rn(Y0, __dmul_rn(M[3], dx1)), W);
    // clamp to the int range, then round half to even (cvRound); cvt.rni.s32.f64 saturates.
    // NaN (only possible with a non-finite H) maps to INT_MIN = "not covered".
    X = (fX == fX) ? __double2int_rn(fX) : INT_MIN;
    Y = (fY == fY) ? __double2int_rn(fY) : INT_MIN;
}

__device__ __forceinline__ int sat_s16(int v)
{
    return max(-32768, min(32767, v));
}

// ---- Appendix A.2-A.4 from global memory ---------------------------------------------------
// Returns true when the pixel is written by this camera; bgr = B | G<<8 | R<<16.
__device__ __forceinline__ bool sample_exact(const CamDev &c, int X, int Y, int legacy,
                                             uint32_t &bgr)
{
    const int sx = sat_s16(X >> kFracBits), sy = sat_s16(Y >> kFracBits);
    if (sx < 0 || sy < 0 || sx > c.w - 1 || sy > c.h - 1) return false;
    const bool ok01 = sx + 1 < c.w, ok10 = sy + 1 < c.h;
    if (legacy && !(ok01 && ok10)) return false;
    const int a = X & (kFracOne - 1), b = Y & (kFracOne - 1);
    // weights / 32; the reference's 15-bit weights are these times 32
    const int w00 = (kFracOne - a) * (kFracOne - b), w01 = a * (kFracOne - b);
    const int w10 = (kFracOne - a) * b, w11 = a * b;
    const uint8_t *p0 = c.ptr + (long long)sy * c.pitch + (long long)sx * 3;
    const uint8_t *p1 = p0 + c.pitch;
    uint32_t out = 0;
    if (ok01 && ok10) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            int v = w00 * __ldg(p0 + ch) + w01 * __ldg(p0 + 3 + ch) + w10 * __ldg(p1 + ch) +
                    w11 * __ldg(p1 + 3 + ch);
            out |= (uint32_t)((v + 512) >> 10) << (8 * ch);   // == (32 v + 2^14) >> 15
        }
    } else {
        // last source row / column, OpenCV 4.x: drop the outside taps, renormalise in fp32
        const int wtot = 32 * (w00 + (ok01 ? w01 : 0) + (ok10 ? w10 : 0));
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            int t0 = w00 * __ldg(p0 + ch);
            if (ok01) t0 += w01 * __ldg(p0 + 3 + ch);
            if (ok10) t0 += w10 * __ldg(p1 + ch);
            t0 *= 32;
            const float q = __fdiv_rn(__fmul_rn((float)t0, 32768.0f), (float)wtot);
            const int t = __float2int_rz(q);
            out |= (uint32_t)max(0, min(255, (t + 16384) >> 15)) << (8 * ch);
        }
    }
    bgr = out;
    return true;
}

// Winner search for one canvas pixel: highest camera index that covers it (Appendix A.5).
// cv::BORDER_CONSTANT with value 0 (lib/Calibrate/utils.cpp:390): every pixel is
// written, taps outside the source count as 0.
__device__ __forceinline__ uint32_t sample_constant(const CamDev &c, int X, int Y)
{
    const int sx = sat_s16(X >> kFracBits), sy = sat_s16(Y >> kFracBits);
    const int a = X & (kFracOne - 1), b = Y & (kFracOne - 1);
    const int wgt[4] = {(kFracOne - a) * (kFracOne - b), a * (kFracOne - b), (kFracOne - a) * b, a * b};
    uint32_t out = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        int acc = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int tx = sx + (t & 1), ty = sy + (t >> 1);
            if (tx >= 0 && tx < c.w && ty >= 0 && ty < c.h) acc += wgt[t] * __ldg(c.ptr + (long long)ty * c.pitch + (long long)tx * 3 + ch);
        }
        out |= (uint32_t)((acc + 512) >> 10) << (8 * ch);   // == (32 acc + 2^14) >> 15
    }
    return out;
}

// BS_FLAG_FEATHER (nothing in the reference does this, the OpenCV call only overwrites):
// every camera that would write the pixel under the active border rule contributes its own bilinear result
// P_i with the integer weight d_i = 1 + its distance to the nearest source edge in 1/32 pixels,
//     d_i = max(1, 1 + min(X, 32 (w - 1) - X, Y, 32 (h - 1) - Y)),
// and the pixel is (sum d_i P_i + (sum d_i) / 2) / sum d_i per channel, integer division.  A pixel one camera
// writes is that camera's pixel unchanged, so blocks without overlap keep the staged path.
// d_i < 2^20, P_i < 2^8, at most 16 cameras: the sums fit 32 bits.  oracle_stitch_feather is the same sequence.
__device__ __forceinline__ bool composite_feather(const FrameArgs &A, int xb, int x1, int y, uint32_t &bgr)
{
    uint32_t acc0 = 0, acc1 = 0, acc2 = 0, wsum = 0;
    for (int i = 0; i < A.n_cams; ++i) {
        int X, Y;
        uint32_t p;
        coord_exact(A.cam[i].M, xb, x1, y, X, Y);
        if (!sample_exact(A.cam[i], X, Y, A.legacy_border, p)) continue;
        const int d = min(min(X, 32 * (A.cam[i].w - 1) - X), min(Y, 32 * (A.cam[i].h - 1) - Y));
        const uint32_t w = (uint32_t)max(1, d + 1);
        acc0 += w * (p & 0xffu);
        acc1 += w * ((p >> 8) & 0xffu);
        acc2 += w * ((p >> 16) & 0xffu);
        wsum += w;
    }
    if (wsum == 0) {
        bgr = 0;
        return false;
    }
    const uint32_t half = wsum >> 1;
    bgr = ((acc0 + half) / wsum) | (((acc1 + half) / wsum) << 8) | (((acc2 + half) / wsum) << 16);
    return true;
}

__device__ __forceinline__ bool composite_exact(const FrameArgs &A, int x, int y, uint32_t &bgr)
{
    const int xb = (x / A.block_w) * A.block_w;
    const int x1 = x - xb;
    if (A.feather) return composite_feather(A, xb, x1, y, bgr);
    if (A.border_constant) {            // the last camera writes every pixel
        int X, Y;
        coord_exact(A.cam[A.n_cams - 1].M, xb, x1, y, X, Y);
        bgr = sample_constant(A.cam[A.n_cams - 1], X, Y);
        return true;
    }
    for (int i = A.n_cams - 1; i >= 0; --i) {
        int X, Y;
        coord_exact(A.cam[i].M, xb, x1, y, X, Y);
        if (sample_exact(A.cam[i], X, Y, A.legacy_border, bgr)) return true;
    }
    bgr = 0;
    return false;
}

}  // namespace bs
