// stitch_undistort.cuh -- cv::undistort, the step the reference runs on every frame right before the
// warp in video mode (/root/reference/tools/stitch/main.cpp:148-149).  SURVEY.md row f1, Appendix A.6.
//
// OpenCV builds a fixed-point map with initUndistortRectifyMap (in row stripes of
// min(max(1, 4096 / w), h) rows, the principal point shifted per stripe) and resamples with the same
// 1/32-pixel bilinear kernel as the warp, BORDER_CONSTANT 0.  The normalised x coordinate is
// advanced along a row by repeated addition of 1/fx, so its value at column j carries j rounding
// steps; that sequence is identical for every row when the camera matrix has the usual form
// [[fx 0 cx] [0 fy cy] [0 0 1]], so the host tabulates it once (xs[j]) together with the per-row
// normalised y (ys[i]); the kernel then evaluates the distortion model per pixel in fp64 with
// individually rounded operations, in the order of the scalar reference (oracle/warp_oracle.c).
#pragma once

#include "stitch_common.cuh"

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

__global__ void __launch_bounds__(256) undistort_kernel(const __grid_constant__ UndistortArgs A)
{
    const int j = blockIdx.x * 64 + (threadIdx.x & 63);
    const int i = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (j >= A.w || i >= A.h) return;
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
    const double fu = __dmul_rn(u, (double)kFracOne), fv = __dmul_rn(v, (double)kFracOne);
    const int iu = (fu == fu) ? __double2int_rn(fu) : INT_MIN, iv = (fv == fv) ? __double2int_rn(fv) : INT_MIN;
    const int sx = (int)(short)(iu >> kFracBits), sy = (int)(short)(iv >> kFracBits);   // OpenCV stores int16
    const int a = iu & (kFracOne - 1), b = iv & (kFracOne - 1);
    const int wgt[4] = {(kFracOne - a) * (kFracOne - b), a * (kFracOne - b), (kFracOne - a) * b, a * b};
    uint8_t *o = A.dst + (long long)i * A.dpitch + (long long)j * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        int acc = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int tx = sx + (t & 1), ty = sy + (t >> 1);
            if (tx >= 0 && tx < A.w && ty >= 0 && ty < A.h)
                acc += wgt[t] * __ldg(A.src + (long long)ty * A.spitch + (long long)tx * 3 + c);   // outside taps are 0
        }
        o[c] = (uint8_t)((acc + 512) >> 10);   // == (32 acc + 2^14) >> 15
    }
}

}  // namespace bs
