/*
 * oracle/warp_oracle.c -- CPU restatement of the stitch hot path.  TEST INFRASTRUCTURE ONLY
 * (see warp_oracle.h for who may call it and how it is pinned).
 *
 * Follows SURVEY.md Appendix A, which specifies what OpenCV's cv::warpPerspective /
 * cv::remap / cv::undistort compute for CV_8UC3 + INTER_LINEAR; the reference calls them at
 * /root/reference/tools/stitch/main.cpp:116-117, :148-151.
 *
 * Build with -ffp-contract=off: the double arithmetic must not be fused.
 */
#include "warp_oracle.h"

#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

enum { FRAC_BITS = 5, FRAC_ONE = 1 << FRAC_BITS, COEF_BITS = 15 };

/* ------------------------------------------------------------------------------------------
 * H -> M = H^-1.  cv::warpPerspective inverts the forward homography it is handed
 * (reference call: tools/stitch/main.cpp:116, no WARP_INVERSE_MAP).  For 3x3 doubles OpenCV
 * uses the cofactor formula scaled by 1/det, not a pivoted LU; restated here term by term so
 * the doubles agree bit for bit (checked against cv2.invert in tests/test_oracle.py).
 * ---------------------------------------------------------------------------------------- */
int oracle_invert3x3(const double H[9], double M[9])
{
    const double a = H[0], b = H[1], c = H[2];
    const double d = H[3], e = H[4], f = H[5];
    const double g = H[6], h = H[7], i = H[8];
    double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (det == 0.0) {
        memset(M, 0, 9 * sizeof(double));
        return 0;
    }
    det = 1.0 / det;
    double t[9];
    t[0] = (e * i - f * h) * det;
    t[1] = (c * h - b * i) * det;
    t[2] = (b * f - c * e) * det;
    t[3] = (f * g - d * i) * det;
    t[4] = (a * i - c * g) * det;
    t[5] = (c * d - a * f) * det;
    t[6] = (d * h - e * g) * det;
    t[7] = (b * g - a * h) * det;
    t[8] = (a * e - b * d) * det;
    memcpy(M, t, sizeof t);
    return 1;
}

/* cvRound on a double already clamped to the int range: round-half-to-even. */
static inline int round_half_even(double v)
{
    return (int)lrint(v);
}

static inline double clamp_int_range(double v)
{
    if (v < (double)INT_MIN) v = (double)INT_MIN;
    if (v > (double)INT_MAX) v = (double)INT_MAX;
    return v;
}

/* Appendix A.1.  OpenCV walks the canvas in blocks and evaluates the x-term relative to the
 * block's first column; `xb` is that column (0 when the caller does not care: the survey
 * found the choice never changes a pixel, but the oracle keeps OpenCV's order anyway). */
static inline void coord_rel(const double M[9], int xb, int x1, int y, int *X, int *Y)
{
    double X0 = M[0] * xb + M[1] * y + M[2];
    double Y0 = M[3] * xb + M[4] * y + M[5];
    double W0 = M[6] * xb + M[7] * y + M[8];
    double W = W0 + M[6] * x1;
    W = W != 0.0 ? (double)FRAC_ONE / W : 0.0;
    double fX = clamp_int_range((X0 + M[0] * x1) * W);
    double fY = clamp_int_range((Y0 + M[3] * x1) * W);
    *X = round_half_even(fX);
    *Y = round_half_even(fY);
}

/* Block width OpenCV's invoker uses for a canvas of dw x dh (32*32 pixels per block,
 * at most 16 rows): the x-term is relative to multiples of this. */
static int block_width(int dw, int dh)
{
    int bh = dh < 16 ? dh : 16;
    if (bh < 1) bh = 1;
    int bw = 1024 / bh;
    if (bw > dw) bw = dw;
    if (bw < 1) bw = 1;
    return bw;
}

void oracle_warp_coord(const double M[9], int x, int y, int *X, int *Y)
{
    /* stand-alone form: block width 64, the value for any canvas at least 64x16 */
    int xb = (x / 64) * 64;
    coord_rel(M, xb, x - xb, y, X, Y);
}

static inline int sat_s16(int v)
{
    return v < -32768 ? -32768 : (v > 32767 ? 32767 : v);
}

static inline uint8_t sat_u8(int v)
{
    return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

/* Appendix A.2-A.4 for one canvas pixel; returns 1 if the pixel was written. */
static inline int sample_transparent(const uint8_t *src, int sw, int sh, size_t spitch,
                                     int X, int Y, int border_rule, uint8_t out[3])
{
    const int sx = sat_s16(X >> FRAC_BITS), sy = sat_s16(Y >> FRAC_BITS);
    const int a = X & (FRAC_ONE - 1), b = Y & (FRAC_ONE - 1);
    const int w00 = (FRAC_ONE - a) * (FRAC_ONE - b) * 32, w01 = a * (FRAC_ONE - b) * 32;
    const int w10 = (FRAC_ONE - a) * b * 32, w11 = a * b * 32;

    if (sx >= 0 && sy >= 0 && sx < sw - 1 && sy < sh - 1) { /* A.3 interior */
        const uint8_t *p0 = src + (size_t)sy * spitch + (size_t)sx * 3;
        const uint8_t *p1 = p0 + spitch;
        for (int c = 0; c < 3; ++c) {
            int v = w00 * p0[c] + w01 * p0[3 + c] + w10 * p1[c] + w11 * p1[3 + c];
            out[c] = (uint8_t)((v + (1 << (COEF_BITS - 1))) >> COEF_BITS);
        }
        return 1;
    }
    if (border_rule == ORACLE_BORDER_LEGACY) return 0;
    if (sx < 0 || sy < 0 || sx > sw - 1 || sy > sh - 1) return 0;

    /* A.4, OpenCV 4.13: on the last row / column the out-of-image taps are dropped and the
     * remaining weights are renormalised with one fp32 divide. */
    const int ok01 = sx + 1 < sw, ok10 = sy + 1 < sh, ok11 = ok01 && ok10;
    const int wtot = w00 + (ok01 ? w01 : 0) + (ok10 ? w10 : 0) + (ok11 ? w11 : 0);
    const uint8_t *p0 = src + (size_t)sy * spitch + (size_t)sx * 3;
    const uint8_t *p1 = p0 + spitch;
    for (int c = 0; c < 3; ++c) {
        int t0 = w00 * p0[c];
        if (ok01) t0 += w01 * p0[3 + c];
        if (ok10) t0 += w10 * p1[c];
        if (ok11) t0 += w11 * p1[3 + c];
        volatile float num = (float)t0 * 32768.0f;
        volatile float q = num / (float)wtot;
        int t = (int)q; /* truncation */
        out[c] = sat_u8((t + (1 << (COEF_BITS - 1))) >> COEF_BITS);
    }
    return 1;
}

/* cv::warpPerspective(..., BORDER_CONSTANT, borderValue = 0) for one canvas pixel: every pixel is
 * written, taps outside the source count as 0 (lib/Calibrate/utils.cpp:390). */
static inline void sample_constant(const uint8_t *src, int sw, int sh, size_t spitch, int X, int Y, uint8_t out[3])
{
    const int sx = sat_s16(X >> FRAC_BITS), sy = sat_s16(Y >> FRAC_BITS);
    const int a = X & (FRAC_ONE - 1), b = Y & (FRAC_ONE - 1);
    const int w[4] = {(FRAC_ONE - a) * (FRAC_ONE - b) * 32, a * (FRAC_ONE - b) * 32, (FRAC_ONE - a) * b * 32, a * b * 32};
    for (int c = 0; c < 3; ++c) {
        int v = 0;
        for (int t = 0; t < 4; ++t) {
            const int tx = sx + (t & 1), ty = sy + (t >> 1);
            if (tx >= 0 && tx < sw && ty >= 0 && ty < sh) v += w[t] * src[(size_t)ty * spitch + (size_t)tx * 3 + c];
        }
        out[c] = (uint8_t)((v + (1 << (COEF_BITS - 1))) >> COEF_BITS);
    }
}

void oracle_warp_perspective(const uint8_t *src, int sw, int sh, size_t spitch, uint8_t *dst,
                             int dw, int dh, size_t dpitch, const double H[9], int border_rule,
                             int nthreads)
{
    double M[9];
    oracle_invert3x3(H, M);
    const int bw = block_width(dw, dh);
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int y = 0; y < dh; ++y) {
        uint8_t *drow = dst + (size_t)y * dpitch;
        for (int x = 0; x < dw; ++x) {
            int X, Y;
            int xb = (x / bw) * bw;
            coord_rel(M, xb, x - xb, y, &X, &Y);
            uint8_t px[3];
            if (border_rule == ORACLE_BORDER_CONSTANT) {
                sample_constant(src, sw, sh, spitch, X, Y, px);
                drow[3 * x + 0] = px[0];
                drow[3 * x + 1] = px[1];
                drow[3 * x + 2] = px[2];
            } else if (sample_transparent(src, sw, sh, spitch, X, Y, border_rule, px)) {
                drow[3 * x + 0] = px[0];
                drow[3 * x + 1] = px[1];
                drow[3 * x + 2] = px[2];
            }
        }
    }
}

void oracle_stitch(int n_cams, const uint8_t *const *srcs, const int *sw, const int *sh,
                   const size_t *spitch, uint8_t *dst, int dw, int dh, size_t dpitch,
                   const double *H, int border_rule, int nthreads)
{
    for (int y = 0; y < dh; ++y) memset(dst + (size_t)y * dpitch, 0, (size_t)dw * 3);
    for (int i = 0; i < n_cams; ++i)
        oracle_warp_perspective(srcs[i], sw[i], sh[i], spitch[i], dst, dw, dh, dpitch, H + 9 * i,
                                border_rule, nthreads);
}

/* BS_FLAG_FEATHER (include/b200_stitch.h).  PARITY UNPINNED: the reference has no blending
 * (tools/stitch/main.cpp:117,151 overwrite), so this function restates the product's own definition and cites
 * nothing in the reference.  Camera i contributes the pixel the OpenCV call would have written with the integer
 * weight d_i = max(1, 1 + distance of the fixed-point source coordinate to the nearest source edge, in 1/32 px);
 * each channel is (sum d_i P_i + (sum d_i) / 2) / sum d_i. */
void oracle_stitch_feather(int n_cams, const uint8_t *const *srcs, const int *sw, const int *sh,
                           const size_t *spitch, uint8_t *dst, int dw, int dh, size_t dpitch,
                           const double *H, int border_rule, int nthreads)
{
    double M[64][9];
    if (n_cams > 64) n_cams = 64;
    for (int i = 0; i < n_cams; ++i) oracle_invert3x3(H + 9 * i, M[i]);
    const int bw = block_width(dw, dh);
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int y = 0; y < dh; ++y) {
        uint8_t *drow = dst + (size_t)y * dpitch;
        for (int x = 0; x < dw; ++x) {
            const int xb = (x / bw) * bw;
            uint64_t acc[3] = {0, 0, 0}, wsum = 0;
            for (int i = 0; i < n_cams; ++i) {
                int X, Y;
                uint8_t px[3];
                coord_rel(M[i], xb, x - xb, y, &X, &Y);
                if (!sample_transparent(srcs[i], sw[i], sh[i], spitch[i], X, Y, border_rule, px)) continue;
                int64_t d = X;
                const int64_t dr = 32 * (int64_t)(sw[i] - 1) - X, db = 32 * (int64_t)(sh[i] - 1) - Y;
                if (dr < d) d = dr;
                if (Y < d) d = Y;
                if (db < d) d = db;
                const uint64_t w = (uint64_t)(d + 1 > 1 ? d + 1 : 1);
                for (int c = 0; c < 3; ++c) acc[c] += w * px[c];
                wsum += w;
            }
            for (int c = 0; c < 3; ++c) drow[3 * x + c] = wsum ? (uint8_t)((acc[c] + wsum / 2) / wsum) : 0;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * cv::undistort (tools/stitch/main.cpp:148-149), Appendix A.6.
 * OpenCV builds the map in row stripes of height min(max(1, 4096 / w), h); inside a stripe
 * the new camera matrix has its principal point shifted up by the stripe's first row, is
 * inverted (same 3x3 cofactor inverse as above) and the normalised coordinates are advanced
 * by repeated addition along the row.  The order of those additions is kept here because
 * it decides the last bit of the doubles.
 * ---------------------------------------------------------------------------------------- */
void oracle_undistort(const uint8_t *src, int w, int h, size_t spitch, uint8_t *dst,
                      size_t dpitch, const double K[9], const double *dist, int n_dist,
                      int nthreads)
{
    double k[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < n_dist && i < 8; ++i) k[i] = dist[i];
    const double k1 = k[0], k2 = k[1], p1 = k[2], p2 = k[3], k3 = k[4], k4 = k[5], k5 = k[6],
                 k6 = k[7];
    const double fx = K[0], fy = K[4], u0 = K[2], v0 = K[5];
    int stripe0 = 4096 / (w > 1 ? w : 1);
    if (stripe0 < 1) stripe0 = 1;
    if (stripe0 > h) stripe0 = h;
    const int n_stripes = (h + stripe0 - 1) / stripe0;
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int s = 0; s < n_stripes; ++s) {
        const int ys = s * stripe0;
        const int rows = (h - ys) < stripe0 ? (h - ys) : stripe0;
        double Ar[9], ir[9];
        memcpy(Ar, K, sizeof Ar);
        Ar[5] = v0 - ys;
        oracle_invert3x3(Ar, ir);
        for (int i = 0; i < rows; ++i) {
            double _x = i * ir[1] + ir[2], _y = i * ir[4] + ir[5], _w = i * ir[7] + ir[8];
            uint8_t *drow = dst + (size_t)(ys + i) * dpitch;
            for (int j = 0; j < w; ++j, _x += ir[0], _y += ir[3], _w += ir[6]) {
                double iw = 1.0 / _w, x = _x * iw, y = _y * iw;
                double x2 = x * x, y2 = y * y;
                double r2 = x2 + y2, _2xy = 2 * x * y;
                double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) /
                            (1 + ((k6 * r2 + k5) * r2 + k4) * r2);
                double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2);
                double yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy;
                double u = fx * xd + u0;
                double v = fy * yd + v0;
                int iu = round_half_even(clamp_int_range(u * FRAC_ONE));
                int iv = round_half_even(clamp_int_range(v * FRAC_ONE));
                int sx = (short)(iu >> FRAC_BITS), sy = (short)(iv >> FRAC_BITS);
                int a = iu & (FRAC_ONE - 1), b = iv & (FRAC_ONE - 1);
                int wgt[4] = {(FRAC_ONE - a) * (FRAC_ONE - b) * 32, a * (FRAC_ONE - b) * 32,
                              (FRAC_ONE - a) * b * 32, a * b * 32};
                for (int c = 0; c < 3; ++c) {
                    int acc = 0;
                    for (int t = 0; t < 4; ++t) {
                        int tx = sx + (t & 1), ty = sy + (t >> 1);
                        if (tx >= 0 && tx < w && ty >= 0 && ty < h)
                            acc += wgt[t] * src[(size_t)ty * spitch + (size_t)tx * 3 + c];
                    }
                    drow[3 * j + c] = (uint8_t)((acc + (1 << (COEF_BITS - 1))) >> COEF_BITS);
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Synthetic frames, SURVEY.md section 8(d).
 * ---------------------------------------------------------------------------------------- */
void oracle_gen_frame(uint8_t *dst, int w, int h, size_t pitch, int cam, int frame, int kind)
{
    const uint32_t salt = (uint32_t)(cam * 131 + frame) * 0x27D4EB2Fu;
    for (int y = 0; y < h; ++y) {
        uint8_t *row = dst + (size_t)y * pitch;
        for (int x = 0; x < w; ++x) {
            for (int c = 0; c < 3; ++c) {
                if (kind == 0) {
                    uint32_t v = ((uint32_t)x * 0x9E3779B1u) ^ ((uint32_t)y * 0x85EBCA77u) ^
                                 ((uint32_t)c * 0xC2B2AE3Du) ^ salt;
                    v ^= v >> 15;
                    v *= 0x2C1B3C6Du;
                    v ^= v >> 12;
                    v *= 0x297A2D39u;
                    v ^= v >> 15;
                    row[3 * x + c] = (uint8_t)(v >> 24);
                } else {
                    row[3 * x + c] = (uint8_t)((x + 2 * y + 64 * c + 7 * frame + 31 * cam) & 255);
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Algorithmic bytes, SURVEY.md section 8(d).
 * ---------------------------------------------------------------------------------------- */
uint64_t oracle_count_texels(int n_cams, const int *sw, const int *sh, int dw, int dh,
                             const double *H, int border_rule, uint64_t *texels_out,
                             uint64_t *wins_out)
{
    const int bw = block_width(dw, dh);
    double *M = (double *)malloc(sizeof(double) * 9 * (size_t)n_cams);
    uint8_t **mark = (uint8_t **)malloc(sizeof(uint8_t *) * (size_t)n_cams);
    for (int i = 0; i < n_cams; ++i) {
        oracle_invert3x3(H + 9 * i, M + 9 * i);
        mark[i] = (uint8_t *)calloc((size_t)sw[i] * sh[i], 1);
        if (wins_out) wins_out[i] = 0;
    }
    for (int y = 0; y < dh; ++y) {
        for (int x = 0; x < dw; ++x) {
            int xb = (x / bw) * bw;
            for (int i = n_cams - 1; i >= 0; --i) {
                int X, Y;
                coord_rel(M + 9 * i, xb, x - xb, y, &X, &Y);
                int sx = sat_s16(X >> FRAC_BITS), sy = sat_s16(Y >> FRAC_BITS);
                int w = sw[i], h = sh[i];
                int covered;
                if (border_rule == ORACLE_BORDER_LEGACY)
                    covered = sx >= 0 && sy >= 0 && sx < w - 1 && sy < h - 1;
                else
                    covered = sx >= 0 && sy >= 0 && sx <= w - 1 && sy <= h - 1;
                if (!covered) continue;
                if (wins_out) wins_out[i]++;
                for (int t = 0; t < 4; ++t) {
                    int tx = sx + (t & 1), ty = sy + (t >> 1);
                    if (tx < w && ty < h) mark[i][(size_t)ty * w + tx] = 1;
                }
                break;
            }
        }
    }
    uint64_t total = 0;
    for (int i = 0; i < n_cams; ++i) {
        uint64_t n = 0;
        size_t cnt = (size_t)sw[i] * sh[i];
        for (size_t j = 0; j < cnt; ++j) n += mark[i][j];
        if (texels_out) texels_out[i] = n;
        total += n;
        free(mark[i]);
    }
    free(mark);
    free(M);
    return total;
}
