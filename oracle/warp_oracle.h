/*
 * oracle/warp_oracle.h -- CPU restatement of the stitch hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the checker, not the product: only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it.  The product path
 * (image-stitching_b200/csrc) never links or calls anything in this directory.
 *
 * What it restates: the per-frame work of the reference's stitch tool,
 *   /root/reference/tools/stitch/main.cpp:113-121 (image mode) and :143-155 (video mode),
 * i.e. N x cv::warpPerspective(src, canvas, H[i], size, INTER_LINEAR, BORDER_TRANSPARENT)
 * into one shared canvas, plus the cv::undistort that precedes it in video mode (:148-149).
 * The arithmetic itself lives in OpenCV imgproc, a third-party dependency that is NOT vendored
 * in the reference (CMakeLists.txt:7 find_package(OpenCV REQUIRED), version unpinned; the
 * de-facto CI pin is Ubuntu xenial's 2.4.9.1, .travis.yml:6,23).  The restatement follows
 * OpenCV's published fixed-point algorithm (INTER_BITS = 5, INTER_REMAP_COEF_BITS = 15) as
 * specified in SURVEY.md Appendix A.
 *
 * Parity pin: the restatement is pinned against outputs of the same OpenCV entry point the
 * reference calls, run in the build container through the cv2 4.13.0 binding
 * (tests/golden/make_golden.py -> the .npz fixtures and known_answers.json under tests/golden), and against the
 * SHA-256 known answers recorded in SURVEY.md Appendix B.3.  The reference's own tests hold
 * no vector for this path.
 */
#ifndef WARP_ORACLE_H
#define WARP_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Border rule for pixels whose integer source coordinate lies on the last source row/column
 * (SURVEY.md Appendix A.4). */
#define ORACLE_BORDER_CV4 0    /* OpenCV 4.13: written, weights renormalised in fp32 */
#define ORACLE_BORDER_LEGACY 1 /* OpenCV 2.4/3.x: left untouched */
#define ORACLE_BORDER_CONSTANT 2 /* not BORDER_TRANSPARENT at all: BORDER_CONSTANT 0, every pixel written
                                  (oracle_warp_perspective only; lib/Calibrate/utils.cpp:390) */

/* 3x3 double inverse the way cv::warpPerspective obtains M = H^-1
 * (call sites tools/stitch/main.cpp:116,150).  Returns 1 on success, 0 if singular
 * (then M is all zeros, as cv::invert leaves it). */
int oracle_invert3x3(const double H[9], double M[9]);

/* Fixed-point source coordinate of canvas pixel (x, y) under inverse map M:
 * X, Y are the 1/32-pixel integers of Appendix A.1 (before the >>5 / &31 split). */
void oracle_warp_coord(const double M[9], int x, int y, int *X, int *Y);

/* One cv::warpPerspective(src, dst, H, dsize, INTER_LINEAR, BORDER_TRANSPARENT) call:
 * dst is read-modify-write, uncovered pixels keep their contents.
 * tools/stitch/main.cpp:116-117, :150-151.  Packed BGR u8, pitches in bytes. */
void oracle_warp_perspective(const uint8_t *src, int sw, int sh, size_t spitch,
                             uint8_t *dst, int dw, int dh, size_t dpitch,
                             const double H[9], int border_rule, int nthreads);

/* The whole per-frame loop (tools/stitch/main.cpp:144-152): canvas zeroed first
 * (the reference leaves it uninitialised; SURVEY.md fact 3 defines uncovered = 0),
 * then cameras 0..n-1 in order. */
void oracle_stitch(int n_cams, const uint8_t *const *srcs, const int *sw, const int *sh,
                   const size_t *spitch, uint8_t *dst, int dw, int dh, size_t dpitch,
                   const double *H /* n_cams x 9 */, int border_rule, int nthreads);

/* BS_FLAG_FEATHER: weighted mean over the cameras that write a pixel instead of last-camera-wins.
 * PARITY UNPINNED -- the reference does not blend; this restates include/b200_stitch.h's definition only and
 * is not part of the parity metric. */
void oracle_stitch_feather(int n_cams, const uint8_t *const *srcs, const int *sw, const int *sh,
                           const size_t *spitch, uint8_t *dst, int dw, int dh, size_t dpitch,
                           const double *H, int border_rule, int nthreads);

/* cv::undistort(src, dst, K, dist) with bilinear interpolation and BORDER_CONSTANT 0
 * (tools/stitch/main.cpp:148-149).  K = fx,0,cx,0,fy,cy,0,0,1
 * row-major; dist has n_dist (0, 4, 5 or 8) coefficients k1 k2 p1 p2 k3 k4 k5 k6. */
void oracle_undistort(const uint8_t *src, int w, int h, size_t spitch, uint8_t *dst,
                      size_t dpitch, const double K[9], const double *dist, int n_dist,
                      int nthreads);

/* Synthetic frame generators of SURVEY.md section 8(d). kind 0 = noise, 1 = smooth. */
void oracle_gen_frame(uint8_t *dst, int w, int h, size_t pitch, int cam, int frame, int kind);

/* Algorithmic source bytes of one frame: for each camera the number
 * of distinct source texels referenced by the canvas pixels that camera wins.
 * texels_out[n_cams], wins_out[n_cams]; returns sum(texels). */
uint64_t oracle_count_texels(int n_cams, const int *sw, const int *sh, int dw, int dh,
                             const double *H, int border_rule, uint64_t *texels_out,
                             uint64_t *wins_out);

#ifdef __cplusplus
}
#endif
#endif
