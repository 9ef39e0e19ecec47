/*
 * b200_stitch.h -- C ABI of the B200-native compositing engine for the stitch stage of
 * AlexeySachkov/image-stitching.
 *
 * The reference has no FFI or plugin interface for this path: its stitch tool calls
 * cv::warpPerspective inline from main() (/root/reference/tools/stitch/main.cpp:116-117 in
 * image mode, :150-151 in video mode, preceded by cv::undistort at :148-149).  The entry points
 * below are what a C/C++ (or cgo/JNI/ctypes) caller binds instead of those OpenCV calls; each
 * one cites the reference lines it replaces.  INTEGRATION.md shows the edit to stitch's main().
 *
 * Conventions
 *   - plain C linkage, no exceptions, no C++ or torch types in any signature;
 *   - images are packed 8-bit BGR (CV_8UC3), row pitch in bytes, as in a continuous cv::Mat;
 *   - homographies are 3x3 row-major doubles, FORWARD (source pixel -> canvas pixel), exactly
 *     as stored under "H" in stitch.conf.xml (tools/calibrate/main.cpp:497-522);
 *   - every function returns BS_OK (0) or a negative BS_ERR_* code; bs_last_error() gives text;
 *   - a handle is single-threaded: use one handle (and one host thread) per GPU;
 *   - output pixels no camera covers are 0 (the reference leaves them uninitialised,
 *     tools/stitch/main.cpp:108,144) unless BS_FLAG_KEEP_CANVAS is given;
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with
 *     BS_ERR_NO_DEVICE.
 */
#ifndef B200_STITCH_H
#define B200_STITCH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BS_OK 0
#define BS_ERR_INVALID_ARG (-1)
#define BS_ERR_CUDA (-2)
#define BS_ERR_NO_DEVICE (-3)
#define BS_ERR_UNSUPPORTED (-4)
#define BS_ERR_NOMEM (-5)
#define BS_ERR_BAD_TICKET (-6)

/* Border rule on the last source row/column: default is OpenCV 4.x (pixel written, weights
 * renormalised); LEGACY is OpenCV 2.4/3.x (pixel left untouched).  SURVEY.md Appendix A.4. */
#define BS_FLAG_LEGACY_BORDER 0x1u
/* Read-modify-write canvas, as one cv::warpPerspective(..., BORDER_TRANSPARENT) call into an
 * existing Mat (tools/stitch/main.cpp:116): uncovered pixels keep their previous contents. */
#define BS_FLAG_KEEP_CANVAS 0x2u
/* Testing aid: run every tile through the exact per-pixel kernel (no tile plan fast path). */
#define BS_FLAG_EXACT_ONLY 0x4u
/* cv::BORDER_CONSTANT with value 0 instead of BORDER_TRANSPARENT: every canvas pixel is written and
 * source taps outside the frame count as 0 -- the variant lib/Calibrate/utils.cpp:390
 * (projectToTheFloor) uses; SURVEY.md row f3.  Meant for bs_warp_perspective; in a handle the last
 * camera then overwrites the whole canvas, as N such OpenCV calls would.  Implies BS_FLAG_EXACT_ONLY. */
#define BS_FLAG_BORDER_CONSTANT 0x8u
/* OPT-IN, NOT the reference's result: where several cameras write one canvas pixel, store
 * their weighted mean instead of the last camera's pixel.  Camera i contributes the pixel P_i the OpenCV call
 * would have written, with the integer weight d_i = max(1, 1 + min(X, 32 (w_i - 1) - X, Y, 32 (h_i - 1) - Y)),
 * its distance to the nearest source edge in 1/32 pixels (X, Y the fixed-point source coordinate of SURVEY.md
 * Appendix A.1); each channel is (sum d_i P_i + (sum d_i) / 2) / sum d_i in integers.  Pixels one camera
 * writes are unchanged, so are uncovered pixels.  The reference has no blending at all
 * (tools/stitch/main.cpp:117,151 overwrite), so there is nothing to be bit-exact with: the CPU restatement
 * oracle_stitch_feather is "parity unpinned" and is excluded from the parity metric.  Blocks of the canvas that
 * two cameras write into are blended by the staged kernel from the weights of their two layers (about four times
 * the cost of a plain block); blocks with three or more cameras on a pixel, with last-row/column pixels or with
 * rows that need exact coordinates take the per-pixel path; at most 16 cameras; not with
 * BS_FLAG_BORDER_CONSTANT; handles only (the one-shot calls and bs_stitch_device_static refuse it). */
#define BS_FLAG_FEATHER 0x10u

#define BS_MAX_CAMERAS 64
#define BS_MAX_DIM 32767 /* OpenCV keeps integer source coordinates in int16 */

typedef struct bs_handle bs_handle;

typedef struct bs_stats {
    int n_cams, canvas_w, canvas_h;
    int tile_w, tile_h;
    int64_t n_tiles;          /* canvas blocks (tile_w x tile_h pixels) in the plan */
    int64_t n_tiles_fast;     /* blocks on the staged path (interior samples of <= 2 cameras) */
    int64_t n_tiles_exact;    /* blocks on the exact per-pixel path (last-row/column pixels, 3+ cameras) */
    int64_t n_tiles_empty;    /* blocks no camera touches */
    int64_t kernel_launches;  /* kernels of this library launched through this handle */
    int64_t frames_submitted;
    uint64_t device_bytes;    /* device memory owned by the handle */
    uint64_t scratch_bytes;   /* of which plan/scratch */
    uint64_t h2d_bytes_last_submit; /* bytes the last bs_submit copied to the device: only the part
                                       of every frame the plan can read is uploaded */
    int frame_kernel;         /* 0: exact per-pixel kernel only, 1: general staged kernel, 2: compact-box staged kernel */
} bs_stats;

/* Number of CUDA devices visible; 0 if none (never negative). */
int bs_device_count(void);

/* Library version string, e.g. "b200-stitch 0.1 (sm_100a)". */
const char *bs_version(void);

/* Last error text of a handle, or of the calling thread when h == NULL. Never NULL. */
const char *bs_last_error(const bs_handle *h);

/*
 * Set-up, once per calibration.  Replaces the state the reference builds at
 * tools/stitch/main.cpp:53-73 (result_size, H[]) and :108 / :144 (canvas allocation):
 * inverts every H in double the way cv::warpPerspective does, builds the tile plan on the
 * device, allocates device frame/canvas rings and the copy/compute streams.
 *   src_w/src_h[n_cams], H[n_cams*9]; flags = BS_FLAG_*.
 */
int bs_create(int device_id, int n_cams, const int *src_w, const int *src_h, int canvas_w,
              int canvas_h, const double *H, unsigned flags, bs_handle **out);

int bs_destroy(bs_handle *h);

/*
 * One panorama frame from HOST buffers, asynchronous.  Replaces the body of the camera loop,
 * tools/stitch/main.cpp:145-152 (video mode) / :113-121 (image mode): N warps into one canvas,
 * highest camera index wins in overlaps.  frames[i]/pitch[i] are packed BGR rows of camera i;
 * `out` receives canvas_h rows of canvas_w*3 bytes at out_pitch.  All host pointers must stay
 * valid until bs_wait(ticket).  Only the region of each frame that the calibration can map into the
 * canvas is copied to the device.  Pinned buffers (bs_host_alloc) are copied by DMA directly;
 * pageable ones work but copy slower.  Up to bs_pipeline_depth() frames may be in flight.
 */
int bs_submit(bs_handle *h, const uint8_t *const *frames, const size_t *pitch, uint8_t *out,
              size_t out_pitch, int64_t *ticket);
int bs_wait(bs_handle *h, int64_t ticket);
int bs_pipeline_depth(const bs_handle *h);

/* bs_submit + bs_wait. */
int bs_stitch(bs_handle *h, const uint8_t *const *frames, const size_t *pitch, uint8_t *out,
              size_t out_pitch);

/*
 * Same work on DEVICE-resident buffers (kernel only, no copies), asynchronous on `stream`
 * (a cudaStream_t; NULL = the legacy default stream, as everywhere in CUDA).  d_frames[i] and d_canvas are device
 * pointers; for the staged fast path they must be 16-byte aligned with pitches that are
 * multiples of 16, otherwise the exact path is used for the whole frame.
 * Replaces tools/stitch/main.cpp:145-152 for callers that decode straight into GPU memory.
 *
 * Ordering: bs_submit_device is an ordinary stream-ordered call -- whatever precedes it in `stream`
 * (copies, the caller's decode kernels, this library's own kernels) has completed, and is visible,
 * before the first source texel is read.
 *
 * bs_submit_device_ex adds submit flags.  BS_SUBMIT_SOURCES_SETTLED: the caller states that no
 * kernel directly before this call in `stream` writes the source frames (they were produced earlier,
 * on another stream ordered by an event, or by copies).  The frame kernel is launched with programmatic
 * stream serialisation in both forms; with this flag it may also prefetch and blend its first blocks while
 * the previous kernel in the stream drains, so back-to-back frames overlap; it still touches the canvas
 * only after that kernel has completed.  Without the flag every source read waits for the previous
 * kernel (griddepcontrol.wait).  A handle with lens models undistorts with kernels of its own in
 * `stream` first and ignores the flag; its intermediate frames are one set per handle, so keep such a
 * handle's device submits on ONE stream.  The caller's current device is not changed.
 */
#define BS_SUBMIT_SOURCES_SETTLED 0x1u
int bs_submit_device(bs_handle *h, const void *const *d_frames, const size_t *pitch,
                     void *d_canvas, size_t canvas_pitch, void *stream);
int bs_submit_device_ex(bs_handle *h, const void *const *d_frames, const size_t *pitch,
                        void *d_canvas, size_t canvas_pitch, void *stream, unsigned submit_flags);

/* Block until everything submitted through the handle has finished. */
int bs_sync(bs_handle *h);

/*
 * Drop-in for ONE call
 *   cv::warpPerspective(src, dst, H, Size(dw, dh), INTER_LINEAR, BORDER_TRANSPARENT)
 * (tools/stitch/main.cpp:116-117, :150-151; also tools/calibrate/main.cpp:484-485) on host
 * buffers: synchronous, dst is read-modify-write.  flags: BS_FLAG_LEGACY_BORDER.  Uses device
 * 0 (or the device of the last bs_create on this thread).  Convenience only: it uploads dst,
 * so the per-frame loop should use bs_create + bs_submit instead.
 */
int bs_warp_perspective(const uint8_t *src, int sw, int sh, size_t spitch, uint8_t *dst, int dw,
                        int dh, size_t dpitch, const double H[9], unsigned flags);

/*
 * Lens model of one camera: from then on every frame of that camera is undistorted on the GPU
 * before it is warped, replacing cv::undistort(frame, undistorted, cameraMatrix[i], distCoeffs[i])
 * at tools/stitch/main.cpp:148-149 (same fixed-point map and bilinear resampling, BORDER_CONSTANT 0,
 * SURVEY.md Appendix A.6).  K = fx 0 cx 0 fy cy 0 0 1 (row-major); dist = k1 k2 p1 p2 [k3 [k4 k5 k6]],
 * n_dist in 0..8 (calibrate writes 8, tools/calibrate/main.cpp:173-178).  A model whose
 * coefficients are all zero (or K == NULL) disables the step: cv::undistort is then an exact copy.
 * The undistorted frame is internal to the handle: inside bs_submit* the lens models of all cameras run as one
 * launch, and only the part of each undistorted frame that the panorama can read is computed (and, from host
 * buffers, only the part of the distorted frame that this needs is uploaded) -- the panorama is the same bit for
 * bit; bs_undistort_device below produces whole frames.
 */
int bs_set_undistort(bs_handle *h, int cam, const double K[9], const double *dist, int n_dist);

/* cv::undistort of one frame of camera `cam` with the model set by bs_set_undistort, between DEVICE
 * buffers (packed BGR, d_src != d_dst), asynchronous on `stream`: tools/stitch/main.cpp:148-149 for callers
 * whose frames are already on the GPU.  A camera without a model is copied.  submit_flags as for
 * bs_submit_device_ex (BS_SUBMIT_SOURCES_SETTLED: no kernel directly before this call writes d_src). */
int bs_undistort_device(bs_handle *h, int cam, const void *d_src, size_t spitch, void *d_dst, size_t dpitch,
                        void *stream, unsigned submit_flags);

/* Drop-in for ONE cv::undistort(src, dst, K, dist) call on host buffers (main.cpp:148-149). */
int bs_undistort(const uint8_t *src, int w, int h, size_t spitch, uint8_t *dst, size_t dpitch,
                 const double K[9], const double *dist, int n_dist);

/* Pinned host memory for frames and panoramas (cudaHostAlloc / cudaFreeHost). */
void *bs_host_alloc(size_t bytes);
void bs_host_free(void *p);

int bs_get_stats(const bs_handle *h, bs_stats *out);

/*
 * Compressed frames in, and optionally a compressed panorama out: the device-side form of
 * imread / VideoCapture at tools/stitch/main.cpp:107,114,127-147 and of imwrite at :125.  jpeg[i] / jpeg_bytes[i]
 * are the JPEG bitstreams of the cameras' frames (baseline JPEG, the handle's frame sizes); they are decoded
 * on the GPU (nvJPEG) straight into the handle's frame ring, so no decoded frame crosses PCIe.  The panorama
 * is returned raw in `out` (may be NULL), and/or as a JPEG in `jout` (may be NULL): jout->data / capacity name
 * the caller's buffer, quality is 1..100 (0 = 95, cv::imwrite's default; 4:2:0 chroma as there), and jout->size
 * holds the bitstream length after bs_wait(ticket) -- only the bitstream is copied from the device.  All
 * buffers (the bitstreams included), and *jout itself, must stay valid until bs_wait.  Decoded pixels are
 * nvJPEG's; they can differ by a unit or two from libjpeg's, as between any two JPEG decoders.
 * Baseline JPEGs with three components sampled 4:2:0 or 4:4:4 (with or without restart markers) are decoded by the
 * library's own GPU decoder -- entropy decoding on the device in self-synchronising chunks, libjpeg's arithmetic:
 * the pixels equal cv::imread's (libjpeg-turbo) bit for bit; all other JPEGs go through nvJPEG.
 * bs_submit_jpeg checks the headers (a JPEG of the camera's frame size) and returns; the entropy (Huffman) stages
 * of the cameras then run on worker threads of the handle and the GPU side is issued when they are done, by the
 * next call that needs the frame (bs_wait of the ticket, a later submit, bs_sync) -- so the host stages of up to
 * bs_pipeline_depth() frames overlap.  A bitstream that turns out to be undecodable is therefore reported by
 * bs_wait of its ticket (or by the next submit that needs its slot), not by bs_submit_jpeg.
 */
typedef struct bs_jpeg_out {
    uint8_t *data;
    size_t capacity;
    size_t size;       /* out: bytes written (or needed, when bs_wait reports that capacity was too small) */
    int quality;
} bs_jpeg_out;
int bs_submit_jpeg(bs_handle *h, const uint8_t *const *jpeg, const size_t *jpeg_bytes, uint8_t *out, size_t out_pitch,
                   bs_jpeg_out *jout, int64_t *ticket);
/* Which decoder the handle's last bs_submit_jpeg used: "b200 ..." (the library's GPU decoder), or nvJPEG's "hardware" (the GPU's JPEG engine, batched),
 * "gpu_hybrid" (GPU-assisted Huffman decoding, batched), "default" (per image, Huffman stage on the host);
 * "" before the first bs_submit_jpeg. */
const char *bs_jpeg_backend(const bs_handle *h);

/*
 * Frame-parallel dispatch over several GPUs of one box.  Replaces the single loop of
 * tools/stitch/main.cpp:143-155 for video: every output frame is independent (the canvas is re-created at
 * :144), so frame f runs on GPU f mod G with a full replica of the calibration; no data is exchanged between
 * GPUs.  A pool owns one handle and one host thread per device (device_ids[n_devices], NULL = 0..n-1; an id may
 * repeat).  Submit frames in order from ONE thread; host buffers must stay valid until bs_pool_wait(frame).
 * bs_pool_submit blocks while bs_pool_depth() frames are outstanding, so wait for old frames as you go:
 * waiting for frames 0, 1, 2, ... in order yields the panoramas in frame order.  bs_pool_wait(f) returns once
 * frame f is in `out`; after an error every later call returns that error (bs_pool_last_error has the text).
 * Lens models (bs_pool_set_undistort) go to every replica and must be set before the first frame.
 */
typedef struct bs_pool bs_pool;
int bs_pool_create(int n_devices, const int *device_ids, int n_cams, const int *src_w, const int *src_h,
                   int canvas_w, int canvas_h, const double *H, unsigned flags, bs_pool **out);
int bs_pool_destroy(bs_pool *p);
int bs_pool_size(const bs_pool *p);                 /* number of replicas */
int bs_pool_depth(const bs_pool *p);                /* frames that may be outstanding: replicas x bs_pipeline_depth */
bs_handle *bs_pool_handle(bs_pool *p, int i);       /* replica i (for bs_get_stats) */
const char *bs_pool_last_error(const bs_pool *p);
int bs_pool_set_undistort(bs_pool *p, int cam, const double K[9], const double *dist, int n_dist);
int bs_pool_submit(bs_pool *p, const uint8_t *const *frames, const size_t *pitch, uint8_t *out, size_t out_pitch,
                   int64_t *frame_index);
int bs_pool_wait(bs_pool *p, int64_t frame_index);
int bs_pool_sync(bs_pool *p);                       /* every submitted frame is done */

/*
 * One panorama frame on DEVICE buffers without allocating anything.  No counterpart in the
 * reference (it has no GPU path and no allocation discipline); added for callers that own all their
 * memory and must not allocate while running -- frame loops with a fixed device-memory budget,
 * hosts that forbid cudaMalloc after start-up (it also synchronises the device), plug-ins living
 * inside someone else's allocator.  Same result as bs_create + bs_submit_device for the same
 * arguments (flags: BS_FLAG_LEGACY_BORDER, BS_FLAG_KEEP_CANVAS); n_cams <= 16.  The tile plan stays in the
 * scratch between calls: a call whose n_cams, sizes, H and flags equal the previous call's on the same device (a
 * video loop) only launches the frame kernel; any other call rebuilds the plan first (about 2 ms at 7680x2160).
 *
 * The staged kernel's scratch (the block plan and the per-warp work split) lives in one static device
 * array of BS_STATIC_SCRATCH_BYTES that every GPU loading the library keeps until exit.  Canvases
 * needing more than BS_STATIC_MAX_BLOCKS blocks of 32x16 pixels (ceil(w/32) * ceil(h/16) > 32768; a
 * 7680x2160 panorama needs 32400) are refused with BS_ERR_UNSUPPORTED.  All calls share that array:
 * calls whose GPU work overlaps (different streams) race on it, so keep them on one stream.  The
 * first call on each device performs the one-time address lookup and must come before any stream
 * capture.  The call returns after its kernels have finished (it synchronises `stream` at the end, and
 * once more for the plan when it rebuilds it).
 */
#define BS_STATIC_MAX_BLOCKS 32768
#define BS_STATIC_SCRATCH_BYTES (BS_STATIC_MAX_BLOCKS * 160 + 32768)
int bs_stitch_device_static(int n_cams, const int *src_w, const int *src_h, const void *const *d_frames,
                            const size_t *pitch, int canvas_w, int canvas_h, void *d_canvas,
                            size_t canvas_pitch, const double *H, unsigned flags, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_STITCH_H */
