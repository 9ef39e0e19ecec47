// stitch_api.cu -- C ABI (include/b200_stitch.h) over the sm_100a kernels.
//
// Host-side responsibilities, all replacing set-up the reference does inline in
// /root/reference/tools/stitch/main.cpp: invert the homographies (what cv::warpPerspective does
// on entry, :116), own the canvas (:108, :144), and run the per-frame camera loop (:145-152) as
// kernel launches fed by copy streams.
#include "../../include/b200_stitch.h"

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>
#include <chrono>
#include <thread>

#include <algorithm>

#include "stitch_fast2.cuh"
#include "stitch_undistort.cuh"

namespace {

// Error text lives in fixed buffers (per thread, and per handle): reporting an error allocates nothing.
constexpr size_t kErrLen = 512;
thread_local char g_last_error[kErrLen] = "";
thread_local int g_last_device = 0;

#ifndef BS_PIPELINE_DEPTH
#define BS_PIPELINE_DEPTH 3
#endif
constexpr int kPipelineDepth = BS_PIPELINE_DEPTH;

struct Slot {
    std::vector<uint8_t *> d_frame;
    std::vector<uint8_t *> d_undist;     // per camera with a lens model: the undistorted frame the warp reads
    uint8_t *d_canvas = nullptr;
    cudaEvent_t ev_h2d = nullptr, ev_kernel = nullptr, ev_done = nullptr;
    int64_t ticket = -1;
    bool busy = false;
    bs_jpeg_out *jout = nullptr;          // the frame in this slot leaves as a JPEG (bs_submit_jpeg): filled by bs_wait
    // bs_submit_jpeg: the frame's host decode stages are still running, its GPU side has not been issued yet
    bool deferred = false;
    int deferred_rc = 0;                  // ... and what issuing it later returned (reported by bs_wait of this ticket)
    char deferred_err[256] = "";
};

// Lens model of one camera (cv::undistort, tools/stitch/main.cpp:148-149).
// One encoded TMA descriptor of a camera's frame buffer (the box shape is fixed per handle).
struct TmapEntry {
    const void *ptr = nullptr;
    size_t pitch = 0;
    int box_h = 0;                        // lens plans: rows of the box the map was encoded with
    CUtensorMap map;
};
constexpr int kTmapCacheWays = 8;

struct JpegCtx;                           // nvJPEG state of a handle (stitch_jpeg.inl), created by the first bs_submit_jpeg

struct LensModel {
    bool enabled = false;                // false: no model, or all coefficients zero (undistort is then a copy)
    double fx = 0, fy = 0, u0 = 0, v0 = 0, winv = 1;
    double k[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // k1 k2 p1 p2 k3 k4 k5 k6
    double *d_xs = nullptr, *d_ys = nullptr;  // device tables, see stitch_undistort.cuh
    uint32_t *d_map = nullptr;                // cached fixed-point map (w * h words), nullptr: evaluate per frame
    uint64_t bytes = 0;                       // device memory behind the pointers (counted in bs_handle::device_bytes)
    // block plan of the lens field for the compact frame kernel (stitch_undistort.cuh: lens_plan_kernel)
    bool plan_ok = false;
    bs::WorkItem *d_items = nullptr;
    bs::CubicCoef *d_cubic = nullptr;         // third-order terms, one per work item
    int *d_warp_start = nullptr;
    int grid = 0, smem_bytes = 0, box_h = 0;
    TmapEntry tmaps[kTmapCacheWays];          // source frames come from small rings
    unsigned tmap_next = 0;
    // what the combined launch over several cameras is built from (LensSet): the plan before it was dealt to warps
    std::vector<bs::WorkItem> items_host;     // two-layer blocks first (a lens plan has none), then one item per block
    std::vector<bs::CubicCoef> cubic_blocks;  // per block, row-major
    int n_items = 0, n_two = 0, blocks_x = 0;
};

// The lens plans of all cameras of a handle dealt to the warps of ONE launch (a 4K frame alone gives a warp only
// 3.4 blocks: four frames in one launch amortise the per-warp set-up and fill the tail).  Rebuilt whenever a
// camera's model changes; valid with two or more plans.
struct LensSet {
    bool valid = false;
    bool trimmed = false;                     // lens blocks the panorama plan never reads were left out
    std::vector<int> cams;
    std::vector<int> src_x0, src_y0, src_x1, src_y1;   // per camera: the distorted-frame rectangle the launch reads (-1: all of it)
    bs::WorkItem *d_items = nullptr;
    bs::CubicCoef *d_cubic = nullptr;
    int *d_warp_start = nullptr;
    int grid = 0, smem_bytes = 0, box_h = 0;
    uint64_t bytes = 0;
};

}  // namespace

struct bs_handle {
    int device = 0;
    int n_cams = 0;
    std::vector<int> sw, sh;
    int W = 0, H = 0;
    unsigned flags = 0;
    std::vector<double> M;               // n_cams x 9, inverse homographies
    std::vector<size_t> d_frame_pitch;   // device pitch of each camera's frame
    size_t d_canvas_pitch = 0;
    cudaStream_t s_h2d = nullptr, s_compute = nullptr, s_d2h = nullptr;
    std::vector<Slot> ring;
    int64_t next_ticket = 0;
    int64_t kernel_launches = 0;
    std::vector<LensModel> lens;         // per camera
    LensSet lens_set;                    // their plans as one launch
    bs::UndistortArgs *d_lens_models = nullptr;   // n_cams entries: model constants the lens launches read (rare paths)
    std::vector<int> up_x0, up_y0, up_x1, up_y1;   // per camera: source texels the plan can touch (inclusive); empty = whole frame
    std::vector<uint8_t *> d_undist;     // undistorted frames for bs_submit_device (one set, stream-ordered)
    // tile plan (static per calibration)
    bool fast_ok = false;                // the staged kernel can be used
    bool compact = false;                // ... and it is the compact-box kernel (stitch_fast2.cuh)
    bs::WorkItem *d_items = nullptr;     // plan: work items grouped by warp (same allocation held the raw records)
    int *d_warp_start = nullptr;         // item range of every warp of the grid
    int n_blocks = 0, blocks_x = 0, blocks_y = 0, n_items = 0;
    int64_t n_fast = 0, n_exact = 0, n_empty = 0;
    int box_we[bs::kMaxCamsPerLaunch] = {0}, box_h[bs::kMaxCamsPerLaunch] = {0};
    int stage_bytes = 0, smem_bytes = 0, grid = 0;
    int remap_grid = 0;                  // persistent CTAs of undistort_remap_kernel (set with the first lens map)
    int rec_slots = bs::kRecSlotsMax;    // record ring slots per warp of the frame kernel
    std::vector<TmapEntry> tmap_cache;   // n_cams x kTmapCacheWays, filled round-robin per camera
    unsigned tmap_next[bs::kMaxCamsPerLaunch] = {0};
    uint64_t device_bytes = 0, scratch_bytes = 0;
    uint64_t h2d_bytes_last = 0;         // bytes the last bs_submit uploaded
    JpegCtx *jpeg = nullptr;
    char err[512] = "";
};

namespace {

int fail(bs_handle *h, int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, kErrLen, fmt, ap);
    va_end(ap);
    if (h) memcpy(h->err, g_last_error, kErrLen);
    return code;
}

#define BS_CUDA(h, call)                                                                      \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return fail(h, BS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                        __FILE__, __LINE__);                                                  \
    } while (0)

#define BS_CUDA_QUIET(call) do { if ((call) != cudaSuccess) cudaGetLastError(); } while (0)

// Host -> device copy that is complete when it returns.  (A plain cudaMemcpy from pageable memory
// may return while the DMA out of the driver's staging buffer is still running, and the handle's
// streams are non-blocking, so a kernel on them would not wait for it.)
cudaError_t upload_now(bs_handle *h, void *dst, const void *src, size_t bytes)
{
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->s_compute);
    return e == cudaSuccess ? cudaStreamSynchronize(h->s_compute) : e;
}

// 3x3 inverse as cv::invert computes it for a 3x3 CV_64F matrix: cofactors times 1/det.
// (cv::warpPerspective inverts H on entry because the reference passes the forward map,
// tools/stitch/main.cpp:116.)  Singular input leaves M = 0, as OpenCV does.
void invert3x3(const double *H, double *M)
{
    const double a = H[0], b = H[1], c = H[2], d = H[3], e = H[4], f = H[5], g = H[6], hh = H[7],
                 i = H[8];
    double det = a * (e * i - f * hh) - b * (d * i - f * g) + c * (d * hh - e * g);
    if (det == 0.0) {
        for (int k = 0; k < 9; ++k) M[k] = 0.0;
        return;
    }
    det = 1.0 / det;
    M[0] = (e * i - f * hh) * det;
    M[1] = (c * hh - b * i) * det;
    M[2] = (b * f - c * e) * det;
    M[3] = (f * g - d * i) * det;
    M[4] = (a * i - c * g) * det;
    M[5] = (c * d - a * f) * det;
    M[6] = (d * hh - e * g) * det;
    M[7] = (b * g - a * hh) * det;
    M[8] = (a * e - b * d) * det;
}

// Column granularity of OpenCV's coordinate blocks (32*32 pixels, at most 16 rows).
int opencv_block_width(int W, int H)
{
    int bh = H < 16 ? H : 16;
    if (bh < 1) bh = 1;
    int bw = 1024 / bh;
    if (bw > W) bw = W;
    if (bw < 1) bw = 1;
    return bw;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int require_device(bs_handle *h)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        return fail(h, BS_ERR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU path",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    return BS_OK;
}


// ---- TMA descriptors ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encoder()
{
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// Source frame as a 2-D tensor of 4-byte elements: [h rows][pitch / 4 elements].
int encode_frame_map(CUtensorMap *m, const void *ptr, size_t pitch, int h, int box_we, int box_h)
{
    EncodeTiledFn enc = get_encoder();
    if (!enc) return -1;
    cuuint64_t gdim[2] = {(cuuint64_t)(pitch / 4), (cuuint64_t)h};
    cuuint64_t gstride[1] = {(cuuint64_t)pitch};
    cuuint32_t box[2] = {(cuuint32_t)box_we, (cuuint32_t)box_h};
    cuuint32_t estr[2] = {1, 1};
    return (int)enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(ptr), gdim, gstride, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

constexpr int kStageBytesMax = 6 * 1024;    // per-warp, per-stage cap on a block's source box

// Dynamic shared memory of one CTA of the frame kernel (+ alignment slack).
int cta_smem_bytes(int stage_bytes, int rec_slots) { return bs::kWarpsPerCta * bs::warp_smem_bytes(stage_bytes, rec_slots) + 128; }

// Result of turning the raw block records into what the frame kernel consumes.
struct PlanLayout {
    int box_we[bs::kMaxCamsPerLaunch] = {0}, box_h[bs::kMaxCamsPerLaunch] = {0};
    int stage_bytes = 0, smem_bytes = 0;
    int n_items = 0, n_two = 0;         // n_two: blocks with two camera layers (they come first)
    int64_t n_fast = 0, n_exact = 0, n_empty = 0;
    bool compact = false;               // every staged box fits the compact kernel (stitch_fast2.cuh): 128 bytes x 18 rows
};

// Host post-processing of the plan, on caller-provided storage (no allocation here):
//   recs[n_blocks]        raw records from plan_kernel (modified: oversized boxes are demoted to EXACT)
//   items[>= 2*n_blocks]  output, heaviest blocks first, a block's layers adjacent
// Sizes one TMA box per camera (the largest source box among its blocks).
// First converted texel column of a box whose leftmost texel is x0 (compact kernel: groups of four texels
// start on word boundaries of the packed rows), and the 16-byte aligned element the TMA box starts at.
inline int compact_t0(int x0) { return x0 & ~3; }
inline int compact_c0(int x0) { return bs::box_elem0(compact_t0(x0)); }

void layout_plan(bs::BlockRec *recs, int n_blocks, int n_cams, bs::WorkItem *items, PlanLayout *L, bool allow_compact, bool feather = false)
{
    int we[bs::kMaxCamsPerLaunch] = {0}, bh[bs::kMaxCamsPerLaunch] = {0};
    // compact: every staged layer's texels fit 40 converted columns, 128 staged bytes per row and 18 rows
    bool compact = allow_compact;
    for (int t = 0; compact && t < n_blocks; ++t) {
        const bs::BlockRec &r = recs[t];
        if (r.type != bs::kBlockFast) continue;
        for (int l = 0; l < r.n_layers; ++l) {
            const int x0 = r.layer[l].x0, x1 = r.layer[l].x1;
            if (x0 < 0 || x1 - compact_t0(x0) >= bs::kConvTexels || 3 * (x1 + 1) - 4 * compact_c0(x0) > bs::kRawPitch ||
                r.layer[l].y1 - r.layer[l].y0 + 1 > bs::kRawRows)
                compact = false;
        }
    }
    L->compact = compact;
    for (int t = 0; t < n_blocks; ++t) {
        bs::BlockRec &r = recs[t];
        if (r.type != bs::kBlockFast) continue;
        // BS_FLAG_FEATHER: two-layer blocks are blended by the compact kernel's run_feather2, which takes the cheap
        // coordinates only; anything else (general kernel, flagged rows) is blended per pixel
        if (feather && r.n_layers == 2 && (!compact || r.layer[0].groups != 0 || r.layer[1].groups != 0)) {
            r.type = bs::kBlockExact;
            continue;
        }
        int twe[bs::kMaxLayers], tbh[bs::kMaxLayers];
        bool ok = true;
        for (int l = 0; l < r.n_layers; ++l) {
            const int e0 = bs::box_elem0(r.layer[l].x0), e1 = (3 * r.layer[l].x1 + 2) >> 2;   // box starts 16-byte aligned
            twe[l] = compact ? bs::kRawPitch / 4 : (e1 - e0 + 1 + 3) & ~3;                     // inner extent: multiple of 16 bytes
            tbh[l] = r.layer[l].y1 - r.layer[l].y0 + 1;
            if (twe[l] > 256 || tbh[l] > 256 || twe[l] * 4 * tbh[l] > kStageBytesMax) ok = false;
        }
        if (!ok) {
            r.type = bs::kBlockExact;
            continue;
        }
        for (int l = 0; l < r.n_layers; ++l) {
            const int c = r.layer[l].cam;
            we[c] = std::max(we[c], twe[l]);
            bh[c] = std::max(bh[c], tbh[l]);
        }
    }
    // the per-camera maxima of width and height come from different blocks; re-check the product
    int stage = 0;
    for (int c = 0; c < n_cams; ++c) {
        if (we[c] * 4 * bh[c] > kStageBytesMax) {
            for (int t = 0; t < n_blocks; ++t) {
                bs::BlockRec &r = recs[t];
                if (r.type != bs::kBlockFast) continue;
                for (int l = 0; l < r.n_layers; ++l)
                    if (r.layer[l].cam == c) r.type = bs::kBlockExact;
            }
            we[c] = bh[c] = 0;
        }
        L->box_we[c] = we[c];
        L->box_h[c] = bh[c];
        stage = std::max(stage, we[c] * 4 * bh[c]);
    }
    if (compact) {      // one box shape for every camera: the frame kernel then expects one byte count
        int rows = 0;
        for (int c = 0; c < n_cams; ++c) rows = std::max(rows, L->box_h[c]);
        for (int c = 0; c < n_cams; ++c) L->box_h[c] = rows;
    }
    L->stage_bytes = (int)align_up((size_t)stage + bs::kStagePad, 128);
    L->smem_bytes = cta_smem_bytes(L->stage_bytes, bs::kRecSlotsMax);
    L->n_fast = L->n_exact = L->n_empty = 0;
    int n = 0;
    auto emit = [&](const bs::BlockRec &r, int l) {
        bs::WorkItem &it = items[n++];
        memset(&it, 0, sizeof it);
        it.bx = r.bx;
        it.by = r.by;
        it.type = r.type;
        const int nl = (r.type == bs::kBlockFast) ? r.n_layers : 1;
        it.flags = (uint8_t)((l == 0 ? bs::kItemFirst : 0) | (l == nl - 1 ? bs::kItemLast : 0));
        if (r.type == bs::kBlockFast) {
            it.flags |= r.layer[l].full ? bs::kItemFull : 0;
            it.cam = r.layer[l].cam;
            it.groups = r.layer[l].groups;
            it.c0 = (int16_t)(L->compact ? compact_c0(r.layer[l].x0) : bs::box_elem0(r.layer[l].x0));
            it.y0 = r.layer[l].y0;
            it.origin_off = (uint32_t)((int)it.y0 * (L->box_we[it.cam] * 4) + 4 * (int)it.c0);
            it.coef = r.coef[l];
            if (L->compact) {
                // compact kernel (stitch_fast2.cuh): origin_off holds the converted-buffer offset of lane u = 0 /
                // row v = -8 (whole-texel parts), aux the byte steps per canvas column and row, flags bits 4-5
                // where the first converted texel sits in the staged row
                const int t0 = compact_t0(r.layer[l].x0);
                const bs::BlockCoef &c = r.coef[l];
                const int ustep = c.nyu * bs::kConvPitch + c.nxu * 4, vstep = c.nyv * bs::kConvPitch + c.nxv * 4;
                const long long run_rel = -((long long)it.y0 * bs::kConvPitch + 4LL * t0) +
                                          (((long long)c.iy - bs::kMagicBias) >> 5) * bs::kConvPitch +
                                          (((long long)c.ix - bs::kMagicBias) >> 5) * 4 - (long long)(bs::kRun / 2) * vstep;
                it.origin_off = (uint32_t)run_rel;
                it.coef.aux = ((uint32_t)ustep & 0xFFFFu) | ((uint32_t)vstep << 16);
                it.flags |= (uint8_t)(((3 * t0 - 4 * (int)it.c0) >> 2) << bs::kItemRawOffShift);
                if (feather && nl == 2) it.flags |= bs::kItemFeather;
                if (r.layer[l].full && r.layer[l].groups == 0 && nl == 1) {
                    it.flags |= bs::kItemDirect;
                    static const bool no_plain = getenv("BS_DEBUG_NO_PLAIN") != nullptr;   // experiment hook: one hot-loop variant only
                    if (!r.layer[l].zero_frac && vstep == bs::kConvPitch && !no_plain) it.flags |= bs::kItemPlain;
                }
            }
        }
    };
    // order: two-layer blocks (two adjacent items each), then the rest heaviest first; within a class
    // raster order, so that blocks processed at the same time are neighbours
    L->n_two = 0;
    for (int t = 0; t < n_blocks; ++t)
        if (recs[t].type == bs::kBlockFast && recs[t].n_layers == 2) {
            emit(recs[t], 0);
            emit(recs[t], 1);
            L->n_fast++;
            L->n_two++;
        }
    // An EXACT block costs a warp about eight plain ones (fp64 coordinates for every camera and pixel):
    // deal it as four items of four rows each (WorkItem::groups = which rows), as long as the item
    // array, sized for two items per block, has room.
    int n_exact_blocks = 0, n_single = 0;
    for (int t = 0; t < n_blocks; ++t) {
        n_exact_blocks += recs[t].type == bs::kBlockExact;
        n_single += recs[t].type != bs::kBlockExact && !(recs[t].type == bs::kBlockFast && recs[t].n_layers == 2);
    }
    const bool split_exact = 2 * L->n_two + n_single + (long long)(bs::kRun / 4) * n_exact_blocks <= 2LL * n_blocks;
    for (int t = 0; t < n_blocks; ++t)
        if (recs[t].type == bs::kBlockExact) {
            if (split_exact) {
                for (int g = 0; g < bs::kRun / 4; ++g) {
                    emit(recs[t], 0);
                    items[n - 1].groups = (uint8_t)(1u << g);
                }
            } else {
                emit(recs[t], 0);
                items[n - 1].groups = (uint8_t)((1u << (bs::kRun / 4)) - 1);
            }
            L->n_exact++;
        }
    for (int t = 0; t < n_blocks; ++t)
        if (recs[t].type == bs::kBlockFast && recs[t].n_layers == 1) emit(recs[t], 0), L->n_fast++;
    for (int t = 0; t < n_blocks; ++t)
        if (recs[t].type == bs::kBlockEmpty) emit(recs[t], 0), L->n_empty++;
    L->n_items = n;
}

// Record ring size and CTAs per SM for a plan: 16 slots per warp unless 8 let one more CTA fit an SM
// (large source boxes, e.g. strong perspective).  Returns CTAs per SM (0: the kernel does not fit).
int choose_rec_slots(int stage_bytes, int *rec_slots, int *smem_bytes)
{
    int per16 = 0, per8 = 0;
    const int s16 = cta_smem_bytes(stage_bytes, bs::kRecSlotsMax), s8 = cta_smem_bytes(stage_bytes, bs::kRecSlotsMax / 2);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per16, bs::stitch_fast_kernel, bs::kFastThreads, s16) != cudaSuccess) per16 = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per8, bs::stitch_fast_kernel, bs::kFastThreads, s8) != cudaSuccess) per8 = 0;
    cudaGetLastError();
    const bool small = per8 > per16;
    *rec_slots = small ? bs::kRecSlotsMax / 2 : bs::kRecSlotsMax;
    *smem_bytes = small ? s8 : s16;
    return small ? per8 : per16;
}

// Deal the plan's blocks round-robin to the n_warps warps of the grid (block u to warp u mod n_warps;
// `in` lists the heaviest blocks first, so they spread evenly) and store every warp's items
// contiguously in `out`.  warp_start has n_warps + 1 entries.  `in` holds the n_two two-layer blocks
// first (two adjacent items each), then one item per block.
void group_items_by_warp(const bs::WorkItem *in, int n_units, int n_two, int n_warps, bs::WorkItem *out, int *warp_start)
{
    int n = 0;
    for (int w = 0; w < n_warps; ++w) {
        warp_start[w] = n;
        for (int u = w; u < n_units; u += n_warps) {
            const int first = u + std::min(u, n_two);
            out[n++] = in[first];
            if (u < n_two) out[n++] = in[first + 1];
        }
    }
    warp_start[n_warps] = n;
}

// Build the block plan on the device and post-process it on the host (once per calibration).
int build_plan(bs_handle *h)
{
    h->fast_ok = false;
    h->blocks_x = (h->W + bs::kBlockW - 1) / bs::kBlockW;
    h->blocks_y = (h->H + bs::kBlockH - 1) / bs::kBlockH;
    h->n_blocks = h->blocks_x * h->blocks_y;
    if ((h->flags & BS_FLAG_EXACT_ONLY) || h->n_cams > bs::kMaxCamsPerLaunch || !get_encoder()) return BS_OK;

    bs::PlanArgs *P = new bs::PlanArgs;
    memset(P, 0, sizeof *P);
    for (int i = 0; i < h->n_cams; ++i) {
        memcpy(P->M[i], &h->M[9 * i], 9 * sizeof(double));
        P->w[i] = h->sw[i];
        P->h[i] = h->sh[i];
    }
    P->n_cams = h->n_cams;
    P->W = h->W;
    P->H = h->H;
    P->block_w = opencv_block_width(h->W, h->H);
    P->legacy_border = (h->flags & BS_FLAG_LEGACY_BORDER) ? 1 : 0;
    P->keep_canvas = (h->flags & BS_FLAG_KEEP_CANVAS) ? 1 : 0;
    P->feather = (h->flags & BS_FLAG_FEATHER) ? 1 : 0;
    P->blocks_x = h->blocks_x;
    P->blocks_y = h->blocks_y;
    // device scratch: raw records first, later overwritten by the item list (at most 2 items of 80 bytes per block)
    const size_t raw_bytes = sizeof(bs::BlockRec) * (size_t)h->n_blocks;
    void *d_plan = nullptr;
    cudaError_t e = cudaMalloc(&d_plan, raw_bytes);
    // per camera, the source rectangle the per-pixel blocks can read (kept beside the staged boxes' rectangles)
    std::vector<int> exact_bbox(4 * (size_t)h->n_cams);
    for (int i = 0; i < h->n_cams; ++i) {
        exact_bbox[4 * i + 0] = exact_bbox[4 * i + 1] = INT_MAX;
        exact_bbox[4 * i + 2] = exact_bbox[4 * i + 3] = -1;
    }
    int *d_bbox = nullptr;
    if (e == cudaSuccess) e = cudaMalloc(&d_bbox, sizeof(int) * exact_bbox.size());
    if (e == cudaSuccess) e = upload_now(h, d_bbox, exact_bbox.data(), sizeof(int) * exact_bbox.size());
    if (e == cudaSuccess) {
        h->d_items = static_cast<bs::WorkItem *>(d_plan);
        P->out = static_cast<bs::BlockRec *>(d_plan);
        P->exact_bbox = d_bbox;
        bs::plan_kernel<<<dim3((h->blocks_x + bs::kPlanWarps - 1) / bs::kPlanWarps, h->blocks_y), 32 * bs::kPlanWarps, 0,
                          h->s_compute>>>(*P);
        e = cudaGetLastError();
    }
    delete P;
    std::vector<bs::BlockRec> recs(h->n_blocks);
    if (e == cudaSuccess) e = cudaMemcpyAsync(recs.data(), d_plan, raw_bytes, cudaMemcpyDeviceToHost, h->s_compute);
    if (e == cudaSuccess) e = cudaMemcpyAsync(exact_bbox.data(), d_bbox, sizeof(int) * exact_bbox.size(), cudaMemcpyDeviceToHost, h->s_compute);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->s_compute);
    cudaFree(d_bbox);
    if (e != cudaSuccess) return fail(h, BS_ERR_CUDA, "block plan failed: %s", cudaGetErrorString(e));
    h->kernel_launches++;
    h->device_bytes += raw_bytes;
    h->scratch_bytes += raw_bytes;

    std::vector<bs::WorkItem> items((size_t)h->n_blocks * bs::kMaxLayers);
    PlanLayout L;
    static const bool no_compact = getenv("BS_DEBUG_NO_COMPACT") != nullptr;   // test hook: always the general staged kernel
    // the compact kernel's fixed TMA box (32 elements x up to 18 rows) must fit inside every frame
    bool frames_fit = true;
    for (int i = 0; i < h->n_cams; ++i) frames_fit = frames_fit && h->sw[i] * 3 >= bs::kRawPitch && h->sh[i] >= bs::kRawRows;
    layout_plan(recs.data(), h->n_blocks, h->n_cams, items.data(), &L, !no_compact && frames_fit, (h->flags & BS_FLAG_FEATHER) != 0);
    h->compact = L.compact;
    memcpy(h->box_we, L.box_we, sizeof h->box_we);
    memcpy(h->box_h, L.box_h, sizeof h->box_h);
    // Source region each camera can contribute: the union of the staged boxes' bounding boxes (blocks the layout
    // demoted to the per-pixel path keep theirs) and of what the plan kernel found for its per-pixel blocks;
    // bs_submit uploads only that part of every frame, and the lens launch computes only that part (build_lens_set).
    static const bool whole_frames = getenv("BS_DEBUG_WHOLE_FRAMES") != nullptr;     // test hook
    h->up_x0.clear();
    if (!whole_frames) {
        h->up_x0.assign(h->n_cams, INT_MAX);
        h->up_y0.assign(h->n_cams, INT_MAX);
        h->up_x1.assign(h->n_cams, -1);
        h->up_y1.assign(h->n_cams, -1);
        for (const bs::BlockRec &r : recs) {
            if (r.type != bs::kBlockFast && r.type != bs::kBlockExact) continue;
            for (int l = 0; l < r.n_layers; ++l) {      // (0 layers in the plan kernel's own per-pixel blocks)
                const int c = r.layer[l].cam;
                h->up_x0[c] = std::min(h->up_x0[c], (int)r.layer[l].x0);
                h->up_y0[c] = std::min(h->up_y0[c], (int)r.layer[l].y0);
                h->up_x1[c] = std::max(h->up_x1[c], (int)r.layer[l].x1);
                h->up_y1[c] = std::max(h->up_y1[c], (int)r.layer[l].y1);
            }
        }
        for (int c = 0; c < h->n_cams; ++c) {
            h->up_x0[c] = std::min(h->up_x0[c], exact_bbox[4 * c + 0]);
            h->up_y0[c] = std::min(h->up_y0[c], exact_bbox[4 * c + 1]);
            h->up_x1[c] = std::max(h->up_x1[c], exact_bbox[4 * c + 2]);
            h->up_y1[c] = std::max(h->up_y1[c], exact_bbox[4 * c + 3]);
        }
    }
    h->stage_bytes = L.stage_bytes;
    h->smem_bytes = L.smem_bytes;
    h->n_items = L.n_items;
    h->n_fast = L.n_fast;
    h->n_exact = L.n_exact;
    h->n_empty = L.n_empty;

    if (getenv("BS_DEBUG_PLAN_SUMMARY")) {
        long flagged = 0, fast_items = 0, full_items = 0;
        for (int i = 0; i < L.n_items; ++i)
            if (items[i].type == bs::kBlockFast) {
                fast_items++;
                flagged += items[i].groups != 0;
                full_items += (items[i].flags & bs::kItemFull) != 0;
            }
        fprintf(stderr, "[plan] blocks %d: fast %lld exact %lld empty %lld; items %d, fast items %ld (full %ld, exact-coordinate groups in %ld)\n",
                h->n_blocks, (long long)L.n_fast, (long long)L.n_exact, (long long)L.n_empty, L.n_items, fast_items,
                full_items, flagged);
        for (int c = 0; c < h->n_cams; ++c)
            fprintf(stderr, "[plan]   cam %d: %dx%d box %d elems x %d rows\n", c, h->sw[c], h->sh[c], h->box_we[c], h->box_h[c]);
        fprintf(stderr, "[plan]   stage %d B, smem/CTA %d B\n", h->stage_bytes, h->smem_bytes);
    }

    int sms = 0, optin = 0, n_warps = 0;
    BS_CUDA(h, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    // The shared-memory limit is a per-device property of a kernel, shared by every handle: always raise it to
    // the device's opt-in maximum, so a handle with a small plan cannot lower it for another one.
    BS_CUDA(h, cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    if (h->compact) {
        h->smem_bytes = bs::kWarps2 * ((h->flags & BS_FLAG_FEATHER) ? bs::kWarp2SmemFeather : bs::kWarp2Smem) + 128;
        int per_sm = 0;
        if (h->smem_bytes > optin) return BS_OK;
        if (h->flags & BS_FLAG_FEATHER) {
            BS_CUDA(h, cudaFuncSetAttribute(bs::stitch_fast2_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
            BS_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast2_kernel<false, true>, bs::kFast2Threads, h->smem_bytes));
        } else {
            BS_CUDA(h, cudaFuncSetAttribute(bs::stitch_fast2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
            BS_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast2_kernel<false>, bs::kFast2Threads, h->smem_bytes));
        }
        if (per_sm < 1) return BS_OK;
        h->grid = std::max(1, std::min((h->n_blocks + bs::kWarps2 - 1) / bs::kWarps2, per_sm * sms));
        if (const char *g = getenv("BS_DEBUG_GRID")) h->grid = std::max(1, atoi(g));
        n_warps = h->grid * bs::kWarps2;
    } else {
        if (cta_smem_bytes(h->stage_bytes, bs::kRecSlotsMax / 2) > optin) return BS_OK;   // does not fit: stay on the exact kernel
        BS_CUDA(h, cudaFuncSetAttribute(bs::stitch_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
        const int per_sm = choose_rec_slots(h->stage_bytes, &h->rec_slots, &h->smem_bytes);
        if (per_sm < 1) return BS_OK;   // does not fit: stay on the exact kernel
        h->grid = std::max(1, std::min((h->n_blocks + bs::kWarpsPerCta - 1) / bs::kWarpsPerCta, per_sm * sms));
        if (const char *g = getenv("BS_DEBUG_GRID")) h->grid = std::max(1, atoi(g));
        n_warps = h->grid * bs::kWarpsPerCta;
    }
    std::vector<bs::WorkItem> grouped(items.size());
    std::vector<int> warp_start(n_warps + 1);
    group_items_by_warp(items.data(), L.n_items - L.n_two, L.n_two, n_warps, grouped.data(), warp_start.data());

    const size_t item_bytes = sizeof(bs::WorkItem) * (size_t)h->n_items;   // <= raw_bytes: 2 x 80 B per block
    BS_CUDA(h, upload_now(h, h->d_items, grouped.data(), item_bytes));
    const size_t ws_bytes = sizeof(int) * (size_t)(n_warps + 1);
    BS_CUDA(h, cudaMalloc(&h->d_warp_start, ws_bytes));
    BS_CUDA(h, upload_now(h, h->d_warp_start, warp_start.data(), ws_bytes));
    h->device_bytes += ws_bytes;
    h->scratch_bytes += ws_bytes;
    h->fast_ok = true;
    return BS_OK;
}

// True when every byte of the canvas has the same upper 32 address bits.
int canvas_in_4g(const void *canvas, size_t pitch, int H)
{
    static const bool never = getenv("BS_DEBUG_NO_DIRECT") != nullptr;   // test hook: always take the staged store path
    const uintptr_t a = reinterpret_cast<uintptr_t>(canvas), b = a + pitch * (size_t)H - 1;
    return (!never && (a >> 32) == (b >> 32)) ? 1 : 0;
}

bool aligned16(const void *p, size_t pitch) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (pitch & 15) == 0; }

// The lens model of camera i as the undistort kernels take it (frame pointers left unset).
bs::UndistortArgs lens_args(const bs_handle *h, const LensModel &m, int i)
{
    bs::UndistortArgs U;
    U.src = nullptr;
    U.dst = nullptr;
    U.spitch = U.dpitch = 0;
    U.w = h->sw[i];
    U.h = h->sh[i];
    U.xs = m.d_xs;
    U.ys = m.d_ys;
    U.winv = m.winv;
    U.fx = m.fx;
    U.fy = m.fy;
    U.u0 = m.u0;
    U.v0 = m.v0;
    U.k1 = m.k[0]; U.k2 = m.k[1]; U.p1 = m.k[2]; U.p2 = m.k[3];
    U.k3 = m.k[4]; U.k4 = m.k[5]; U.k5 = m.k[6]; U.k6 = m.k[7];
    return U;
}

// One launch of the compact frame kernel over lens plans (stitch_undistort.cuh): cv::undistort of one frame of each
// camera in `cams`, src[k] -> dst[k].  `items` / `warp_start` / `cubic` are either one camera's plan or the
// handle's LensSet; every item names its camera.  All buffers 16-byte aligned (the callers check).
int launch_lens(bs_handle *h, const int *cams, int n, const bs::WorkItem *d_items, const int *d_warp_start, const bs::CubicCoef *d_cubic,
                int grid, int smem_bytes, int box_h, const void *const *d_src, const size_t *spitch, uint8_t *const *d_dst,
                const size_t *dpitch, cudaStream_t stream, bool sources_from_kernel)
{
    bs::Fast2Args FA;
    memset(&FA, 0, sizeof FA);
    FA.F.n_cams = h->n_cams < bs::kMaxCamsPerLaunch ? h->n_cams : bs::kMaxCamsPerLaunch;
    FA.F.block_w = 1;
    FA.F.canvas_in_4g = 1;
    for (int k = 0; k < n; ++k) {
        const int c = cams[k];
        LensModel &m = h->lens[c];
        FA.F.cam[c].ptr = static_cast<const uint8_t *>(d_src[k]);
        FA.F.cam[c].pitch = (long long)spitch[k];
        FA.F.cam[c].w = h->sw[c];
        FA.F.cam[c].h = h->sh[c];
        FA.lens_dst[c] = d_dst[k];
        FA.lens_dpitch[c] = (long long)dpitch[k];
        if (!canvas_in_4g(d_dst[k], dpitch[k], h->sh[c])) FA.F.canvas_in_4g = 0;
        TmapEntry *hit = nullptr;
        for (int w = 0; w < kTmapCacheWays && !hit; ++w)
            if (m.tmaps[w].ptr == d_src[k] && m.tmaps[w].pitch == spitch[k] && m.tmaps[w].box_h == box_h) hit = &m.tmaps[w];
        if (!hit) {
            hit = &m.tmaps[m.tmap_next++ % kTmapCacheWays];
            hit->ptr = nullptr;
            const int er = encode_frame_map(&hit->map, d_src[k], spitch[k], h->sh[c], bs::kRawPitch / 4, box_h);
            if (er != 0) return fail(h, BS_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for the lens plan of camera %d", er, c);
            hit->ptr = d_src[k];
            hit->pitch = spitch[k];
            hit->box_h = box_h;
        }
        FA.tmap[c] = hit->map;
    }
    FA.F.canvas = FA.lens_dst[cams[0]];      // not used by lens launches; kept valid for debuggers
    FA.F.cpitch = FA.lens_dpitch[cams[0]];
    FA.F.W = h->sw[cams[0]];
    FA.F.H = h->sh[cams[0]];
    FA.items = d_items;
    FA.warp_start = d_warp_start;
    FA.box_bytes = bs::kRawPitch * box_h;
    FA.wait_sources = sources_from_kernel ? 1 : 0;
    FA.lens_cubic = d_cubic;
    FA.lens_models = h->d_lens_models;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(bs::kFast2Threads);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    BS_CUDA(h, cudaLaunchKernelEx(&cfg, bs::stitch_fast2_kernel<true>, FA));
    static const bool dbg_sync_lens = getenv("BS_DEBUG_SYNC") != nullptr;
    if (dbg_sync_lens) {
        cudaError_t e1 = cudaDeviceSynchronize();
        if (e1 != cudaSuccess)
            return fail(h, BS_ERR_CUDA, "[debug] lens kernel over %d camera(s) (first %d, grid %d, smem %d, box rows %d) faulted: %s", n, cams[0],
                        grid, smem_bytes, box_h, cudaGetErrorString(e1));
    }
    h->kernel_launches++;
    return BS_OK;
}

// cv::undistort of one frame of camera `cam` (which has a lens model) between device buffers.
// `sources_from_kernel`: see launch_frame.
int undistort_one(bs_handle *h, int cam, const void *d_src, size_t spitch, uint8_t *d_dst, size_t dpitch, cudaStream_t stream,
                  bool sources_from_kernel)
{
    LensModel &m = h->lens[cam];
    const bool aligned4 = (reinterpret_cast<uintptr_t>(d_src) & 3) == 0 && (spitch & 3) == 0 &&
                          (reinterpret_cast<uintptr_t>(d_dst) & 3) == 0 && (dpitch & 3) == 0;
    if (m.plan_ok && aligned16(d_src, spitch) && aligned16(d_dst, dpitch)) {
        // the lens field as a block plan: the compact frame kernel resamples the frame (stitch_undistort.cuh)
        return launch_lens(h, &cam, 1, m.d_items, m.d_warp_start, m.d_cubic, m.grid, m.smem_bytes, m.box_h, &d_src, &spitch, &d_dst, &dpitch,
                           stream, sources_from_kernel);
    } else if (m.d_map && aligned4) {
        // cached map: one word per pixel instead of the fp64 lens model (stitch_undistort.cuh)
        bs::RemapArgs R;
        R.src = static_cast<const uint8_t *>(d_src);
        R.dst = d_dst;
        R.spitch = (long long)spitch;
        R.dpitch = (long long)dpitch;
        R.w = h->sw[cam];
        R.h = h->sh[cam];
        R.map = m.d_map;
        bs::undistort_remap_kernel<<<h->remap_grid, 256, 0, stream>>>(R);
    } else {
        bs::UndistortArgs U = lens_args(h, m, cam);
        U.src = static_cast<const uint8_t *>(d_src);
        U.dst = d_dst;
        U.spitch = (long long)spitch;
        U.dpitch = (long long)dpitch;
        bs::undistort_kernel<<<dim3((U.w + 63) / 64, (U.h + 3) / 4), 256, 0, stream>>>(U);
    }
    BS_CUDA(h, cudaGetLastError());
    h->kernel_launches++;
    return BS_OK;
}

// The buffer an undistorted frame of camera `cam` is written to (one per user of the lens model).
int alloc_undist(bs_handle *h, std::vector<uint8_t *> &tmp, int cam)
{
    if ((int)tmp.size() < h->n_cams) tmp.resize(h->n_cams, nullptr);
    if (tmp[cam]) return BS_OK;
    const size_t bytes = h->d_frame_pitch[cam] * (size_t)h->sh[cam];
    BS_CUDA(h, cudaMalloc(&tmp[cam], bytes));
    h->device_bytes += bytes;
    return BS_OK;
}

// cv::undistort for the cameras that have a lens model: src frame -> tmp[i]; the pointers the warp
// should read are returned in out_frames / out_pitch.  Stream-ordered before the frame kernel.
int undistort_frames(bs_handle *h, const void *const *d_frames, const size_t *pitch, std::vector<uint8_t *> &tmp,
                     const void **out_frames, size_t *out_pitch, cudaStream_t stream, bool sources_from_kernel, bool *launched = nullptr)
{
    if (launched) *launched = false;
    // all lens plans in one launch when every member's buffers qualify
    const LensSet &S = h->lens_set;
    bool use_set = S.valid;
    for (size_t k = 0; k < S.cams.size() && use_set; ++k) {
        const int c = S.cams[k];
        use_set = (int)tmp.size() >= h->n_cams && tmp[c] && aligned16(d_frames[c], pitch[c]) && aligned16(tmp[c], h->d_frame_pitch[c]);
    }
    if (use_set) {
        const void *src[bs::kMaxCamsPerLaunch];
        size_t sp[bs::kMaxCamsPerLaunch], dp[bs::kMaxCamsPerLaunch];
        uint8_t *dst[bs::kMaxCamsPerLaunch];
        const int n = (int)S.cams.size();
        for (int k = 0; k < n; ++k) {
            const int c = S.cams[k];
            src[k] = d_frames[c];
            sp[k] = pitch[c];
            dst[k] = tmp[c];
            dp[k] = h->d_frame_pitch[c];
        }
        const int rc = launch_lens(h, S.cams.data(), n, S.d_items, S.d_warp_start, S.d_cubic, S.grid, S.smem_bytes, S.box_h, src, sp, dst, dp,
                                   stream, sources_from_kernel);
        if (rc != BS_OK) return rc;
        if (launched) *launched = true;
    }
    for (int i = 0; i < h->n_cams; ++i) {
        out_frames[i] = d_frames[i];
        out_pitch[i] = pitch[i];
        if (h->lens.empty() || !h->lens[i].enabled) continue;
        const LensModel &m = h->lens[i];
        if ((int)tmp.size() < h->n_cams || !tmp[i])     // allocated by bs_set_undistort / ensure_ring, never here
            return fail(h, BS_ERR_INVALID_ARG, "internal: no undistort buffer for camera %d", i);
        out_frames[i] = tmp[i];
        out_pitch[i] = h->d_frame_pitch[i];
        if (use_set && m.plan_ok && std::find(S.cams.begin(), S.cams.end(), i) != S.cams.end()) continue;   // done above
        if (use_set && S.trimmed && h->up_x1[i] < 0) continue;                 // no panorama pixel reads this camera
        const int rc = undistort_one(h, i, d_frames[i], pitch[i], tmp[i], h->d_frame_pitch[i], stream, sources_from_kernel);
        if (rc != BS_OK) return rc;
        if (launched) *launched = true;
    }
    return BS_OK;
}

// Host-side tables of the lens model: OpenCV's stripe-wise inverse camera matrix and the running
// sum it forms along a row (see stitch_undistort.cuh).  Only the usual camera-matrix form is accepted.
int build_lens_tables(int w, int h, const double K[9], LensModel &m, std::vector<double> &xs, std::vector<double> &ys)
{
    if (K[1] != 0.0 || K[3] != 0.0 || K[6] != 0.0 || K[7] != 0.0 || K[8] != 1.0 || K[0] == 0.0 || K[4] == 0.0) return BS_ERR_UNSUPPORTED;
    m.fx = K[0];
    m.fy = K[4];
    m.u0 = K[2];
    m.v0 = K[5];
    int stripe0 = 4096 / (w > 1 ? w : 1);
    if (stripe0 < 1) stripe0 = 1;
    if (stripe0 > h) stripe0 = h;
    xs.resize(w);
    ys.resize(h);
    for (int ys0 = 0; ys0 < h; ys0 += stripe0) {
        double Ar[9], ir[9];
        memcpy(Ar, K, sizeof Ar);
        Ar[5] = K[5] - ys0;
        invert3x3(Ar, ir);
        const int rows = std::min(stripe0, h - ys0);
        for (int i = 0; i < rows; ++i) ys[ys0 + i] = i * ir[4] + ir[5];
        if (ys0 == 0) {
            double _x = 0 * ir[1] + ir[2];
            for (int j = 0; j < w; ++j, _x += ir[0]) xs[j] = _x;
            m.winv = 1.0 / (0 * ir[7] + ir[8]);
        }
    }
    return BS_OK;
}

// Launch the kernels of one frame.  Cameras beyond the per-launch limit go in further
// launches that keep what is already on the canvas (later cameras win, Appendix A.5).
// `sources_from_kernel`: the frames may have been written by the kernel directly before this launch in
// `stream` (an undistort kernel of ours, a caller's decoder): every source read is then ordered behind it.
int launch_frame(bs_handle *h, const void *const *d_frames, const size_t *pitch, uint8_t *d_canvas,
                 size_t cpitch, bool keep, cudaStream_t stream, bool sources_from_kernel)
{
    static const bool dbg_sync = getenv("BS_DEBUG_SYNC") != nullptr;
    if (dbg_sync) {
        cudaError_t e0 = cudaDeviceSynchronize();
        if (e0 != cudaSuccess) return fail(h, BS_ERR_CUDA, "[debug] error pending BEFORE launch_frame: %s", cudaGetErrorString(e0));
    }
    bool fast = h->fast_ok && aligned16(d_canvas, cpitch);
    for (int i = 0; fast && i < h->n_cams; ++i) fast = aligned16(d_frames[i], pitch[i]);
    if (fast) {
        // both staged kernels take the same leading members; the general one adds its box geometry
        bs::FastArgs FA1;
        bs::Fast2Args FA2;
        memset(&FA1, 0, sizeof FA1);
        memset(&FA2, 0, sizeof FA2);
        bs::FrameArgs &A = h->compact ? FA2.F : FA1.F;
        CUtensorMap *tmaps = h->compact ? FA2.tmap : FA1.tmap;
        A.n_cams = h->n_cams;
        for (int i = 0; i < A.n_cams; ++i) {
            memcpy(A.cam[i].M, &h->M[9 * i], 9 * sizeof(double));
            A.cam[i].ptr = static_cast<const uint8_t *>(d_frames[i]);
            A.cam[i].pitch = (long long)pitch[i];
            A.cam[i].w = h->sw[i];
            A.cam[i].h = h->sh[i];
            if (!h->compact) {
                FA1.pitch[i] = h->box_we[i] * 4;
                FA1.box_bytes[i] = h->box_we[i] * 4 * h->box_h[i];
            }
            if (h->box_we[i] > 0) {
                // frame buffers come from small rings (ours or the caller's): keep the last few encoded maps
                if (h->tmap_cache.empty()) h->tmap_cache.resize((size_t)h->n_cams * kTmapCacheWays);
                TmapEntry *ways = &h->tmap_cache[(size_t)i * kTmapCacheWays], *hit = nullptr;
                for (int k = 0; k < kTmapCacheWays && !hit; ++k)
                    if (ways[k].ptr == d_frames[i] && ways[k].pitch == pitch[i]) hit = &ways[k];
                if (!hit) {
                    hit = &ways[h->tmap_next[i]++ % kTmapCacheWays];
                    hit->ptr = nullptr;
                    const int er = encode_frame_map(&hit->map, d_frames[i], pitch[i], h->sh[i], h->box_we[i], h->box_h[i]);
                    if (er != 0)
                        return fail(h, BS_ERR_CUDA,
                                    "cuTensorMapEncodeTiled failed (%d) for camera %d: ptr=%p pitch=%zu h=%d box=%dx%d", er,
                                    i, d_frames[i], pitch[i], h->sh[i], h->box_we[i], h->box_h[i]);
                    hit->ptr = d_frames[i];
                    hit->pitch = pitch[i];
                }
                tmaps[i] = hit->map;
            }
        }
        A.canvas = d_canvas;
        A.cpitch = (long long)cpitch;
        A.W = h->W;
        A.H = h->H;
        A.block_w = opencv_block_width(h->W, h->H);
        A.legacy_border = (h->flags & BS_FLAG_LEGACY_BORDER) ? 1 : 0;
        A.keep_canvas = keep ? 1 : 0;
        A.feather = (h->flags & BS_FLAG_FEATHER) ? 1 : 0;
        A.canvas_in_4g = canvas_in_4g(d_canvas, cpitch, h->H);
        // Programmatic dependent launch: the prologue and first block of this frame may overlap the
        // tail of the previous kernel on the same stream (see stitch_fast.cuh).  When that kernel may
        // have produced the source frames, the kernel waits for it before its first source read.
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3(h->grid);
        cfg.dynamicSmemBytes = (size_t)h->smem_bytes;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (h->compact) {
            FA2.items = h->d_items;
            FA2.warp_start = h->d_warp_start;
            FA2.wait_sources = sources_from_kernel ? 1 : 0;
            FA2.box_bytes = bs::kRawPitch * h->box_h[0];
            cfg.blockDim = dim3(bs::kFast2Threads);
            if (h->flags & BS_FLAG_FEATHER) BS_CUDA(h, cudaLaunchKernelEx(&cfg, bs::stitch_fast2_kernel<false, true>, FA2));
            else BS_CUDA(h, cudaLaunchKernelEx(&cfg, bs::stitch_fast2_kernel<false>, FA2));
        } else {
            FA1.items = h->d_items;
            FA1.warp_start = h->d_warp_start;
            FA1.stage_bytes = h->stage_bytes;
            FA1.rec_slots = h->rec_slots;
            FA1.wait_sources = sources_from_kernel ? 1 : 0;
            cfg.blockDim = dim3(bs::kFastThreads);
            BS_CUDA(h, cudaLaunchKernelEx(&cfg, bs::stitch_fast_kernel, FA1));
        }
        BS_CUDA(h, cudaGetLastError());
        if (dbg_sync) {
            cudaError_t e1 = cudaDeviceSynchronize();
            if (e1 != cudaSuccess)
                return fail(h, BS_ERR_CUDA, "[debug] staged kernel (compact %d, grid %d, smem %d) faulted: %s", (int)h->compact,
                            h->grid, h->smem_bytes, cudaGetErrorString(e1));
        }
        h->kernel_launches++;
        return BS_OK;
    }
    for (int base = 0; base < h->n_cams; base += bs::kMaxCamsPerLaunch) {
        bs::FrameArgs A;
        memset(&A, 0, sizeof A);
        A.n_cams = std::min(bs::kMaxCamsPerLaunch, h->n_cams - base);
        for (int i = 0; i < A.n_cams; ++i) {
            memcpy(A.cam[i].M, &h->M[9 * (base + i)], 9 * sizeof(double));
            A.cam[i].ptr = static_cast<const uint8_t *>(d_frames[base + i]);
            A.cam[i].pitch = (long long)pitch[base + i];
            A.cam[i].w = h->sw[base + i];
            A.cam[i].h = h->sh[base + i];
        }
        A.canvas = d_canvas;
        A.cpitch = (long long)cpitch;
        A.W = h->W;
        A.H = h->H;
        A.block_w = opencv_block_width(h->W, h->H);
        A.legacy_border = (h->flags & BS_FLAG_LEGACY_BORDER) ? 1 : 0;
        A.keep_canvas = (keep || base > 0) ? 1 : 0;
        A.cam_base = base;
        A.border_constant = (h->flags & BS_FLAG_BORDER_CONSTANT) ? 1 : 0;
        A.feather = (h->flags & BS_FLAG_FEATHER) ? 1 : 0;
        dim3 grid((h->W + bs::kExactTileW - 1) / bs::kExactTileW,
                  (h->H + bs::kExactTileH - 1) / bs::kExactTileH);
        bs::stitch_exact_kernel<<<grid, dim3(32, bs::kExactTileH), 0, stream>>>(A);
        BS_CUDA(h, cudaGetLastError());
        h->kernel_launches++;
    }
    return BS_OK;
}

int ensure_ring(bs_handle *h)
{
    if (!h->ring.empty()) return BS_OK;
    h->ring.resize(kPipelineDepth);
    for (Slot &s : h->ring) {
        s.d_frame.assign(h->n_cams, nullptr);
        for (int i = 0; i < h->n_cams; ++i) {
            size_t bytes = h->d_frame_pitch[i] * (size_t)h->sh[i];
            BS_CUDA(h, cudaMalloc(&s.d_frame[i], bytes));
            h->device_bytes += bytes;
        }
        size_t cbytes = h->d_canvas_pitch * (size_t)h->H;
        BS_CUDA(h, cudaMalloc(&s.d_canvas, cbytes));
        h->device_bytes += cbytes;
        BS_CUDA(h, cudaEventCreateWithFlags(&s.ev_h2d, cudaEventDisableTiming));
        BS_CUDA(h, cudaEventCreateWithFlags(&s.ev_kernel, cudaEventDisableTiming));
        BS_CUDA(h, cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming));
        for (int i = 0; i < h->n_cams && !h->lens.empty(); ++i)
            if (h->lens[i].enabled) {
                const int rc = alloc_undist(h, s.d_undist, i);
                if (rc != BS_OK) return rc;
            }
    }
    return BS_OK;
}

// Queue one frame from host buffers on slot `s`: uploads on s_h2d, kernels on s_compute, download on s_d2h.
int issue_frame(bs_handle *h, Slot &s, const uint8_t *const *frames, const size_t *pitch, uint8_t *out, size_t out_pitch)
{
    h->h2d_bytes_last = 0;
    for (int i = 0; i < h->n_cams; ++i) {
        int x0 = 0, y0 = 0, x1 = h->sw[i] - 1, y1 = h->sh[i] - 1;
        const bool has_lens = !h->lens.empty() && h->lens[i].enabled;
        const LensSet &S = h->lens_set;
        bool whole = !h->fast_ok || h->up_x0.empty() || has_lens;
        if (has_lens && S.valid && S.trimmed) {
            // the lens launch reads only the boxes of the blocks the panorama needs (build_lens_set)
            if (h->up_x1[i] < 0) continue;           // no panorama pixel reads this camera
            if (S.src_x1[i] >= 0) {
                x0 = (S.src_x0[i] / 64) * 64;
                y0 = S.src_y0[i];
                x1 = std::min(x1, S.src_x1[i]);
                y1 = std::min(y1, S.src_y1[i]);
            }
        } else if (!whole) {
            if (h->up_x1[i] < 0) continue;           // this camera contributes nothing to the canvas
            x0 = (h->up_x0[i] / 64) * 64;            // rows start on a 64-byte boundary of the frame (3 * 64 texel bytes)
            y0 = h->up_y0[i];
            x1 = std::min(x1, h->up_x1[i]);
            y1 = std::min(y1, h->up_y1[i]);
        }
        BS_CUDA(h, cudaMemcpy2DAsync(s.d_frame[i] + (size_t)y0 * h->d_frame_pitch[i] + (size_t)x0 * 3, h->d_frame_pitch[i],
                                     frames[i] + (size_t)y0 * pitch[i] + (size_t)x0 * 3, pitch[i], (size_t)(x1 - x0 + 1) * 3,
                                     (size_t)(y1 - y0 + 1), cudaMemcpyHostToDevice, h->s_h2d));
        h->h2d_bytes_last += (uint64_t)(x1 - x0 + 1) * 3 * (uint64_t)(y1 - y0 + 1);
    }
    const bool keep = (h->flags & BS_FLAG_KEEP_CANVAS) != 0;
    if (keep)
        BS_CUDA(h, cudaMemcpy2DAsync(s.d_canvas, h->d_canvas_pitch, out, out_pitch, (size_t)h->W * 3,
                                     h->H, cudaMemcpyHostToDevice, h->s_h2d));
    BS_CUDA(h, cudaEventRecord(s.ev_h2d, h->s_h2d));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_compute, s.ev_h2d, 0));
    std::vector<const void *> d_frames(s.d_frame.begin(), s.d_frame.end()), warp_frames(h->n_cams);
    std::vector<size_t> warp_pitch(h->n_cams);
    bool lens_ran = false;
    int rc = undistort_frames(h, d_frames.data(), h->d_frame_pitch.data(), s.d_undist, warp_frames.data(), warp_pitch.data(),
                              h->s_compute, false, &lens_ran);
    if (rc != BS_OK) return rc;
    // without a lens kernel the frames reach the device by copies on another stream, ordered by an event
    rc = launch_frame(h, warp_frames.data(), warp_pitch.data(), s.d_canvas, h->d_canvas_pitch, keep, h->s_compute, lens_ran);
    if (rc != BS_OK) return rc;
    BS_CUDA(h, cudaEventRecord(s.ev_kernel, h->s_compute));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_d2h, s.ev_kernel, 0));
    BS_CUDA(h, cudaMemcpy2DAsync(out, out_pitch, s.d_canvas, h->d_canvas_pitch, (size_t)h->W * 3, h->H,
                                 cudaMemcpyDeviceToHost, h->s_d2h));
    BS_CUDA(h, cudaEventRecord(s.ev_done, h->s_d2h));
    return BS_OK;
}

int retrieve_jpeg(bs_handle *h, Slot &s, int slot_index);    // stitch_jpeg.inl
int flush_deferred_jpeg(bs_handle *h, int64_t upto, bool only_ready = false);
void jpeg_ctx_destroy(JpegCtx *c);
}  // namespace

extern "C" {

int bs_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char *bs_version(void) { return "b200-stitch 0.1 (sm_100a)"; }

const char *bs_last_error(const bs_handle *h) { return h ? h->err : g_last_error; }

int bs_create(int device_id, int n_cams, const int *src_w, const int *src_h, int canvas_w,
              int canvas_h, const double *H, unsigned flags, bs_handle **out)
{
    if (!out) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: out is NULL");
    *out = nullptr;
    if (n_cams < 1 || n_cams > BS_MAX_CAMERAS)
        return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: n_cams=%d outside [1, %d]", n_cams,
                    BS_MAX_CAMERAS);
    if (!src_w || !src_h || !H) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: NULL array");
    if (canvas_w < 1 || canvas_h < 1 || canvas_w > 65535 || canvas_h > 65535)
        return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: canvas %dx%d outside [1, 65535]",
                    canvas_w, canvas_h);
    for (int i = 0; i < n_cams; ++i)
        if (src_w[i] < 1 || src_h[i] < 1 || src_w[i] > BS_MAX_DIM || src_h[i] > BS_MAX_DIM)
            return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: camera %d is %dx%d, outside [1, %d]",
                        i, src_w[i], src_h[i], BS_MAX_DIM);
    if (flags & ~(BS_FLAG_LEGACY_BORDER | BS_FLAG_KEEP_CANVAS | BS_FLAG_EXACT_ONLY | BS_FLAG_BORDER_CONSTANT | BS_FLAG_FEATHER))
        return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: unknown flags 0x%x", flags);
    if (flags & BS_FLAG_FEATHER) {
        if (flags & BS_FLAG_BORDER_CONSTANT)
            return fail(nullptr, BS_ERR_UNSUPPORTED, "bs_create: BS_FLAG_FEATHER with BS_FLAG_BORDER_CONSTANT: every camera covers every pixel there");
        if (n_cams > bs::kMaxCamsPerLaunch)
            return fail(nullptr, BS_ERR_UNSUPPORTED, "bs_create: BS_FLAG_FEATHER takes at most %d cameras (one launch must see all of them), got %d",
                        bs::kMaxCamsPerLaunch, n_cams);
    }
    int rc = require_device(nullptr);
    if (rc != BS_OK) return rc;
    int ndev = bs_device_count();
    if (device_id < 0 || device_id >= ndev)
        return fail(nullptr, BS_ERR_INVALID_ARG, "bs_create: device %d not in [0, %d)", device_id, ndev);

    bs_handle *h = new (std::nothrow) bs_handle;
    if (!h) return fail(nullptr, BS_ERR_NOMEM, "bs_create: out of host memory");
    h->device = device_id;
    h->n_cams = n_cams;
    h->sw.assign(src_w, src_w + n_cams);
    h->sh.assign(src_h, src_h + n_cams);
    h->W = canvas_w;
    h->H = canvas_h;
    h->flags = (flags & BS_FLAG_BORDER_CONSTANT) ? (flags | BS_FLAG_EXACT_ONLY) : flags;   // no staged path for that mode
    h->M.resize(9 * (size_t)n_cams);
    for (int i = 0; i < n_cams; ++i) invert3x3(H + 9 * i, &h->M[9 * i]);
    h->d_frame_pitch.resize(n_cams);
    for (int i = 0; i < n_cams; ++i) h->d_frame_pitch[i] = align_up((size_t)src_w[i] * 3, 128);
    h->d_canvas_pitch = align_up((size_t)canvas_w * 3, 128);

    cudaError_t e = cudaSetDevice(device_id);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->s_compute, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        fail(nullptr, BS_ERR_CUDA, "bs_create: stream set-up failed: %s", cudaGetErrorString(e));
        bs_destroy(h);
        return BS_ERR_CUDA;
    }
    g_last_device = device_id;
    rc = build_plan(h);
    if (rc != BS_OK) {
        memcpy(g_last_error, h->err, kErrLen);
        bs_destroy(h);
        return rc;
    }
    *out = h;
    return BS_OK;
}

int bs_destroy(bs_handle *h)
{
    if (!h) return BS_OK;
    cudaSetDevice(h->device);
    if (h->s_h2d) cudaStreamSynchronize(h->s_h2d);
    if (h->s_compute) cudaStreamSynchronize(h->s_compute);
    if (h->s_d2h) cudaStreamSynchronize(h->s_d2h);
    for (LensModel &m : h->lens) {
        cudaFree(m.d_xs);
        cudaFree(m.d_ys);
        cudaFree(m.d_map);
        cudaFree(m.d_items);
        cudaFree(m.d_cubic);
        cudaFree(m.d_warp_start);
    }
    jpeg_ctx_destroy(h->jpeg);
    for (uint8_t *p : h->d_undist) cudaFree(p);
    for (Slot &s : h->ring) {
        for (uint8_t *p : s.d_undist) cudaFree(p);
        for (uint8_t *p : s.d_frame) cudaFree(p);
        cudaFree(s.d_canvas);
        if (s.ev_h2d) cudaEventDestroy(s.ev_h2d);
        if (s.ev_kernel) cudaEventDestroy(s.ev_kernel);
        if (s.ev_done) cudaEventDestroy(s.ev_done);
    }
    cudaFree(h->d_items);
    cudaFree(h->d_warp_start);
    cudaFree(h->lens_set.d_items);
    cudaFree(h->lens_set.d_cubic);
    cudaFree(h->lens_set.d_warp_start);
    cudaFree(h->d_lens_models);
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_compute) cudaStreamDestroy(h->s_compute);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    delete h;
    return BS_OK;
}

int bs_pipeline_depth(const bs_handle *) { return kPipelineDepth; }

int bs_submit(bs_handle *h, const uint8_t *const *frames, const size_t *pitch, uint8_t *out,
              size_t out_pitch, int64_t *ticket)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_submit: NULL handle");
    if (!frames || !pitch || !out) return fail(h, BS_ERR_INVALID_ARG, "bs_submit: NULL argument");
    for (int i = 0; i < h->n_cams; ++i) {
        if (!frames[i]) return fail(h, BS_ERR_INVALID_ARG, "bs_submit: frame %d is NULL", i);
        if (pitch[i] < (size_t)h->sw[i] * 3)
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit: pitch[%d]=%zu < %d*3", i, pitch[i], h->sw[i]);
    }
    if (out_pitch < (size_t)h->W * 3)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit: out_pitch=%zu < %d*3", out_pitch, h->W);
    BS_CUDA(h, cudaSetDevice(h->device));
    int rc = ensure_ring(h);
    if (rc != BS_OK) return rc;

    flush_deferred_jpeg(h, h->next_ticket - 1);     // compressed frames submitted earlier go to the streams first
    const int64_t t = h->next_ticket;
    Slot &s = h->ring[t % kPipelineDepth];
    if (s.busy || s.deferred_rc != BS_OK) {   // the slot's previous frame must have left the device (or report why it did not)
        rc = bs_wait(h, s.ticket);
        if (rc != BS_OK) return rc;
    }
    rc = issue_frame(h, s, frames, pitch, out, out_pitch);
    if (rc != BS_OK) {
        // Part of the frame may have been queued: let the three streams run dry so that nothing still
        // refers to the slot or to the caller's buffers, and leave the slot free.  The handle stays usable.
        cudaStreamSynchronize(h->s_h2d);
        cudaStreamSynchronize(h->s_compute);
        cudaStreamSynchronize(h->s_d2h);
        cudaGetLastError();
        s.busy = false;
        s.ticket = -1;
        return rc;
    }
    s.ticket = t;
    s.busy = true;
    h->next_ticket++;
    if (ticket) *ticket = t;
    return BS_OK;
}

int bs_wait(bs_handle *h, int64_t ticket)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_wait: NULL handle");
    if (ticket < 0 || ticket >= h->next_ticket)
        return fail(h, BS_ERR_BAD_TICKET, "bs_wait: ticket %lld was never issued", (long long)ticket);
    if (ticket < h->next_ticket - kPipelineDepth) return BS_OK;   // its slot was already recycled => done
    Slot &s = h->ring[ticket % kPipelineDepth];
    if (s.ticket != ticket) return BS_OK;
    BS_CUDA(h, cudaSetDevice(h->device));
    if (s.deferred) flush_deferred_jpeg(h, ticket);             // a compressed frame whose GPU side is still to be issued
    if (s.deferred_rc != BS_OK) {                               // ... and could not be
        const int rc = s.deferred_rc;
        s.deferred_rc = BS_OK;
        return fail(h, rc, "%s", s.deferred_err);
    }
    if (!s.busy) return BS_OK;
    // While this frame finishes, later compressed frames whose host stages complete go to the GPU (they would
    // otherwise wait for the caller's next call): poll instead of blocking as long as such frames exist.
    for (;;) {
        bool later_deferred = false;
        for (const Slot &o : h->ring) later_deferred = later_deferred || (o.deferred && o.ticket > ticket);
        if (!later_deferred) break;
        flush_deferred_jpeg(h, h->next_ticket - 1, true);
        const cudaError_t q = cudaEventQuery(s.ev_done);
        if (q == cudaSuccess) break;
        if (q != cudaErrorNotReady) return fail(h, BS_ERR_CUDA, "bs_wait: %s", cudaGetErrorString(q));
        std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    BS_CUDA(h, cudaEventSynchronize(s.ev_done));
    s.busy = false;
    return retrieve_jpeg(h, s, (int)(ticket % kPipelineDepth));
}

int bs_stitch(bs_handle *h, const uint8_t *const *frames, const size_t *pitch, uint8_t *out,
              size_t out_pitch)
{
    int64_t t = -1;
    int rc = bs_submit(h, frames, pitch, out, out_pitch, &t);
    if (rc != BS_OK) return rc;
    return bs_wait(h, t);
}

int bs_submit_device_ex(bs_handle *h, const void *const *d_frames, const size_t *pitch, void *d_canvas,
                        size_t canvas_pitch, void *stream, unsigned submit_flags)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_submit_device: NULL handle");
    if (!d_frames || !pitch || !d_canvas)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: NULL argument");
    if (submit_flags & ~BS_SUBMIT_SOURCES_SETTLED)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: unknown submit flags 0x%x", submit_flags);
    for (int i = 0; i < h->n_cams; ++i) {
        if (!d_frames[i]) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: frame %d is NULL", i);
        if (pitch[i] < (size_t)h->sw[i] * 3)
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: pitch[%d]=%zu < %d*3", i, pitch[i],
                        h->sw[i]);
    }
    if (canvas_pitch < (size_t)h->W * 3)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: canvas_pitch=%zu < %d*3", canvas_pitch, h->W);
    // the caller's current device is left as it was
    int prev_dev = -1;
    BS_CUDA(h, cudaGetDevice(&prev_dev));
    if (prev_dev != h->device) BS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);   // NULL = the legacy default stream
    std::vector<const void *> warp_frames(h->n_cams);
    std::vector<size_t> warp_pitch(h->n_cams);
    bool lens_ran = false;
    const bool from_kernel = !(submit_flags & BS_SUBMIT_SOURCES_SETTLED);
    int rc = undistort_frames(h, d_frames, pitch, h->d_undist, warp_frames.data(), warp_pitch.data(), st, from_kernel, &lens_ran);
    if (rc == BS_OK)
        rc = launch_frame(h, warp_frames.data(), warp_pitch.data(), static_cast<uint8_t *>(d_canvas), canvas_pitch,
                          (h->flags & BS_FLAG_KEEP_CANVAS) != 0, st, lens_ran || !(submit_flags & BS_SUBMIT_SOURCES_SETTLED));
    if (prev_dev != h->device) cudaSetDevice(prev_dev);
    return rc;
}

int bs_submit_device(bs_handle *h, const void *const *d_frames, const size_t *pitch, void *d_canvas,
                     size_t canvas_pitch, void *stream)
{
    return bs_submit_device_ex(h, d_frames, pitch, d_canvas, canvas_pitch, stream, 0u);
}

int bs_sync(bs_handle *h)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_sync: NULL handle");
    BS_CUDA(h, cudaSetDevice(h->device));
    flush_deferred_jpeg(h, h->next_ticket - 1);
    BS_CUDA(h, cudaStreamSynchronize(h->s_h2d));
    BS_CUDA(h, cudaStreamSynchronize(h->s_compute));
    BS_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    for (Slot &s : h->ring) s.busy = false;
    return BS_OK;
}

}  // extern "C"

namespace {
// Drop a lens model and give its device memory back (the bytes leave the handle's account).
void clear_lens(bs_handle *h, LensModel &m)
{
    cudaFree(m.d_xs);
    cudaFree(m.d_ys);
    cudaFree(m.d_map);
    cudaFree(m.d_items);
    cudaFree(m.d_cubic);
    cudaFree(m.d_warp_start);
    h->device_bytes -= m.bytes;
    m = LensModel();
}
// Block plan of camera `cam`'s lens field (lens_plan_kernel) for the compact frame kernel.  Leaves m.plan_ok
// false when the field does not fit it (source boxes too large, frame too small): the caller then falls back
// to the cached per-pixel map.
int build_lens_plan(bs_handle *h, LensModel &m, int cam)
{
    const int w = h->sw[cam], hh = h->sh[cam];
    // (an item names its camera, and the launch carries 16 cameras: later cameras keep the per-pixel map)
    if (!get_encoder() || w * 3 < bs::kRawPitch || hh < bs::kRawRows || cam >= bs::kMaxCamsPerLaunch) return BS_OK;
    const int blocks_x = (w + bs::kBlockW - 1) / bs::kBlockW, blocks_y = (hh + bs::kBlockH - 1) / bs::kBlockH;
    const int n_blocks = blocks_x * blocks_y;
    const size_t raw_bytes = sizeof(bs::BlockRec) * (size_t)n_blocks;
    void *d_plan = nullptr;
    bs::CubicCoef *d_cubic_blocks = nullptr;
    BS_CUDA(h, cudaMalloc(&d_plan, raw_bytes));
    if (cudaMalloc(&d_cubic_blocks, sizeof(bs::CubicCoef) * (size_t)n_blocks) != cudaSuccess) {
        cudaFree(d_plan);
        return fail(h, BS_ERR_CUDA, "lens plan: out of device memory");
    }
    bs::LensPlanArgs P;
    P.U = lens_args(h, m, cam);
    P.blocks_x = blocks_x;
    P.blocks_y = blocks_y;
    P.out = static_cast<bs::BlockRec *>(d_plan);
    P.cubic = d_cubic_blocks;
    bs::lens_plan_kernel<<<dim3((blocks_x + bs::kPlanWarps - 1) / bs::kPlanWarps, blocks_y), 32 * bs::kPlanWarps, 0, h->s_compute>>>(P);
    std::vector<bs::BlockRec> recs(n_blocks);
    std::vector<bs::CubicCoef> cubic_blocks(n_blocks);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(recs.data(), d_plan, raw_bytes, cudaMemcpyDeviceToHost, h->s_compute);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(cubic_blocks.data(), d_cubic_blocks, sizeof(bs::CubicCoef) * (size_t)n_blocks, cudaMemcpyDeviceToHost, h->s_compute);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->s_compute);
    cudaFree(d_cubic_blocks);
    if (e != cudaSuccess) {
        cudaFree(d_plan);
        return fail(h, BS_ERR_CUDA, "lens plan failed: %s", cudaGetErrorString(e));
    }
    h->kernel_launches++;
    std::vector<bs::WorkItem> items((size_t)n_blocks * bs::kMaxLayers);
    PlanLayout L;
    layout_plan(recs.data(), n_blocks, 1, items.data(), &L, true);
    for (int i = 0; i < L.n_items; ++i) items[i].cam = (uint8_t)cam;
    if (getenv("BS_DEBUG_PLAN_SUMMARY")) {
        long flagged = 0, direct = 0, plain = 0;
        for (int i = 0; i < L.n_items; ++i)
            if (items[i].type == bs::kBlockFast) {
                flagged += items[i].groups != 0;
                direct += (items[i].flags & bs::kItemDirect) != 0;
                plain += (items[i].flags & bs::kItemPlain) != 0;
            }
        fprintf(stderr, "[lens plan] cam %d: blocks %d: fast %lld exact %lld empty %lld; compact %d; flagged %ld direct %ld plain %ld; box rows %d\n",
                cam, n_blocks, (long long)L.n_fast, (long long)L.n_exact, (long long)L.n_empty, (int)L.compact, flagged, direct, plain,
                L.box_h[0]);
    }
    int optin = 0, sms = 0, per_sm = 0;
    const int smem = bs::kWarps2 * bs::kWarp2SmemLens + 128;
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    if (!L.compact || L.n_fast == 0 || smem > optin ||
        cudaFuncSetAttribute(bs::stitch_fast2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast2_kernel<true>, bs::kFast2Threads, smem) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        cudaFree(d_plan);
        return BS_OK;
    }
    m.grid = std::max(1, std::min((n_blocks + bs::kWarps2 - 1) / bs::kWarps2, per_sm * sms));
    m.smem_bytes = smem;
    m.box_h = L.box_h[0];
    const int n_warps = m.grid * bs::kWarps2;
    std::vector<bs::WorkItem> grouped(items.size());
    std::vector<int> warp_start(n_warps + 1);
    group_items_by_warp(items.data(), L.n_items - L.n_two, L.n_two, n_warps, grouped.data(), warp_start.data());
    const size_t ws_bytes = sizeof(int) * (size_t)(n_warps + 1), cub_bytes = sizeof(bs::CubicCoef) * (size_t)L.n_items;
    std::vector<bs::CubicCoef> cubic_items(L.n_items);       // the third-order terms in item order
    for (int k = 0; k < L.n_items; ++k) cubic_items[k] = cubic_blocks[(size_t)grouped[k].by * blocks_x + grouped[k].bx];
    e = upload_now(h, d_plan, grouped.data(), sizeof(bs::WorkItem) * (size_t)L.n_items);     // <= raw_bytes (at most two items per block)
    if (e == cudaSuccess) e = cudaMalloc(&m.d_warp_start, ws_bytes);
    if (e == cudaSuccess) e = upload_now(h, m.d_warp_start, warp_start.data(), ws_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&m.d_cubic, cub_bytes);
    if (e == cudaSuccess) e = upload_now(h, m.d_cubic, cubic_items.data(), cub_bytes);
    if (e != cudaSuccess) {
        cudaFree(d_plan);
        cudaFree(m.d_warp_start);
        cudaFree(m.d_cubic);
        m.d_warp_start = nullptr;
        m.d_cubic = nullptr;
        return fail(h, BS_ERR_CUDA, "lens plan upload failed: %s", cudaGetErrorString(e));
    }
    m.d_items = static_cast<bs::WorkItem *>(d_plan);
    m.bytes += raw_bytes + ws_bytes + cub_bytes;
    h->device_bytes += raw_bytes + ws_bytes + cub_bytes;
    items.resize(L.n_items);
    m.items_host.swap(items);
    m.cubic_blocks.swap(cubic_blocks);
    m.n_items = L.n_items;
    m.n_two = L.n_two;
    m.blocks_x = blocks_x;
    m.plan_ok = true;
    return BS_OK;
}

void clear_lens_set(bs_handle *h)
{
    LensSet &S = h->lens_set;
    cudaFree(S.d_items);
    cudaFree(S.d_cubic);
    cudaFree(S.d_warp_start);
    h->device_bytes -= S.bytes;
    S = LensSet();
}

// Deal the lens plans of all cameras that have one to the warps of one launch (see LensSet).
int build_lens_set(bs_handle *h)
{
    clear_lens_set(h);
    LensSet &S = h->lens_set;
    static const bool never = getenv("BS_DEBUG_NO_LENS_SET") != nullptr;     // test hook: one launch per camera
    if (never) return BS_OK;
    // Only the part of an undistorted frame that the panorama plan can read has to exist: when that plan knows the
    // source rectangle of every camera (no per-pixel blocks, which gather anywhere), the lens blocks outside it are
    // left out of the launch -- at cfg3 a camera wins 47 % of its frame, so half of the lens work disappears.  (The
    // frame kernel's TMA boxes reach a little beyond the rectangle; what they load from there is never used.)
    static const bool no_trim = getenv("BS_DEBUG_NO_LENS_TRIM") != nullptr;        // test hook: whole frames
    const bool trim = h->fast_ok && !h->up_x0.empty() && !no_trim;
    auto needed = [&](int c, const bs::WorkItem &it) {
        if (!trim) return true;
        if (h->up_x1[c] < 0) return false;
        const int x0 = it.bx * bs::kBlockW, y0 = it.by * bs::kBlockH;
        return x0 <= h->up_x1[c] && x0 + bs::kBlockW - 1 >= h->up_x0[c] && y0 <= h->up_y1[c] && y0 + bs::kBlockH - 1 >= h->up_y0[c];
    };
    const int n_two = 0;        // a lens plan has one camera: no two-layer blocks
    std::vector<bs::WorkItem> in;
    S.src_x0.assign(h->n_cams, -1);
    S.src_y0.assign(h->n_cams, -1);
    S.src_x1.assign(h->n_cams, -1);
    S.src_y1.assign(h->n_cams, -1);
    for (int c = 0; c < h->n_cams && c < bs::kMaxCamsPerLaunch; ++c)
        if (h->lens[c].enabled && h->lens[c].plan_ok) {
            const LensModel &m = h->lens[c];
            if (m.n_two != 0) {
                S = LensSet();
                return BS_OK;
            }
            // the distorted-frame rectangle the kept blocks read (their whole TMA boxes): what bs_submit must upload
            int sx0 = INT_MAX, sy0 = INT_MAX, sx1 = -1, sy1 = -1;
            bool any = false, gathers = false;
            for (const bs::WorkItem &it : m.items_host) {
                if (!needed(c, it)) continue;
                in.push_back(it);
                any = true;
                if (it.type == bs::kBlockFast) {
                    sx0 = std::min(sx0, 4 * (int)it.c0 / 3);
                    sx1 = std::max(sx1, (4 * (int)it.c0 + bs::kRawPitch - 1) / 3);
                    sy0 = std::min(sy0, (int)it.y0);
                    sy1 = std::max(sy1, (int)it.y0 + m.box_h - 1);
                } else if (it.type == bs::kBlockExact) {
                    gathers = true;          // evaluated tap by tap from anywhere in the frame
                }
            }
            if (!any) continue;              // nothing of this camera's frame is ever read
            S.cams.push_back(c);
            S.box_h = std::max(S.box_h, m.box_h);
            S.smem_bytes = m.smem_bytes;
            if (trim && !gathers && sx1 >= 0) {
                S.src_x0[c] = std::max(0, sx0);
                S.src_y0[c] = std::max(0, sy0);
                S.src_x1[c] = std::min(h->sw[c] - 1, sx1);
                S.src_y1[c] = std::min(h->sh[c] - 1, sy1);
            }
        }
    const int n_items = (int)in.size(), n_units = n_items, n_blocks = n_items;
    if (S.cams.empty() || (S.cams.size() < 2 && !trim)) {
        S = LensSet();
        return BS_OK;
    }
    S.trimmed = trim;
    int sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast2_kernel<true>, bs::kFast2Threads, S.smem_bytes) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        S = LensSet();
        return BS_OK;
    }
    S.grid = std::max(1, std::min((n_blocks + bs::kWarps2 - 1) / bs::kWarps2, per_sm * sms));
    const int n_warps = S.grid * bs::kWarps2;
    std::vector<bs::WorkItem> grouped(in.size());
    std::vector<int> warp_start(n_warps + 1);
    group_items_by_warp(in.data(), n_units, n_two, n_warps, grouped.data(), warp_start.data());
    std::vector<bs::CubicCoef> cubic(n_items);
    for (int k = 0; k < n_items; ++k) {
        const LensModel &m = h->lens[grouped[k].cam];
        cubic[k] = m.cubic_blocks[(size_t)grouped[k].by * m.blocks_x + grouped[k].bx];
    }
    const size_t it_bytes = sizeof(bs::WorkItem) * (size_t)n_items, ws_bytes = sizeof(int) * (size_t)(n_warps + 1),
                 cub_bytes = sizeof(bs::CubicCoef) * (size_t)n_items;
    cudaError_t e = cudaMalloc(&S.d_items, it_bytes);
    if (e == cudaSuccess) e = upload_now(h, S.d_items, grouped.data(), it_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&S.d_warp_start, ws_bytes);
    if (e == cudaSuccess) e = upload_now(h, S.d_warp_start, warp_start.data(), ws_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&S.d_cubic, cub_bytes);
    if (e == cudaSuccess) e = upload_now(h, S.d_cubic, cubic.data(), cub_bytes);
    if (e != cudaSuccess) {
        cudaFree(S.d_items);
        cudaFree(S.d_warp_start);
        cudaFree(S.d_cubic);
        S = LensSet();
        return fail(h, BS_ERR_CUDA, "lens set upload failed: %s", cudaGetErrorString(e));
    }
    S.bytes = it_bytes + ws_bytes + cub_bytes;
    h->device_bytes += S.bytes;
    S.valid = true;
    return BS_OK;
}
}  // namespace

extern "C" {

int bs_set_undistort(bs_handle *h, int cam, const double K[9], const double *dist, int n_dist)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_set_undistort: NULL handle");
    if (cam < 0 || cam >= h->n_cams) return fail(h, BS_ERR_INVALID_ARG, "bs_set_undistort: camera %d not in [0, %d)", cam, h->n_cams);
    if (n_dist < 0 || n_dist > 8 || (n_dist > 0 && !dist))
        return fail(h, BS_ERR_UNSUPPORTED, "bs_set_undistort: %d distortion coefficients (0..8 supported: k1 k2 p1 p2 k3 k4 k5 k6)", n_dist);
    BS_CUDA(h, cudaSetDevice(h->device));
    if (h->lens.empty()) h->lens.resize(h->n_cams);
    LensModel &m = h->lens[cam];
    flush_deferred_jpeg(h, h->next_ticket - 1);          // frames submitted before this call use the old model
    BS_CUDA(h, cudaStreamSynchronize(h->s_compute));     // frames in flight may still be reading the old tables
    clear_lens(h, m);
    bool nonzero = false;
    for (int i = 0; i < n_dist; ++i) {
        m.k[i] = dist[i];
        nonzero = nonzero || dist[i] != 0.0;
    }
    if (!K || !nonzero) return build_lens_set(h);     // cv::undistort with a zero model is an exact copy: nothing to run
    std::vector<double> xs, ys;
    if (build_lens_tables(h->sw[cam], h->sh[cam], K, m, xs, ys) != BS_OK)
        return fail(h, BS_ERR_UNSUPPORTED, "bs_set_undistort: camera matrix must be [[fx 0 cx] [0 fy cy] [0 0 1]]");
    int *d_overflow = nullptr, overflow = 0;
    const size_t tab_bytes = (xs.size() + ys.size()) * sizeof(double);
    const size_t map_bytes = sizeof(uint32_t) * (size_t)h->sw[cam] * (size_t)h->sh[cam];
    const bool want_map = !(h->flags & BS_FLAG_EXACT_ONLY);      // one-shot handles (bs_undistort): one pass, no map
    cudaError_t e = cudaMalloc(&m.d_xs, xs.size() * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&m.d_ys, ys.size() * sizeof(double));
    if (e == cudaSuccess) e = upload_now(h, m.d_xs, xs.data(), xs.size() * sizeof(double));
    if (e == cudaSuccess) e = upload_now(h, m.d_ys, ys.data(), ys.size() * sizeof(double));
    if (e == cudaSuccess) {
        m.bytes = tab_bytes;
        h->device_bytes += tab_bytes;
    }
    // the model constants as the lens launches read them (per camera, in device memory: the launch carries pointers only)
    if (e == cudaSuccess && !h->d_lens_models) {
        e = cudaMalloc(&h->d_lens_models, sizeof(bs::UndistortArgs) * (size_t)h->n_cams);
        if (e == cudaSuccess) h->device_bytes += sizeof(bs::UndistortArgs) * (size_t)h->n_cams;
    }
    if (e == cudaSuccess) {
        const bs::UndistortArgs U = lens_args(h, m, cam);
        e = upload_now(h, h->d_lens_models + cam, &U, sizeof U);
    }
    // where this camera's undistorted frames go: the device-submit buffer now, the ring's slots if they exist
    int rc = BS_OK;
    if (e == cudaSuccess) rc = alloc_undist(h, h->d_undist, cam);
    for (size_t k = 0; rc == BS_OK && e == cudaSuccess && k < h->ring.size(); ++k) rc = alloc_undist(h, h->ring[k].d_undist, cam);
    // The lens field depends on the calibration only.  Preferred form: a block plan for the compact frame kernel.
    if (e == cudaSuccess && rc == BS_OK && want_map && !getenv("BS_DEBUG_NO_LENS_PLAN")) rc = build_lens_plan(h, m, cam);
    if (e == cudaSuccess && rc == BS_OK && want_map && !m.plan_ok) {
        // Otherwise: one packed map word per pixel, evaluated once, every frame resampled from it.
        e = cudaMalloc(&m.d_map, map_bytes);
        if (e == cudaSuccess) e = cudaMalloc(&d_overflow, sizeof(int));
        if (e == cudaSuccess) e = cudaMemsetAsync(d_overflow, 0, sizeof(int), h->s_compute);
        if (e == cudaSuccess && h->remap_grid == 0) {
            int per_sm = 0, sms = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::undistort_remap_kernel, 256, 0);
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
            h->remap_grid = std::max(1, per_sm * sms);
        }
        if (e == cudaSuccess) {
            const bs::UndistortArgs U = lens_args(h, m, cam);
            bs::undistort_map_kernel<<<dim3((U.w + 63) / 64, (U.h + 3) / 4), 256, 0, h->s_compute>>>(U, m.d_map, d_overflow);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(&overflow, d_overflow, sizeof(int), cudaMemcpyDeviceToHost, h->s_compute);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->s_compute);
        cudaFree(d_overflow);
        if (e == cudaSuccess && !overflow) {
            m.bytes += map_bytes;
            h->device_bytes += map_bytes;
        } else {                                             // displacement beyond the packed range: keep the one-pass kernel
            cudaFree(m.d_map);
            m.d_map = nullptr;
        }
    }
    if (e != cudaSuccess || rc != BS_OK) {
        clear_lens(h, m);                                    // never leave a half-built model behind
        build_lens_set(h);
        if (e != cudaSuccess) return fail(h, BS_ERR_CUDA, "bs_set_undistort: %s", cudaGetErrorString(e));
        return rc;
    }
    m.enabled = true;
    return build_lens_set(h);
}

int bs_undistort_device(bs_handle *h, int cam, const void *d_src, size_t spitch, void *d_dst, size_t dpitch, void *stream,
                        unsigned submit_flags)
{
    if (h && (submit_flags & ~BS_SUBMIT_SOURCES_SETTLED))
        return fail(h, BS_ERR_INVALID_ARG, "bs_undistort_device: unknown submit flags 0x%x", submit_flags);
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_undistort_device: NULL handle");
    if (cam < 0 || cam >= h->n_cams || !d_src || !d_dst || d_src == d_dst)
        return fail(h, BS_ERR_INVALID_ARG, "bs_undistort_device: bad camera index or buffers");
    if (spitch < (size_t)h->sw[cam] * 3 || dpitch < (size_t)h->sw[cam] * 3)
        return fail(h, BS_ERR_INVALID_ARG, "bs_undistort_device: pitch below %d*3", h->sw[cam]);
    int prev_dev = -1;
    BS_CUDA(h, cudaGetDevice(&prev_dev));
    if (prev_dev != h->device) BS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = BS_OK;
    if (h->lens.empty() || !h->lens[cam].enabled) {          // zero model: cv::undistort is an exact copy
        if (cudaMemcpy2DAsync(d_dst, dpitch, d_src, spitch, (size_t)h->sw[cam] * 3, h->sh[cam], cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            rc = fail(h, BS_ERR_CUDA, "bs_undistort_device: copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    } else {
        rc = undistort_one(h, cam, d_src, spitch, static_cast<uint8_t *>(d_dst), dpitch, st, !(submit_flags & BS_SUBMIT_SOURCES_SETTLED));
    }
    if (prev_dev != h->device) cudaSetDevice(prev_dev);
    return rc;
}

int bs_undistort(const uint8_t *src, int w, int h, size_t spitch, uint8_t *dst, size_t dpitch, const double K[9],
                 const double *dist, int n_dist)
{
    if (!src || !dst || !K) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_undistort: NULL argument");
    if (w < 1 || h < 1 || w > BS_MAX_DIM || h > BS_MAX_DIM || spitch < (size_t)w * 3 || dpitch < (size_t)w * 3)
        return fail(nullptr, BS_ERR_INVALID_ARG, "bs_undistort: bad size %dx%d or pitch", w, h);
    int rc = require_device(nullptr);
    if (rc != BS_OK) return rc;
    // a one-camera handle whose canvas is never used: it owns the tables and the buffers
    const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    bs_handle *hd = nullptr;
    rc = bs_create(g_last_device, 1, &w, &h, 32, 16, I, BS_FLAG_EXACT_ONLY, &hd);
    if (rc != BS_OK) return rc;
    rc = bs_set_undistort(hd, 0, K, dist, n_dist);
    uint8_t *d_src = nullptr;
    const size_t pitch = hd->d_frame_pitch[0];
    cudaError_t e = cudaSuccess;
    if (rc == BS_OK) {
        e = cudaMalloc(&d_src, pitch * (size_t)h);
        if (e == cudaSuccess) e = cudaMemcpy2DAsync(d_src, pitch, src, spitch, (size_t)w * 3, h, cudaMemcpyHostToDevice, hd->s_compute);
        const void *in[1] = {d_src};
        const void *out[1] = {nullptr};
        size_t out_pitch[1] = {0};
        if (e == cudaSuccess) rc = undistort_frames(hd, in, &pitch, hd->d_undist, out, out_pitch, hd->s_compute, true);
        if (e == cudaSuccess && rc == BS_OK) e = cudaStreamSynchronize(hd->s_compute);
        if (e == cudaSuccess && rc == BS_OK)
            e = cudaMemcpy2D(dst, dpitch, out[0], out_pitch[0], (size_t)w * 3, h, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(nullptr, BS_ERR_CUDA, "bs_undistort: %s", cudaGetErrorString(e));
    } else {
        memcpy(g_last_error, hd->err, kErrLen);
    }
    cudaFree(d_src);
    bs_destroy(hd);
    return rc;
}

int bs_warp_perspective(const uint8_t *src, int sw, int sh, size_t spitch, uint8_t *dst, int dw, int dh,
                        size_t dpitch, const double H[9], unsigned flags)
{
    if (!src || !dst || !H) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_warp_perspective: NULL argument");
    if (flags & BS_FLAG_FEATHER) return fail(nullptr, BS_ERR_UNSUPPORTED, "bs_warp_perspective: BS_FLAG_FEATHER needs a handle with all cameras");
    int rc = require_device(nullptr);
    if (rc != BS_OK) return rc;
    bs_handle *h = nullptr;
    // BORDER_TRANSPARENT is read-modify-write on dst; BORDER_CONSTANT overwrites all of it
    rc = bs_create(g_last_device, 1, &sw, &sh, dw, dh, H, (flags & BS_FLAG_BORDER_CONSTANT) ? flags : (flags | BS_FLAG_KEEP_CANVAS), &h);
    if (rc != BS_OK) return rc;
    const uint8_t *frames[1] = {src};
    size_t pitches[1] = {spitch};
    rc = bs_stitch(h, frames, pitches, dst, dpitch);
    if (rc != BS_OK) memcpy(g_last_error, h->err, kErrLen);
    bs_destroy(h);
    return rc;
}

void *bs_host_alloc(size_t bytes)
{
    void *p = nullptr;
    // BS_HOST_ALLOC_WC=1 (experiment hook): write-combined pages -- fast for the host to fill and for
    // the device to read, very slow for the host to read back, so only for buffers that are uploaded
    static const bool wc = getenv("BS_HOST_ALLOC_WC") != nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable | (wc ? cudaHostAllocWriteCombined : 0)) != cudaSuccess) {
        cudaGetLastError();
        fail(nullptr, BS_ERR_NOMEM, "bs_host_alloc: cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}

void bs_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int bs_get_stats(const bs_handle *h, bs_stats *out)
{
    if (!h || !out) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_get_stats: NULL argument");
    memset(out, 0, sizeof *out);
    out->n_cams = h->n_cams;
    out->canvas_w = h->W;
    out->canvas_h = h->H;
    if (h->fast_ok) {
        out->tile_w = bs::kBlockW;
        out->tile_h = bs::kBlockH;
        out->n_tiles = h->n_blocks;
        out->n_tiles_fast = h->n_fast;
        out->n_tiles_exact = h->n_exact;
        out->n_tiles_empty = h->n_empty;
    } else {
        out->tile_w = bs::kExactTileW;
        out->tile_h = bs::kExactTileH;
        out->n_tiles = (int64_t)((h->W + bs::kExactTileW - 1) / bs::kExactTileW) *
                       ((h->H + bs::kExactTileH - 1) / bs::kExactTileH);
        out->n_tiles_exact = out->n_tiles;
    }
    out->kernel_launches = h->kernel_launches;
    out->frames_submitted = h->next_ticket;
    out->device_bytes = h->device_bytes;
    out->scratch_bytes = h->scratch_bytes;
    out->h2d_bytes_last_submit = h->h2d_bytes_last;
    out->frame_kernel = !h->fast_ok ? 0 : (h->compact ? 2 : 1);
    return BS_OK;
}

}  // extern "C"

// ================================================================================================
// bs_stitch_device_static -- one panorama frame without allocating anything.
//
// Who this is for: callers that hold every buffer themselves and are not allowed to allocate while
// they run -- a capture/render loop with a fixed memory budget, code running under an allocator that
// forbids cudaMalloc after start-up (cudaMalloc also synchronises the whole device, which a frame loop
// cannot afford), or a process whose memory was handed to it by a host application.  bs_create and
// bs_submit allocate (plan, rings, streams); this entry point does not call cudaMalloc, cudaFree,
// cudaHostAlloc, or any host heap routine (error texts go to fixed per-thread buffers).
//
// The staged kernel needs scratch memory: the block plan and the per-warp work split.  Here they live
// in ONE file-scope device array of a fixed size.  A call whose calibration arguments (cameras, sizes, homographies,
// flags) equal those of the previous call on the same device finds its plan still there and only launches the frame
// kernel; any other call rebuilds the plan in full.  Consequences a caller must know:
//   * every GPU that loads this library holds the array (BS_STATIC_SCRATCH_BYTES) until the process exits;
//   * canvases that need more than BS_STATIC_MAX_BLOCKS blocks of 32x16 pixels are refused;
//   * all calls share the array and it keeps the previous call's contents.  A call with other calibration
//     arguments rewrites every record and the whole work split (nothing stale is ever read; the remembered
//     arguments are compared bit for bit, and only a call that completed is remembered), so two
//     calls whose GPU work overlaps (different streams, or different host threads on different
//     streams) race on it unless their calibrations are equal.  Issue calls on one stream, or order the
//     streams yourself.
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress) and
//     sets the kernel's shared-memory limit; make that first call before any stream capture begins.
// ================================================================================================
#include <mutex>

namespace {

constexpr int kStaticMaxWarps = 8191;   // warps of the grid the table below has room for (B200: 148 x 2 x 8 = 2368)
struct StaticScratch {
    // raw block records from the plan kernel; then overwritten by the work items (at most 2 x 80 bytes per block)
    bs::BlockRec plan[BS_STATIC_MAX_BLOCKS];
    int warp_start[kStaticMaxWarps + 1];
};
static_assert(sizeof(StaticScratch) == BS_STATIC_SCRATCH_BYTES, "BS_STATIC_SCRATCH_BYTES is out of date");
static_assert(sizeof(bs::WorkItem) * bs::kMaxLayers <= sizeof(bs::BlockRec), "items must fit where the raw records were");

// The scratch itself.  In host code this name is only a stand-in for the device object: its device
// address comes from cudaGetSymbolAddress (static_device_state), never from &g_static_scratch.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

// What the plan in a device's scratch was built from.  A call with the same calibration arguments as the previous
// call on that device (a video loop) finds its plan and work split already there and only launches the frame kernel.
struct StaticKey {
    int n_cams, W, H;
    unsigned flags;
    int sw[bs::kMaxCamsPerLaunch], sh[bs::kMaxCamsPerLaunch];
    double Hm[bs::kMaxCamsPerLaunch * 9];
};
struct StaticDeviceState {
    bool ready = false;
    StaticScratch *scratch = nullptr;   // device address of g_static_scratch on this device
    int sms = 0;
    int smem_optin = 0;
    // the plan the scratch holds (valid only after a staged call completed)
    bool plan_valid = false;
    StaticKey key;
    PlanLayout L;
    int grid = 0, rec_slots = 0, per_sm = 0;
};
constexpr int kStaticMaxDevices = 64;
StaticDeviceState g_static_dev[kStaticMaxDevices];
std::mutex g_static_mutex;                            // host side of the sharing; the GPU side is the caller's job
bs::BlockRec g_static_recs[BS_STATIC_MAX_BLOCKS];     // host staging for the plan, fixed size (no heap)
bs::WorkItem g_static_items[BS_STATIC_MAX_BLOCKS * bs::kMaxLayers];
bs::WorkItem g_static_grouped[BS_STATIC_MAX_BLOCKS * bs::kMaxLayers];
int g_static_warp_start[kStaticMaxWarps + 1];
bs::PlanArgs g_static_plan_args;

// Once per device, with that device current.
int static_device_state(int dev, StaticDeviceState **out)
{
    if (dev < 0 || dev >= kStaticMaxDevices)
        return fail(nullptr, BS_ERR_UNSUPPORTED, "bs_stitch_device_static: device ordinal %d not supported", dev);
    StaticDeviceState &st = g_static_dev[dev];
    if (!st.ready) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(nullptr, BS_ERR_CUDA, "bs_stitch_device_static: cudaGetSymbolAddress failed: %s",
                        cudaGetErrorString(e));
        st.scratch = static_cast<StaticScratch *>(p);
        BS_CUDA(nullptr, cudaDeviceGetAttribute(&st.sms, cudaDevAttrMultiProcessorCount, dev));
        BS_CUDA(nullptr, cudaDeviceGetAttribute(&st.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        BS_CUDA(nullptr, cudaFuncSetAttribute(bs::stitch_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              st.smem_optin));
        st.ready = true;
    }
    *out = &st;
    return BS_OK;
}

}  // namespace

extern "C" int bs_stitch_device_static(int n_cams, const int *src_w, const int *src_h, const void *const *d_frames,
                                       const size_t *pitch, int canvas_w, int canvas_h, void *d_canvas,
                                       size_t canvas_pitch, const double *H, unsigned flags, void *stream)
{
    const char *me = "bs_stitch_device_static";
    if (!src_w || !src_h || !d_frames || !pitch || !d_canvas || !H)
        return fail(nullptr, BS_ERR_INVALID_ARG, "%s: NULL argument", me);
    if (n_cams < 1 || n_cams > bs::kMaxCamsPerLaunch)
        return fail(nullptr, BS_ERR_UNSUPPORTED, "%s: n_cams=%d outside [1, %d]", me, n_cams, bs::kMaxCamsPerLaunch);
    if (canvas_w < 1 || canvas_h < 1 || canvas_w > 65535 || canvas_h > 65535)
        return fail(nullptr, BS_ERR_INVALID_ARG, "%s: canvas %dx%d outside [1, 65535]", me, canvas_w, canvas_h);
    for (int i = 0; i < n_cams; ++i) {
        if (src_w[i] < 1 || src_h[i] < 1 || src_w[i] > BS_MAX_DIM || src_h[i] > BS_MAX_DIM)
            return fail(nullptr, BS_ERR_INVALID_ARG, "%s: camera %d is %dx%d, outside [1, %d]", me, i, src_w[i], src_h[i],
                        BS_MAX_DIM);
        if (!d_frames[i] || pitch[i] < (size_t)src_w[i] * 3)
            return fail(nullptr, BS_ERR_INVALID_ARG, "%s: frame %d is NULL or its pitch is below %d*3", me, i, src_w[i]);
    }
    if (canvas_pitch < (size_t)canvas_w * 3)
        return fail(nullptr, BS_ERR_INVALID_ARG, "%s: canvas_pitch=%zu < %d*3", me, canvas_pitch, canvas_w);
    if (flags & ~(BS_FLAG_LEGACY_BORDER | BS_FLAG_KEEP_CANVAS))
        return fail(nullptr, BS_ERR_UNSUPPORTED, "%s: flags 0x%x: only BS_FLAG_LEGACY_BORDER and BS_FLAG_KEEP_CANVAS are accepted here",
                    me, flags);
    const int blocks_x = (canvas_w + bs::kBlockW - 1) / bs::kBlockW, blocks_y = (canvas_h + bs::kBlockH - 1) / bs::kBlockH;
    const long long n_blocks_ll = (long long)blocks_x * blocks_y;
    if (n_blocks_ll > BS_STATIC_MAX_BLOCKS)
        return fail(nullptr, BS_ERR_UNSUPPORTED,
                    "%s: a %dx%d canvas needs %lld blocks of %dx%d pixels, the static scratch holds %d; use bs_create", me,
                    canvas_w, canvas_h, n_blocks_ll, bs::kBlockW, bs::kBlockH, BS_STATIC_MAX_BLOCKS);
    const int n_blocks = (int)n_blocks_ll;
    int rc = require_device(nullptr);
    if (rc != BS_OK) return rc;
    int dev = 0;
    BS_CUDA(nullptr, cudaGetDevice(&dev));

    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticDeviceState *st = nullptr;
    rc = static_device_state(dev, &st);
    if (rc != BS_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool keep = (flags & BS_FLAG_KEEP_CANVAS) != 0;
    const int legacy = (flags & BS_FLAG_LEGACY_BORDER) ? 1 : 0;
    const int block_w = opencv_block_width(canvas_w, canvas_h);

    // ---- the plan: still in the scratch from the previous call, or rebuilt in full ---------------------
    static StaticKey key;                       // (under g_static_mutex; static: no large object on the stack)
    memset(&key, 0, sizeof key);
    key.n_cams = n_cams;
    key.W = canvas_w;
    key.H = canvas_h;
    key.flags = flags;
    for (int i = 0; i < n_cams; ++i) {
        key.sw[i] = src_w[i];
        key.sh[i] = src_h[i];
    }
    memcpy(key.Hm, H, sizeof(double) * 9 * (size_t)n_cams);
    const bool hit = st->plan_valid && memcmp(&key, &st->key, sizeof key) == 0;

    bs::FastArgs FA;
    memset(&FA, 0, sizeof FA);
    bs::FrameArgs &A = FA.F;
    A.n_cams = n_cams;
    for (int i = 0; i < n_cams; ++i) {
        invert3x3(H + 9 * i, A.cam[i].M);
        A.cam[i].ptr = static_cast<const uint8_t *>(d_frames[i]);
        A.cam[i].pitch = (long long)pitch[i];
        A.cam[i].w = src_w[i];
        A.cam[i].h = src_h[i];
    }
    A.canvas = static_cast<uint8_t *>(d_canvas);
    A.cpitch = (long long)canvas_pitch;
    A.W = canvas_w;
    A.H = canvas_h;
    A.block_w = block_w;
    A.legacy_border = legacy;
    A.keep_canvas = keep ? 1 : 0;
    A.canvas_in_4g = canvas_in_4g(d_canvas, canvas_pitch, canvas_h);

    PlanLayout &L = st->L;
    if (!hit) {
        // every record is rewritten: nothing stale survives
        st->plan_valid = false;
        bs::PlanArgs &P = g_static_plan_args;
        memset(&P, 0, sizeof P);
        for (int i = 0; i < n_cams; ++i) {
            memcpy(P.M[i], A.cam[i].M, 9 * sizeof(double));
            P.w[i] = src_w[i];
            P.h[i] = src_h[i];
        }
        P.n_cams = n_cams;
        P.W = canvas_w;
        P.H = canvas_h;
        P.block_w = block_w;
        P.legacy_border = legacy;
        P.keep_canvas = keep ? 1 : 0;
        P.blocks_x = blocks_x;
        P.blocks_y = blocks_y;
        P.out = st->scratch->plan;
        bs::plan_kernel<<<dim3((blocks_x + bs::kPlanWarps - 1) / bs::kPlanWarps, blocks_y), 32 * bs::kPlanWarps, 0, s>>>(P);
        BS_CUDA(nullptr, cudaGetLastError());
        const size_t raw_bytes = sizeof(bs::BlockRec) * (size_t)n_blocks;
        BS_CUDA(nullptr, cudaMemcpyAsync(g_static_recs, st->scratch->plan, raw_bytes, cudaMemcpyDeviceToHost, s));
        BS_CUDA(nullptr, cudaStreamSynchronize(s));
        // size one TMA box per camera, flatten to work items
        L = PlanLayout();
        layout_plan(g_static_recs, n_blocks, n_cams, g_static_items, &L, false);
        st->per_sm = 0;
        st->rec_slots = bs::kRecSlotsMax;
        if (L.n_fast > 0 && get_encoder() && cta_smem_bytes(L.stage_bytes, bs::kRecSlotsMax / 2) <= st->smem_optin)
            st->per_sm = choose_rec_slots(L.stage_bytes, &st->rec_slots, &L.smem_bytes);
    }
    bool fast = st->per_sm >= 1 && aligned16(d_canvas, canvas_pitch);
    for (int i = 0; fast && i < n_cams; ++i) fast = aligned16(d_frames[i], pitch[i]);

    if (!fast) {
        // no staged blocks (or unaligned buffers): the per-pixel kernel needs no scratch at all
        dim3 grid((canvas_w + bs::kExactTileW - 1) / bs::kExactTileW, (canvas_h + bs::kExactTileH - 1) / bs::kExactTileH);
        bs::stitch_exact_kernel<<<grid, dim3(32, bs::kExactTileH), 0, s>>>(A);
        BS_CUDA(nullptr, cudaGetLastError());
        BS_CUDA(nullptr, cudaStreamSynchronize(s));
        return BS_OK;
    }

    for (int i = 0; i < n_cams; ++i) {
        FA.pitch[i] = L.box_we[i] * 4;
        FA.box_bytes[i] = L.box_we[i] * 4 * L.box_h[i];
        if (L.box_we[i] > 0) {
            const int er = encode_frame_map(&FA.tmap[i], d_frames[i], pitch[i], src_h[i], L.box_we[i], L.box_h[i]);
            if (er != 0) return fail(nullptr, BS_ERR_CUDA, "%s: cuTensorMapEncodeTiled failed (%d) for camera %d", me, er, i);
        }
    }
    if (!hit) {
        // split the items over the grid's warps; the item list replaces the raw records in the scratch,
        // stream-ordered after the previous call's kernels
        int grid = std::max(1, std::min((n_blocks + bs::kWarpsPerCta - 1) / bs::kWarpsPerCta, st->per_sm * st->sms));
        grid = std::min(grid, kStaticMaxWarps / bs::kWarpsPerCta);
        st->grid = grid;
        const int n_warps = grid * bs::kWarpsPerCta;
        group_items_by_warp(g_static_items, L.n_items - L.n_two, L.n_two, n_warps, g_static_grouped, g_static_warp_start);
        BS_CUDA(nullptr, cudaMemcpyAsync(st->scratch->plan, g_static_grouped, sizeof(bs::WorkItem) * (size_t)L.n_items,
                                         cudaMemcpyHostToDevice, s));
        BS_CUDA(nullptr, cudaMemcpyAsync(st->scratch->warp_start, g_static_warp_start, sizeof(int) * (size_t)(n_warps + 1),
                                         cudaMemcpyHostToDevice, s));
    }
    FA.items = reinterpret_cast<const bs::WorkItem *>(st->scratch->plan);
    FA.warp_start = st->scratch->warp_start;
    FA.stage_bytes = L.stage_bytes;
    FA.rec_slots = st->rec_slots;
    bs::stitch_fast_kernel<<<st->grid, bs::kFastThreads, L.smem_bytes, s>>>(FA);
    BS_CUDA(nullptr, cudaGetLastError());
    // the host staging arrays are reused by the next call: everything queued above must have left them
    BS_CUDA(nullptr, cudaStreamSynchronize(s));
    if (!hit) {
        st->key = key;
        st->plan_valid = true;
    }
    return BS_OK;
}

#include "stitch_jpeg.inl"
