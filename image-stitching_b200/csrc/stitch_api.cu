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

#include <algorithm>

#include "stitch_fast.cuh"

namespace {

thread_local std::string g_last_error = "";
thread_local int g_last_device = 0;

constexpr int kPipelineDepth = 3;

struct Slot {
    std::vector<uint8_t *> d_frame;
    uint8_t *d_canvas = nullptr;
    cudaEvent_t ev_h2d = nullptr, ev_kernel = nullptr, ev_done = nullptr;
    int64_t ticket = -1;
    bool busy = false;
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
    // tile plan (static per calibration)
    bool fast_ok = false;                // the staged kernel can be used
    bs::TileRec *d_tiles = nullptr;      // plan, ordered: exact tiles, fast tiles, empty tiles
    unsigned *d_sched = nullptr;         // tile scheduler counters of the frame kernel
    int n_tiles = 0, tiles_x = 0, tiles_y = 0;
    int64_t n_fast = 0, n_exact = 0, n_empty = 0;
    int box_we[bs::kMaxCamsPerLaunch] = {0}, box_h[bs::kMaxCamsPerLaunch] = {0};
    int stage_bytes = 0, smem_bytes = 0, grid = 0;
    uint64_t device_bytes = 0, scratch_bytes = 0;
    std::string err;
};

namespace {

int fail(bs_handle *h, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    if (h) h->err = buf;
    return code;
}

#define BS_CUDA(h, call)                                                                      \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return fail(h, BS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                        __FILE__, __LINE__);                                                  \
    } while (0)

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

constexpr int kStageBytesMax = 32 * 1024;   // per-stage cap on a camera's source box

// Build the tile plan on the device and post-process it on the host (once per calibration).
int build_plan(bs_handle *h)
{
    h->fast_ok = false;
    h->tiles_x = (h->W + bs::kTileW - 1) / bs::kTileW;
    h->tiles_y = (h->H + bs::kTileH - 1) / bs::kTileH;
    h->n_tiles = h->tiles_x * h->tiles_y;
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
    P->tiles_x = h->tiles_x;
    P->tiles_y = h->tiles_y;
    const size_t bytes = sizeof(bs::TileRec) * (size_t)h->n_tiles;
    cudaError_t e = cudaMalloc(&h->d_tiles, bytes);
    if (e == cudaSuccess) {
        P->out = h->d_tiles;
        bs::plan_kernel<<<dim3(h->tiles_x, h->tiles_y), bs::kThreads, 0, h->s_compute>>>(*P);
        e = cudaGetLastError();
    }
    delete P;
    std::vector<bs::TileRec> recs(h->n_tiles);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(recs.data(), h->d_tiles, bytes, cudaMemcpyDeviceToHost, h->s_compute);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->s_compute);
    if (e != cudaSuccess) return fail(h, BS_ERR_CUDA, "tile plan failed: %s", cudaGetErrorString(e));
    h->kernel_launches++;
    h->device_bytes += bytes;
    h->scratch_bytes += bytes;

    // per-camera TMA box = largest source box among its tiles; tiles whose box is too big go exact
    int we[bs::kMaxCamsPerLaunch] = {0}, bh[bs::kMaxCamsPerLaunch] = {0};
    for (bs::TileRec &r : recs) {
        if (r.type != bs::kTileFast) continue;
        bool ok = true;
        int twe[bs::kMaxLayers], tbh[bs::kMaxLayers];
        for (int l = 0; l < r.n_layers; ++l) {
            const int e0 = bs::box_elem0(r.bx0[l]), e1 = (3 * r.bx1[l] + 2) >> 2;   // box starts 16-byte aligned
            twe[l] = (e1 - e0 + 1 + 3) & ~3;            // inner extent must be a multiple of 16 bytes
            tbh[l] = r.by1[l] - r.by0[l] + 1;
            if (twe[l] > 256 || tbh[l] > 256 || twe[l] * 4 * tbh[l] > kStageBytesMax) ok = false;
        }
        if (!ok) {
            r.type = bs::kTileExact;
            continue;
        }
        for (int l = 0; l < r.n_layers; ++l) {
            we[r.cam[l]] = std::max(we[r.cam[l]], twe[l]);
            bh[r.cam[l]] = std::max(bh[r.cam[l]], tbh[l]);
        }
    }
    // the per-camera maxima of width and height come from different tiles; re-check the product
    int stage = 0;
    for (int c = 0; c < h->n_cams; ++c) {
        if (we[c] * 4 * bh[c] > kStageBytesMax) {
            // shrink by sending this camera's largest-box tiles to the exact path
            for (bs::TileRec &r : recs) {
                if (r.type != bs::kTileFast) continue;
                for (int l = 0; l < r.n_layers; ++l)
                    if (r.cam[l] == c) r.type = bs::kTileExact;
            }
            we[c] = bh[c] = 0;
        }
        h->box_we[c] = we[c];
        h->box_h[c] = bh[c];
        stage = std::max(stage, we[c] * 4 * bh[c]);
    }
    h->stage_bytes = (int)align_up((size_t)stage + bs::kStagePad, 128);

    std::vector<bs::TileRec> ordered;
    ordered.reserve(recs.size());
    h->n_fast = h->n_exact = h->n_empty = 0;
    for (const bs::TileRec &r : recs)
        if (r.type == bs::kTileExact) ordered.push_back(r), h->n_exact++;
    for (int layers = bs::kMaxLayers; layers >= 1; --layers)   // heaviest tiles first
        for (const bs::TileRec &r : recs)
            if (r.type == bs::kTileFast && r.n_layers == layers) ordered.push_back(r), h->n_fast++;
    for (const bs::TileRec &r : recs)
        if (r.type == bs::kTileEmpty) ordered.push_back(r), h->n_empty++;
    BS_CUDA(h, cudaMemcpy(h->d_tiles, ordered.data(), bytes, cudaMemcpyHostToDevice));
    BS_CUDA(h, cudaMalloc(&h->d_sched, 2 * sizeof(unsigned)));
    BS_CUDA(h, cudaMemset(h->d_sched, 0, 2 * sizeof(unsigned)));   // [0] is set to the grid size below
    h->device_bytes += 8;
    h->scratch_bytes += 8;

    if (getenv("BS_DEBUG_PLAN_SUMMARY")) {
        long groups = 0, flagged_groups = 0, blocks = 0, flagged_blocks = 0;
        long per_cam[bs::kMaxCamsPerLaunch] = {0};
        for (const bs::TileRec &r : ordered) {
            if (r.type != bs::kTileFast) continue;
            for (int l = 0; l < r.n_layers; ++l) {
                for (int w = 0; w < 8; ++w) {
                    const unsigned g = (r.exact_mask[l] >> (4 * w)) & 15u;
                    blocks++;
                    groups += 4;
                    flagged_groups += __builtin_popcount(g);
                    if (g) flagged_blocks++, per_cam[r.cam[l]]++;
                }
            }
        }
        fprintf(stderr, "[plan] blocks %ld flagged %ld (%.2f%%), 32x4 groups %ld flagged %ld (%.2f%%)\n", blocks,
                flagged_blocks, 100.0 * flagged_blocks / (blocks ? blocks : 1), groups, flagged_groups,
                100.0 * flagged_groups / (groups ? groups : 1));
        for (int c = 0; c < h->n_cams; ++c) fprintf(stderr, "[plan]   cam %d flagged blocks %ld\n", c, per_cam[c]);
        int shown = 0;
        for (const bs::TileRec &r : ordered)
            for (int l = 0; l < r.n_layers && shown < 12; ++l)
                if (r.type == bs::kTileFast && r.exact_mask[l]) {
                    fprintf(stderr, "[plan]   tile (%d,%d) cam %d mask %08x\n", r.tx, r.ty, r.cam[l], r.exact_mask[l]);
                    shown++;
                }
    }
    if (getenv("BS_DEBUG_PLAN")) {
        for (int c = 0; c < h->n_cams; ++c)
            fprintf(stderr, "[plan] cam %d: %dx%d box %d elems x %d rows\n", c, h->sw[c], h->sh[c], h->box_we[c],
                    h->box_h[c]);
        for (const bs::TileRec &r : ordered) {
            fprintf(stderr, "[plan] tile (%d,%d) type %d layers %d unc %d:", r.tx, r.ty, r.type, r.n_layers,
                    r.has_uncovered);
            for (int l = 0; l < r.n_layers; ++l)
                fprintf(stderr, " [cam %d x %d..%d y %d..%d exact %x all %02x none %02x]", r.cam[l], r.bx0[l], r.bx1[l],
                        r.by0[l], r.by1[l], r.exact_mask[l], r.all_mask[l], r.none_mask[l]);
            fprintf(stderr, "\n");
        }
    }
    h->smem_bytes = bs::kStages * h->stage_bytes + bs::kConsumerWarps * bs::kWarpOutBytes + 2 * bs::kLutBytes + bs::kStages * bs::kMetaBytes + 64;   // + 8 KB slack to align the table
    {
        // The limit is a per-device property of the kernel, shared by every handle: always raise it to
        // the device's opt-in maximum, so a handle with a small plan cannot lower it for another one.
        int optin = 0;
        BS_CUDA(h, cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
        if (h->smem_bytes > optin) return BS_OK;   // does not fit: stay on the exact kernel
        BS_CUDA(h, cudaFuncSetAttribute(bs::stitch_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    }
    int per_sm = 0, sms = 0;
    BS_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast_kernel, bs::kFastThreads,
                                                             h->smem_bytes));
    BS_CUDA(h, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    if (per_sm < 1) return BS_OK;   // does not fit: stay on the exact kernel
    h->grid = std::min(h->n_tiles, per_sm * sms);
    if (const char *g = getenv("BS_DEBUG_GRID")) h->grid = std::max(1, std::min(h->n_tiles, atoi(g)));
    {   // the first tile of every CTA is static, so the dynamic counter starts at the grid size
        const unsigned first_dynamic = (unsigned)h->grid;
        BS_CUDA(h, cudaMemcpy(h->d_sched, &first_dynamic, sizeof first_dynamic, cudaMemcpyHostToDevice));
    }
    h->fast_ok = true;
    return BS_OK;
}

bool aligned16(const void *p, size_t pitch) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (pitch & 15) == 0; }

// Launch the kernels of one frame.  Cameras beyond the per-launch limit go in further
// launches that keep what is already on the canvas (later cameras win, Appendix A.5).
int launch_frame(bs_handle *h, const void *const *d_frames, const size_t *pitch, uint8_t *d_canvas,
                 size_t cpitch, bool keep, cudaStream_t stream)
{
    static const bool dbg_sync = getenv("BS_DEBUG_SYNC") != nullptr;
    if (dbg_sync) {
        cudaError_t e0 = cudaDeviceSynchronize();
        if (e0 != cudaSuccess) return fail(h, BS_ERR_CUDA, "[debug] error pending BEFORE launch_frame: %s", cudaGetErrorString(e0));
    }
    bool fast = h->fast_ok && aligned16(d_canvas, cpitch);
    for (int i = 0; fast && i < h->n_cams; ++i) fast = aligned16(d_frames[i], pitch[i]);
    if (fast) {
        bs::FastArgs FA;
        memset(&FA, 0, sizeof FA);
        bs::FrameArgs &A = FA.F;
        A.n_cams = h->n_cams;
        for (int i = 0; i < A.n_cams; ++i) {
            memcpy(A.cam[i].M, &h->M[9 * i], 9 * sizeof(double));
            A.cam[i].ptr = static_cast<const uint8_t *>(d_frames[i]);
            A.cam[i].pitch = (long long)pitch[i];
            A.cam[i].w = h->sw[i];
            A.cam[i].h = h->sh[i];
            FA.box_we[i] = h->box_we[i];
            FA.box_h[i] = h->box_h[i];
            if (h->box_we[i] > 0) {
                const int er = encode_frame_map(&FA.tmap[i], d_frames[i], pitch[i], h->sh[i], h->box_we[i],
                                                h->box_h[i]);
                if (er != 0)
                    return fail(h, BS_ERR_CUDA,
                                "cuTensorMapEncodeTiled failed (%d) for camera %d: ptr=%p pitch=%zu h=%d box=%dx%d", er,
                                i, d_frames[i], pitch[i], h->sh[i], h->box_we[i], h->box_h[i]);
            }
        }
        A.canvas = d_canvas;
        A.cpitch = (long long)cpitch;
        A.W = h->W;
        A.H = h->H;
        A.block_w = opencv_block_width(h->W, h->H);
        A.legacy_border = (h->flags & BS_FLAG_LEGACY_BORDER) ? 1 : 0;
        A.keep_canvas = keep ? 1 : 0;
        FA.tiles = h->d_tiles;
        FA.n_tiles = h->n_tiles;
        FA.stage_bytes = h->stage_bytes;
        FA.sched = h->d_sched;
        {
            // programmatic dependent launch: the prologue and first tile of this frame may overlap the
            // tail of the previous frame's kernel on the same stream (see stitch_fast.cuh)
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof cfg);
            cfg.gridDim = dim3(h->grid);
            cfg.blockDim = dim3(bs::kFastThreads);
            cfg.dynamicSmemBytes = (size_t)h->smem_bytes;
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            BS_CUDA(h, cudaLaunchKernelEx(&cfg, bs::stitch_fast_kernel, FA));
        }
        BS_CUDA(h, cudaGetLastError());
        if (dbg_sync) {
            cudaError_t e1 = cudaDeviceSynchronize();
            if (e1 != cudaSuccess)
                return fail(h, BS_ERR_CUDA, "[debug] stitch_fast_kernel (grid %d, smem %d) faulted: %s", h->grid,
                            h->smem_bytes, cudaGetErrorString(e1));
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
    }
    return BS_OK;
}

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

const char *bs_last_error(const bs_handle *h) { return h ? h->err.c_str() : g_last_error.c_str(); }

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
    h->flags = flags;
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
        g_last_error = h->err;
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
    for (Slot &s : h->ring) {
        for (uint8_t *p : s.d_frame) cudaFree(p);
        cudaFree(s.d_canvas);
        if (s.ev_h2d) cudaEventDestroy(s.ev_h2d);
        if (s.ev_kernel) cudaEventDestroy(s.ev_kernel);
        if (s.ev_done) cudaEventDestroy(s.ev_done);
    }
    cudaFree(h->d_tiles);
    cudaFree(h->d_sched);
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

    const int64_t t = h->next_ticket;
    Slot &s = h->ring[t % kPipelineDepth];
    if (s.busy) {   // the slot's previous frame must have left the device
        BS_CUDA(h, cudaEventSynchronize(s.ev_done));
        s.busy = false;
    }
    for (int i = 0; i < h->n_cams; ++i)
        BS_CUDA(h, cudaMemcpy2DAsync(s.d_frame[i], h->d_frame_pitch[i], frames[i], pitch[i],
                                     (size_t)h->sw[i] * 3, h->sh[i], cudaMemcpyHostToDevice, h->s_h2d));
    const bool keep = (h->flags & BS_FLAG_KEEP_CANVAS) != 0;
    if (keep)
        BS_CUDA(h, cudaMemcpy2DAsync(s.d_canvas, h->d_canvas_pitch, out, out_pitch, (size_t)h->W * 3,
                                     h->H, cudaMemcpyHostToDevice, h->s_h2d));
    BS_CUDA(h, cudaEventRecord(s.ev_h2d, h->s_h2d));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_compute, s.ev_h2d, 0));
    std::vector<const void *> d_frames(s.d_frame.begin(), s.d_frame.end());
    rc = launch_frame(h, d_frames.data(), h->d_frame_pitch.data(), s.d_canvas, h->d_canvas_pitch, keep,
                      h->s_compute);
    if (rc != BS_OK) return rc;
    BS_CUDA(h, cudaEventRecord(s.ev_kernel, h->s_compute));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_d2h, s.ev_kernel, 0));
    BS_CUDA(h, cudaMemcpy2DAsync(out, out_pitch, s.d_canvas, h->d_canvas_pitch, (size_t)h->W * 3, h->H,
                                 cudaMemcpyDeviceToHost, h->s_d2h));
    BS_CUDA(h, cudaEventRecord(s.ev_done, h->s_d2h));
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
    if (s.ticket != ticket || !s.busy) return BS_OK;
    BS_CUDA(h, cudaSetDevice(h->device));
    BS_CUDA(h, cudaEventSynchronize(s.ev_done));
    s.busy = false;
    return BS_OK;
}

int bs_stitch(bs_handle *h, const uint8_t *const *frames, const size_t *pitch, uint8_t *out,
              size_t out_pitch)
{
    int64_t t = -1;
    int rc = bs_submit(h, frames, pitch, out, out_pitch, &t);
    if (rc != BS_OK) return rc;
    return bs_wait(h, t);
}

int bs_submit_device(bs_handle *h, const void *const *d_frames, const size_t *pitch, void *d_canvas,
                     size_t canvas_pitch, void *stream)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_submit_device: NULL handle");
    if (!d_frames || !pitch || !d_canvas)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: NULL argument");
    for (int i = 0; i < h->n_cams; ++i) {
        if (!d_frames[i]) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: frame %d is NULL", i);
        if (pitch[i] < (size_t)h->sw[i] * 3)
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: pitch[%d]=%zu < %d*3", i, pitch[i],
                        h->sw[i]);
    }
    if (canvas_pitch < (size_t)h->W * 3)
        return fail(h, BS_ERR_INVALID_ARG, "bs_submit_device: canvas_pitch=%zu < %d*3", canvas_pitch, h->W);
    BS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);   // NULL = the legacy default stream
    return launch_frame(h, d_frames, pitch, static_cast<uint8_t *>(d_canvas), canvas_pitch,
                        (h->flags & BS_FLAG_KEEP_CANVAS) != 0, st);
}

int bs_sync(bs_handle *h)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_sync: NULL handle");
    BS_CUDA(h, cudaSetDevice(h->device));
    BS_CUDA(h, cudaStreamSynchronize(h->s_h2d));
    BS_CUDA(h, cudaStreamSynchronize(h->s_compute));
    BS_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    for (Slot &s : h->ring) s.busy = false;
    return BS_OK;
}

int bs_warp_perspective(const uint8_t *src, int sw, int sh, size_t spitch, uint8_t *dst, int dw, int dh,
                        size_t dpitch, const double H[9], unsigned flags)
{
    if (!src || !dst || !H) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_warp_perspective: NULL argument");
    int rc = require_device(nullptr);
    if (rc != BS_OK) return rc;
    bs_handle *h = nullptr;
    rc = bs_create(g_last_device, 1, &sw, &sh, dw, dh, H, flags | BS_FLAG_KEEP_CANVAS, &h);
    if (rc != BS_OK) return rc;
    const uint8_t *frames[1] = {src};
    size_t pitches[1] = {spitch};
    rc = bs_stitch(h, frames, pitches, dst, dpitch);
    if (rc != BS_OK) g_last_error = h->err;
    bs_destroy(h);
    return rc;
}

void *bs_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
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
        out->tile_w = bs::kTileW;
        out->tile_h = bs::kTileH;
        out->n_tiles = h->n_tiles;
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
// cudaHostAlloc, or any host heap routine.
//
// The staged kernel needs scratch memory: the tile plan and its scheduler counters.  Here they live
// in ONE file-scope device array of a fixed size.  Consequences a caller must know:
//   * every GPU that loads this library holds the array (BS_STATIC_SCRATCH_BYTES) until the process exits;
//   * canvases that need more than BS_STATIC_MAX_TILES tiles of 128x32 pixels are refused;
//   * all calls share the array and it keeps the previous call's contents.  The plan is rebuilt in full
//     and the scheduler counters are re-armed by every call, so nothing stale is ever read, but two
//     calls whose GPU work overlaps (different streams, or different host threads on different
//     streams) race on it.  Issue calls on one stream, or order the streams yourself.
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress) and
//     sets the kernel's shared-memory limit; make that first call before any stream capture begins.
// ================================================================================================
#include <mutex>

namespace {

struct StaticScratch {
    bs::TileRec tiles[BS_STATIC_MAX_TILES];
    unsigned sched[64];          // [0] next dynamic tile, [1] CTAs that ran dry; padded to 256 bytes
};
static_assert(sizeof(StaticScratch) == BS_STATIC_SCRATCH_BYTES, "BS_STATIC_SCRATCH_BYTES is out of date");

// The scratch itself.  In host code this name is only a stand-in for the device object: its device
// address comes from cudaGetSymbolAddress (static_device_state), never from &g_static_scratch.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

struct StaticDeviceState {
    bool ready = false;
    StaticScratch *scratch = nullptr;   // device address of g_static_scratch on this device
    int sms = 0;
    int smem_optin = 0;
};
constexpr int kStaticMaxDevices = 64;
StaticDeviceState g_static_dev[kStaticMaxDevices];
std::mutex g_static_mutex;                            // host side of the sharing; the GPU side is the caller's job
bs::TileRec g_static_recs[BS_STATIC_MAX_TILES];       // host staging for the plan, fixed size (no heap)
bs::TileRec g_static_ordered[BS_STATIC_MAX_TILES];
bs::PlanArgs g_static_plan_args;

__global__ void static_sched_arm(unsigned *sched, unsigned first_dynamic)
{
    sched[0] = first_dynamic;   // the first tile of every CTA is static
    sched[1] = 0u;
}

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
    const int tiles_x = (canvas_w + bs::kTileW - 1) / bs::kTileW, tiles_y = (canvas_h + bs::kTileH - 1) / bs::kTileH;
    const long long n_tiles_ll = (long long)tiles_x * tiles_y;
    if (n_tiles_ll > BS_STATIC_MAX_TILES)
        return fail(nullptr, BS_ERR_UNSUPPORTED,
                    "%s: a %dx%d canvas needs %lld tiles of %dx%d pixels, the static scratch holds %d; use bs_create", me,
                    canvas_w, canvas_h, n_tiles_ll, bs::kTileW, bs::kTileH, BS_STATIC_MAX_TILES);
    const int n_tiles = (int)n_tiles_ll;
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

    // ---- plan into the static scratch -------------------------------------------------------------
    bs::PlanArgs &P = g_static_plan_args;
    memset(&P, 0, sizeof P);
    for (int i = 0; i < n_cams; ++i) {
        invert3x3(H + 9 * i, P.M[i]);
        P.w[i] = src_w[i];
        P.h[i] = src_h[i];
    }
    P.n_cams = n_cams;
    P.W = canvas_w;
    P.H = canvas_h;
    P.block_w = block_w;
    P.legacy_border = legacy;
    P.tiles_x = tiles_x;
    P.tiles_y = tiles_y;
    P.out = st->scratch->tiles;
    bs::plan_kernel<<<dim3(tiles_x, tiles_y), bs::kThreads, 0, s>>>(P);
    BS_CUDA(nullptr, cudaGetLastError());
    const size_t plan_bytes = sizeof(bs::TileRec) * (size_t)n_tiles;
    BS_CUDA(nullptr, cudaMemcpyAsync(g_static_recs, st->scratch->tiles, plan_bytes, cudaMemcpyDeviceToHost, s));
    BS_CUDA(nullptr, cudaStreamSynchronize(s));

    // ---- size one TMA box per camera; tiles whose source box is too big go to the exact path -------
    int we[bs::kMaxCamsPerLaunch] = {0}, bh[bs::kMaxCamsPerLaunch] = {0};
    for (int t = 0; t < n_tiles; ++t) {
        bs::TileRec &r = g_static_recs[t];
        if (r.type != bs::kTileFast) continue;
        bool ok = true;
        int twe[bs::kMaxLayers], tbh[bs::kMaxLayers];
        for (int l = 0; l < r.n_layers; ++l) {
            const int e0 = bs::box_elem0(r.bx0[l]), e1 = (3 * r.bx1[l] + 2) >> 2;
            twe[l] = (e1 - e0 + 1 + 3) & ~3;
            tbh[l] = r.by1[l] - r.by0[l] + 1;
            if (twe[l] > 256 || tbh[l] > 256 || twe[l] * 4 * tbh[l] > kStageBytesMax) ok = false;
        }
        if (!ok) {
            r.type = bs::kTileExact;
            continue;
        }
        for (int l = 0; l < r.n_layers; ++l) {
            we[r.cam[l]] = std::max(we[r.cam[l]], twe[l]);
            bh[r.cam[l]] = std::max(bh[r.cam[l]], tbh[l]);
        }
    }
    int stage = 0;
    for (int c = 0; c < n_cams; ++c) {
        if (we[c] * 4 * bh[c] > kStageBytesMax) {
            for (int t = 0; t < n_tiles; ++t) {
                bs::TileRec &r = g_static_recs[t];
                if (r.type != bs::kTileFast) continue;
                for (int l = 0; l < r.n_layers; ++l)
                    if (r.cam[l] == c) r.type = bs::kTileExact;
            }
            we[c] = bh[c] = 0;
        }
        stage = std::max(stage, we[c] * 4 * bh[c]);
    }
    const int stage_bytes = (int)align_up((size_t)stage + bs::kStagePad, 128);
    int n_ordered = 0, n_fast = 0;
    for (int t = 0; t < n_tiles; ++t)
        if (g_static_recs[t].type == bs::kTileExact) g_static_ordered[n_ordered++] = g_static_recs[t];
    for (int layers = bs::kMaxLayers; layers >= 1; --layers)
        for (int t = 0; t < n_tiles; ++t)
            if (g_static_recs[t].type == bs::kTileFast && g_static_recs[t].n_layers == layers)
                g_static_ordered[n_ordered++] = g_static_recs[t], ++n_fast;
    for (int t = 0; t < n_tiles; ++t)
        if (g_static_recs[t].type == bs::kTileEmpty) g_static_ordered[n_ordered++] = g_static_recs[t];

    const int smem_bytes = bs::kStages * stage_bytes + bs::kConsumerWarps * bs::kWarpOutBytes + 2 * bs::kLutBytes +
                           bs::kStages * bs::kMetaBytes + 64;
    bool fast = n_fast > 0 && get_encoder() && smem_bytes <= st->smem_optin && aligned16(d_canvas, canvas_pitch);
    for (int i = 0; fast && i < n_cams; ++i) fast = aligned16(d_frames[i], pitch[i]);
    int per_sm = 0;
    if (fast) {
        BS_CUDA(nullptr, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bs::stitch_fast_kernel, bs::kFastThreads,
                                                                       smem_bytes));
        fast = per_sm >= 1;
    }

    bs::FastArgs FA;
    memset(&FA, 0, sizeof FA);
    bs::FrameArgs &A = FA.F;
    A.n_cams = n_cams;
    for (int i = 0; i < n_cams; ++i) {
        memcpy(A.cam[i].M, P.M[i], 9 * sizeof(double));
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

    if (!fast) {
        // no staged tiles (or unaligned buffers): the per-pixel kernel needs no scratch at all
        dim3 grid((canvas_w + bs::kExactTileW - 1) / bs::kExactTileW, (canvas_h + bs::kExactTileH - 1) / bs::kExactTileH);
        bs::stitch_exact_kernel<<<grid, dim3(32, bs::kExactTileH), 0, s>>>(A);
        BS_CUDA(nullptr, cudaGetLastError());
        return BS_OK;
    }

    for (int i = 0; i < n_cams; ++i) {
        FA.box_we[i] = we[i];
        FA.box_h[i] = bh[i];
        if (we[i] > 0) {
            const int er = encode_frame_map(&FA.tmap[i], d_frames[i], pitch[i], src_h[i], we[i], bh[i]);
            if (er != 0) return fail(nullptr, BS_ERR_CUDA, "%s: cuTensorMapEncodeTiled failed (%d) for camera %d", me, er, i);
        }
    }
    const int grid = std::min(n_tiles, per_sm * st->sms);
    // the ordered plan replaces the raw one in the scratch, and the scheduler is re-armed: both are
    // stream-ordered after the previous call's kernels on this stream
    BS_CUDA(nullptr, cudaMemcpyAsync(st->scratch->tiles, g_static_ordered, plan_bytes, cudaMemcpyHostToDevice, s));
    static_sched_arm<<<1, 1, 0, s>>>(st->scratch->sched, (unsigned)grid);
    BS_CUDA(nullptr, cudaGetLastError());
    FA.tiles = st->scratch->tiles;
    FA.n_tiles = n_tiles;
    FA.stage_bytes = stage_bytes;
    FA.sched = st->scratch->sched;
    bs::stitch_fast_kernel<<<grid, bs::kFastThreads, smem_bytes, s>>>(FA);
    BS_CUDA(nullptr, cudaGetLastError());
    // the host staging arrays are reused by the next call: the copy above must have left them
    BS_CUDA(nullptr, cudaStreamSynchronize(s));
    return BS_OK;
}
