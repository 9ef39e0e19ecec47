// stitch_jpeg.inl -- compressed frames in, compressed panorama out.  Included
// at the end of stitch_api.cu (same translation unit: it works on the handle's frame ring).
//
// The reference decodes with cv::imread / cv::VideoCapture and encodes with cv::imwrite on the host
// (/root/reference/tools/stitch/main.cpp:107,114,125,127-147).  Here a JPEG frame crosses PCIe as it is and is
// decoded by nvJPEG straight into the slot's device frame -- no decoded frame ever exists on the host -- and
// the panorama can leave as a JPEG encoded from the device canvas, so that only bitstreams cross the link.
// nvJPEG is the codec, as libjpeg is the reference's: its pixels may differ by a unit or two from libjpeg's,
// which is why the parity tests feed raw pixels and the JPEG tests compare against nvJPEG's own decode.
#include <nvjpeg.h>

#include "stitch_jpegdec.cuh"

#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>

namespace {

// The entropy (Huffman) stage of a baseline JPEG is sequential and, on this image's nvJPEG, runs on the host:
// ~20 ms for a 4K frame.  The cameras of one frame are independent, so bs_submit_jpeg runs their host stages on
// a few worker threads of its own (nvJPEG's decoupled API: parse + nvjpegDecodeJpegHost per camera in parallel,
// then transfer + device stage on the upload stream in camera order).
// ---- the repo's own JPEG decoder on the GPU (stitch_jpegdec.cuh) ---------------------------------------------------
// The chunked, self-synchronising form (stitch_jpegdec.cuh).  Every restart interval (a marker-less scan is one
// interval) is cut into `T` chunks of its stuffing-free data; one thread per chunk, grid = (T / 128, intervals).
// A ROUND is one launch: every chunk decodes from the state the chunk before it left in the previous round.
struct ChunkRec {
    jd::SyncState st[2];     // exit state of the chunk in even / odd rounds
    int started;             // blocks started in the chunk (last round); after jpeg_scan_kernel: index of its first block
    int dc_sum[3];           // DC differences of those blocks per component; after the scan: predictors at the chunk's entry
    int moved[2];            // the exit state changed in the last even / odd round
};
constexpr int kSyncThreads = 128;
struct SyncShared {
    jd::Header hdr;
    uint8_t nat[64];
};
__device__ __forceinline__ void sync_load_tables(SyncShared &sh, const jd::Header *H)
{
    const uint32_t *src = reinterpret_cast<const uint32_t *>(H);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&sh.hdr);
    for (int i = (int)threadIdx.x; i < (int)(sizeof(jd::Header) / 4); i += (int)blockDim.x) dst[i] = src[i];
    for (int i = (int)threadIdx.x; i < 64; i += (int)blockDim.x) sh.nat[i] = (uint8_t)jd::natural(i);
    __syncthreads();
}
__device__ __forceinline__ uint32_t sync_chunk_bits(uint32_t data_bits, int T)
{
    const uint32_t c = ((data_bits + (uint32_t)T - 1) / (uint32_t)T + 7u) & ~7u;
    return c < 64u ? 64u : c;
}
// round 0: guesses (chunk start, block 0, index 0).  round r > 0: from the exit states of round r - 1; `changed` counts
// the chunks whose exit state moved.
__global__ void __launch_bounds__(kSyncThreads) jpeg_sync_round_kernel(const jd::Header *__restrict__ H, const uint8_t *__restrict__ data,
                                                                      const uint32_t *__restrict__ offs, int T, int round, ChunkRec *recs,
                                                                      int *changed)
{
    __shared__ SyncShared sh;
    sync_load_tables(sh, H);
    const int j = (int)(blockIdx.x * blockDim.x + threadIdx.x), r = (int)blockIdx.y;
    if (j >= T) return;
    const uint8_t *base = data + offs[r];
    const uint32_t data_bits = (offs[r + 1] - offs[r]) * 8u, chunk_bits = sync_chunk_bits(data_bits, T);
    ChunkRec *rec = recs + (size_t)r * T;
    jd::SyncState s;
    if (round == 0 || j == 0) {
        s.pos = round == 0 ? (uint32_t)j * chunk_bits : 0u;
        s.blk = 0;
        s.k = 0;
    } else {
        s = rec[j - 1].st[(round - 1) & 1];
    }
    // a chunk whose entry state did not move in the previous round leaves as it did: later rounds cost almost nothing
    // once the stream has settled, so there can be plenty of them
    if (round >= 2 && (j == 0 || !rec[j - 1].moved[(round - 1) & 1])) {
        rec[j].st[round & 1] = rec[j].st[(round - 1) & 1];
        rec[j].moved[round & 1] = 0;
        return;
    }
    jd::ChunkOut co;
    int16_t *const none[3] = {nullptr, nullptr, nullptr};
    const jd::SyncState ns = jd::run_chunk<false>(sh.hdr, sh.hdr.dc, sh.nat, base, data_bits, s, (uint32_t)(j + 1) * chunk_bits, 0, 0x7fffffff, 0, none, co);
    const int moved = round == 0 || !jd::same(ns, rec[j].st[(round - 1) & 1]);
    if (round > 0 && moved) atomicAdd(changed + r, 1);
    rec[j].moved[round & 1] = moved;
    rec[j].st[round & 1] = ns;
    rec[j].started = co.started;
    rec[j].dc_sum[0] = co.dc_sum[0];
    rec[j].dc_sum[1] = co.dc_sum[1];
    rec[j].dc_sum[2] = co.dc_sum[2];
}
// Exclusive prefix sums over the chunks of every interval (one block per interval): started -> first block index,
// dc_sum -> DC predictors at the chunk's entry.
__global__ void __launch_bounds__(256) jpeg_scan_kernel(ChunkRec *recs, int T)
{
    __shared__ int part[256][4];
    ChunkRec *rec = recs + (size_t)blockIdx.x * T;
    const int t = (int)threadIdx.x, per = (T + 255) / 256, lo = min(T, t * per), hi = min(T, lo + per);
    int acc[4] = {0, 0, 0, 0};
    for (int i = lo; i < hi; ++i) {
        acc[0] += rec[i].started;
        acc[1] += rec[i].dc_sum[0];
        acc[2] += rec[i].dc_sum[1];
        acc[3] += rec[i].dc_sum[2];
    }
    for (int q = 0; q < 4; ++q) part[t][q] = acc[q];
    __syncthreads();
    if (t < 4) {
        int run = 0;
        for (int i = 0; i < 256; ++i) {
            const int v = part[i][t];
            part[i][t] = run;
            run += v;
        }
    }
    __syncthreads();
    for (int q = 0; q < 4; ++q) acc[q] = part[t][q];
    for (int i = lo; i < hi; ++i) {
        const int v0 = rec[i].started, v1 = rec[i].dc_sum[0], v2 = rec[i].dc_sum[1], v3 = rec[i].dc_sum[2];
        rec[i].started = acc[0];
        rec[i].dc_sum[0] = acc[1];
        rec[i].dc_sum[1] = acc[2];
        rec[i].dc_sum[2] = acc[3];
        acc[0] += v0;
        acc[1] += v1;
        acc[2] += v2;
        acc[3] += v3;
    }
}
// The write pass: every chunk decodes once more from its (now final) entry state, coefficients to their blocks.
__global__ void __launch_bounds__(kSyncThreads) jpeg_sync_write_kernel(const jd::Header *__restrict__ H, const uint8_t *__restrict__ data,
                                                                      const uint32_t *__restrict__ offs, int T, int final_round, int dri,
                                                                      const ChunkRec *__restrict__ recs, int16_t *c0, int16_t *c1, int16_t *c2,
                                                                      int *written, const int *__restrict__ unsettled)
{
    __shared__ SyncShared sh;
    sync_load_tables(sh, H);
    const int j = (int)(blockIdx.x * blockDim.x + threadIdx.x), r = (int)blockIdx.y;
    if (j >= T || unsettled[r]) return;          // (an interval whose chunks still moved is decoded by jpeg_sync_serial_kernel)
    const uint8_t *base = data + offs[r];
    const uint32_t data_bits = (offs[r + 1] - offs[r]) * 8u, chunk_bits = sync_chunk_bits(data_bits, T);
    const ChunkRec *rec = recs + (size_t)r * T;
    jd::SyncState s;
    if (j == 0) {
        s.pos = 0;
        s.blk = 0;
        s.k = 0;
    } else {
        s = rec[j - 1].st[final_round & 1];
    }
    const int total = sh.hdr.mcus_x * sh.hdr.mcus_y, m_first = r * dri, nb = sh.hdr.hs * sh.hdr.vs + 2;
    const int max_block = min(dri, total - m_first) * nb;
    jd::ChunkOut co;
    co.dc_base[0] = rec[j].dc_sum[0];
    co.dc_base[1] = rec[j].dc_sum[1];
    co.dc_base[2] = rec[j].dc_sum[2];
    int16_t *const coef[3] = {c0, c1, c2};
    jd::run_chunk<true>(sh.hdr, sh.hdr.dc, sh.nat, base, data_bits, s, (uint32_t)(j + 1) * chunk_bits, rec[j].started, max_block, m_first, coef, co);
    if (co.started) atomicAdd(written, co.started);
}
// The way out for streams that do not re-synchronise (a long flat area is a periodic bit pattern that a decoder can
// follow at the wrong phase for ever; the rounds then advance one chunk at a time): an interval whose chunks still
// moved in the last round is decoded by ONE thread from its start.  Slow and rare; never wrong.
__global__ void __launch_bounds__(32) jpeg_sync_serial_kernel(const jd::Header *__restrict__ H, const uint8_t *__restrict__ data,
                                                             const uint32_t *__restrict__ offs, int n_intervals, int dri, int16_t *c0,
                                                             int16_t *c1, int16_t *c2, int *written, const int *__restrict__ unsettled)
{
    __shared__ SyncShared sh;
    sync_load_tables(sh, H);
    const int r = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (r >= n_intervals || !unsettled[r]) return;
    const int total = sh.hdr.mcus_x * sh.hdr.mcus_y, m_first = r * dri, nb = sh.hdr.hs * sh.hdr.vs + 2;
    const uint32_t data_bits = (offs[r + 1] - offs[r]) * 8u;
    jd::SyncState s;
    s.pos = 0;
    s.blk = 0;
    s.k = 0;
    jd::ChunkOut co;
    co.dc_base[0] = co.dc_base[1] = co.dc_base[2] = 0;
    int16_t *const coef[3] = {c0, c1, c2};
    jd::run_chunk<true>(sh.hdr, sh.hdr.dc, sh.nat, data + offs[r], data_bits, s, data_bits, 0, min(dri, total - m_first) * nb, m_first, coef, co);
    atomicAdd(written, co.started);
}
// flag[0] <- 1 unless every block of the frame was written (flag[1] counts them): the data is damaged.
__global__ void jpeg_sync_check_kernel(const jd::Header *__restrict__ H, int *flag)
{
    const int expected = H->mcus_x * H->mcus_y * (H->hs * H->vs + 2);
    flag[0] = flag[1] != expected ? 1 : 0;
}

// One thread per 8x8 block of any component: dequantise, inverse DCT, samples into the component's plane.
__global__ void __launch_bounds__(128) jpeg_idct_kernel(const jd::Header *__restrict__ H, const int16_t *c0, const int16_t *c1, const int16_t *c2,
                                                       uint8_t *p0, uint8_t *p1, uint8_t *p2)
{
    const int n0 = jd::blocks_w(*H, 0) * jd::blocks_h(*H, 0), n1 = jd::blocks_w(*H, 1) * jd::blocks_h(*H, 1);
    int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n0 + 2 * n1) return;
    const int c = i < n0 ? 0 : (i < n0 + n1 ? 1 : 2);
    i -= c == 0 ? 0 : (c == 1 ? n0 : n0 + n1);
    const int bw = jd::blocks_w(*H, c);
    const size_t pitch = (size_t)bw * 8;
    const int16_t *coef = (c == 0 ? c0 : (c == 1 ? c1 : c2)) + 64 * (size_t)i;
    uint8_t *plane = c == 0 ? p0 : (c == 1 ? p1 : p2);
    jd::idct_islow(coef, H->quant[c], plane + (size_t)(i / bw) * 8 * pitch + (size_t)(i % bw) * 8, pitch);
}
// One thread per four horizontally adjacent pixels: chroma up-sampling, colour conversion, packed BGR into the frame.
__global__ void __launch_bounds__(256) jpeg_color_kernel(const jd::Header *__restrict__ H, const uint8_t *__restrict__ p0, const uint8_t *__restrict__ p1,
                                                        const uint8_t *__restrict__ p2, uint8_t *dst, size_t dpitch)
{
    const int w = H->w, h = H->h, hs = H->hs;
    const int x0 = 4 * (int)(blockIdx.x * blockDim.x + threadIdx.x), y = (int)blockIdx.y;
    if (x0 >= w || y >= h) return;
    const size_t pitch0 = (size_t)jd::blocks_w(*H, 0) * 8, pitch1 = (size_t)jd::blocks_w(*H, 1) * 8;
    const int cw = (w + hs - 1) / hs, ch = (h + H->vs - 1) / H->vs;
    uint32_t px[4];
    const int n = min(4, w - x0);
    for (int k = 0; k < n; ++k) {
        const int x = x0 + k;
        const int Y = p0[(size_t)y * pitch0 + x];
        int cb, cr;
        if (hs == 2) {
            cb = jd::chroma_h2v2(p1, pitch1, cw, ch, x, y);
            cr = jd::chroma_h2v2(p2, pitch1, cw, ch, x, y);
        } else {
            cb = p1[(size_t)y * pitch1 + x];
            cr = p2[(size_t)y * pitch1 + x];
        }
        px[k] = jd::ycc_to_bgr(Y, cb, cr);
    }
    uint8_t *o = dst + (size_t)y * dpitch + (size_t)x0 * 3;
    if (n == 4 && ((reinterpret_cast<uintptr_t>(o) & 3) == 0)) {
        uint32_t *o32 = reinterpret_cast<uint32_t *>(o);
        o32[0] = px[0] | (px[1] << 24);
        o32[1] = (px[1] >> 8) | (px[2] << 16);
        o32[2] = (px[2] >> 16) | (px[3] << 8);
    } else {
        for (int k = 0; k < n; ++k) {
            o[3 * k + 0] = (uint8_t)px[k];
            o[3 * k + 1] = (uint8_t)(px[k] >> 8);
            o[3 * k + 2] = (uint8_t)(px[k] >> 16);
        }
    }
}

// Buffers of the GPU decoder for one camera of one pipeline slot (allocated by the first frame that takes this path).
struct DevDec {
    bool ready = false;
    jd::Header *h_hdr = nullptr, *d_hdr = nullptr;      // pinned / device
    uint8_t *h_scan = nullptr, *d_scan = nullptr;       // entropy-coded data
    size_t scan_cap = 0;
    uint32_t *h_offs = nullptr, *d_offs = nullptr;      // restart interval offsets
    int offs_cap = 0, n_intervals = 0;
    uint32_t clean_bytes = 0;                           // stuffing-free scan data in h_scan
    int *h_flag = nullptr, *d_flag = nullptr;           // [0] != 0 after the launches: the chunks did not synchronise / a block is missing
    ChunkRec *d_recs = nullptr;                         // one record per chunk
    size_t recs_cap = 0;
    int16_t *d_coef = nullptr;                          // the three components' coefficient blocks, one allocation
    size_t coef_bytes[3] = {0, 0, 0};
    uint8_t *d_planes = nullptr;                        // the three components' sample planes, one allocation
    size_t plane_bytes[3] = {0, 0, 0};
    uint64_t device_bytes = 0;
};

struct JpegLane {                              // one camera of one pipeline slot
    nvjpegJpegState_t state = nullptr;
    nvjpegBufferPinned_t pinned = nullptr;
    nvjpegBufferDevice_t device = nullptr;
    nvjpegJpegStream_t stream = nullptr;
    DevDec dev;
    // GPU decoder: every camera of every pipeline slot decodes on a stream of its own (the chains of consecutive frames
    // and of the cameras of one frame overlap: a chain is a few hundred single-lane warps), joined into the upload stream
    cudaStream_t dec_stream = nullptr;
    cudaEvent_t dec_done = nullptr;
};

// A frame whose host stages are running: everything bs_submit_jpeg was given, kept until the GPU part is issued.
struct JpegPending {
    std::vector<const uint8_t *> jpeg;
    std::vector<size_t> bytes;
    std::vector<int> status;                   // per camera: nvjpegStatus_t of the host stage
    uint8_t *out = nullptr;
    size_t out_pitch = 0;
    bs_jpeg_out *jout = nullptr;
    bool use_dev = false;                      // every frame has a restart interval: entropy decoding on the GPU
    int remaining = 0;                         // host stages still running (guarded by JpegWorkers::m)
};

struct JpegWorkers {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv_job, cv_done;
    std::deque<std::pair<int, int>> tasks;     // (pipeline slot, camera)
    std::function<int(int, int)> stage;        // the host stage of (slot, camera) -> status
    JpegPending *pending = nullptr;            // kPipelineDepth entries
    bool stop = false;

    void start(int n, int device)
    {
        for (int k = 0; k < n; ++k)
            th.emplace_back([this, device] {
                cudaSetDevice(device);
                std::unique_lock<std::mutex> lk(m);
                for (;;) {
                    cv_job.wait(lk, [&] { return stop || !tasks.empty(); });
                    if (stop) return;
                    const std::pair<int, int> t = tasks.front();
                    tasks.pop_front();
                    lk.unlock();
                    const int st = stage(t.first, t.second);
                    lk.lock();
                    pending[t.first].status[t.second] = st;
                    if (--pending[t.first].remaining == 0) cv_done.notify_all();
                }
            });
    }
    void post(int slot, int n_cams)            // queue the host stages of one frame; returns at once
    {
        {
            std::lock_guard<std::mutex> lk(m);
            pending[slot].remaining = n_cams;
            for (int i = 0; i < n_cams; ++i) tasks.emplace_back(slot, i);
        }
        cv_job.notify_all();
    }
    bool slot_done(int slot)                   // every host stage of that frame has finished (does not block)
    {
        std::lock_guard<std::mutex> lk(m);
        return pending[slot].remaining == 0;
    }
    void wait_slot(int slot)                   // returns when every host stage of that frame has finished
    {
        std::unique_lock<std::mutex> lk(m);
        cv_done.wait(lk, [&] { return pending[slot].remaining == 0; });
    }
    ~JpegWorkers()
    {
        {
            std::lock_guard<std::mutex> lk(m);
            stop = true;
        }
        cv_job.notify_all();
        for (auto &t : th) t.join();
    }
};

struct JpegCtx {
    nvjpegHandle_t lib = nullptr;              // per-image decoder (host Huffman stage) and the encoder
    std::vector<nvjpegJpegState_t> dec;        // kPipelineDepth x n_cams decode states (one per frame in flight)
    // Batched decoder, one batch = the cameras of one frame: the GPU's JPEG engine when the library drives one
    // on this device, else GPU-assisted Huffman decoding; nullptr when neither initialises.
    nvjpegHandle_t blib = nullptr;
    nvjpegJpegState_t bdec[kPipelineDepth] = {};
    bool batched_ok = false;
    nvjpegEncoderState_t enc[kPipelineDepth] = {};
    nvjpegEncoderParams_t enc_params[kPipelineDepth] = {};
    const char *backend = "default";
    // host entropy stage on worker threads (see JpegLane)
    nvjpegJpegDecoder_t hdec = nullptr;
    nvjpegDecodeParams_t hparams = nullptr;
    std::vector<JpegLane> lanes;               // kPipelineDepth x n_cams
    JpegWorkers *workers = nullptr;
    JpegPending pending[kPipelineDepth];
    bool threaded_ok = false;
    bool last_use_dev = false;                 // the last submitted frame took the GPU entropy decoder
};

void jpeg_lanes_destroy(JpegCtx *c)
{
    delete c->workers;                         // joins the threads
    c->workers = nullptr;
    for (JpegLane &l : c->lanes) {
        if (l.stream) nvjpegJpegStreamDestroy(l.stream);
        if (l.state) nvjpegJpegStateDestroy(l.state);
        if (l.pinned) nvjpegBufferPinnedDestroy(l.pinned);
        if (l.device) nvjpegBufferDeviceDestroy(l.device);
        cudaFreeHost(l.dev.h_hdr);
        cudaFreeHost(l.dev.h_scan);
        cudaFreeHost(l.dev.h_offs);
        cudaFree(l.dev.d_hdr);
        cudaFree(l.dev.d_scan);
        cudaFree(l.dev.d_offs);
        cudaFree(l.dev.d_coef);
        cudaFree(l.dev.d_planes);
        cudaFree(l.dev.d_flag);
        cudaFree(l.dev.d_recs);
        cudaFreeHost(l.dev.h_flag);
        if (l.dec_stream) cudaStreamDestroy(l.dec_stream);
        if (l.dec_done) cudaEventDestroy(l.dec_done);
    }
    c->lanes.clear();
    if (c->hparams) nvjpegDecodeParamsDestroy(c->hparams);
    if (c->hdec) nvjpegDecoderDestroy(c->hdec);
    c->hparams = nullptr;
    c->hdec = nullptr;
    c->threaded_ok = false;
}

// Decoupled decoder with the Huffman stage on the host + one lane per camera and slot + the worker threads.
void jpeg_lanes_create(bs_handle *h, JpegCtx *c)
{
    static const char *off = getenv("BS_DEBUG_JPEG_THREADS");       // test hook: "0" keeps the one-thread decoders
    if (off && off[0] == '0') return;
    bool ok = nvjpegDecoderCreate(c->lib, NVJPEG_BACKEND_HYBRID, &c->hdec) == NVJPEG_STATUS_SUCCESS &&
              nvjpegDecodeParamsCreate(c->lib, &c->hparams) == NVJPEG_STATUS_SUCCESS &&
              nvjpegDecodeParamsSetOutputFormat(c->hparams, NVJPEG_OUTPUT_BGRI) == NVJPEG_STATUS_SUCCESS;
    c->lanes.assign((size_t)kPipelineDepth * h->n_cams, JpegLane());
    for (size_t k = 0; k < c->lanes.size() && ok; ++k) {
        JpegLane &l = c->lanes[k];
        ok = nvjpegDecoderStateCreate(c->lib, c->hdec, &l.state) == NVJPEG_STATUS_SUCCESS &&
             nvjpegBufferPinnedCreate(c->lib, nullptr, &l.pinned) == NVJPEG_STATUS_SUCCESS &&
             nvjpegBufferDeviceCreate(c->lib, nullptr, &l.device) == NVJPEG_STATUS_SUCCESS &&
             nvjpegStateAttachPinnedBuffer(l.state, l.pinned) == NVJPEG_STATUS_SUCCESS &&
             nvjpegStateAttachDeviceBuffer(l.state, l.device) == NVJPEG_STATUS_SUCCESS &&
             nvjpegJpegStreamCreate(c->lib, &l.stream) == NVJPEG_STATUS_SUCCESS;
    }
    if (ok) {
        c->workers = new (std::nothrow) JpegWorkers;
        ok = c->workers != nullptr;
        if (ok) {
            const int n_cams = h->n_cams;
            c->workers->pending = c->pending;
            c->workers->stage = [c, n_cams](int slot, int i) {
                JpegLane &l = c->lanes[(size_t)slot * n_cams + i];
                const JpegPending &p = c->pending[slot];
                if (p.use_dev) {
                    // GPU entropy decoding: the host only removes the byte stuffing and the restart markers while it
                    // moves the scan data into pinned memory (memchr + memcpy, ~0.5 ms per 4K frame)
                    DevDec &d = l.dev;
                    const jd::Header &H = *d.h_hdr;
                    const uint8_t *scan = p.jpeg[i] + H.scan_off;
                    d.n_intervals = jd::unstuff_scan(scan, H.scan_len, H, d.h_scan, d.h_offs, d.offs_cap - 1);
                    if (d.n_intervals < 0) return -1;
                    d.clean_bytes = d.h_offs[d.n_intervals];
                    return 0;
                }
                nvjpegStatus_t st = nvjpegJpegStreamParse(c->lib, p.jpeg[i], p.bytes[i], 0, 0, l.stream);
                if (st == NVJPEG_STATUS_SUCCESS) st = nvjpegDecodeJpegHost(c->lib, c->hdec, l.state, c->hparams, l.stream);
                return (int)st;
            };
            // enough threads for the frames the pipeline holds, but not more than the machine has cores to spare
            const int hw = (int)std::thread::hardware_concurrency();
            const int want = std::min(kPipelineDepth * n_cams, 16);
            c->workers->start(std::max(std::min(want, hw > 2 ? hw - 1 : 1), std::min(n_cams, 4)), h->device);
        }
    }
    for (size_t k = 0; k < c->lanes.size() && ok; ++k)
        ok = cudaStreamCreateWithFlags(&c->lanes[k].dec_stream, cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->lanes[k].dec_done, cudaEventDisableTiming) == cudaSuccess;
    cudaGetLastError();
    if (!ok) jpeg_lanes_destroy(c);
    c->threaded_ok = ok;
    if (ok) c->backend = "hybrid, host entropy stage of the cameras on worker threads";
}

int jpeg_ctx(bs_handle *h, JpegCtx **out)
{
    if (h->jpeg) {
        *out = h->jpeg;
        return BS_OK;
    }
    JpegCtx *c = new (std::nothrow) JpegCtx;
    if (!c) return fail(h, BS_ERR_NOMEM, "nvJPEG context: out of host memory");
    // GPU-assisted Huffman decoding where the library offers it for this GPU, its default otherwise
    if (nvjpegCreateEx(NVJPEG_BACKEND_GPU_HYBRID, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &c->lib) == NVJPEG_STATUS_SUCCESS) {
        c->backend = "gpu_hybrid";
    } else if (nvjpegCreateSimple(&c->lib) != NVJPEG_STATUS_SUCCESS) {
        delete c;
        return fail(h, BS_ERR_CUDA, "nvjpegCreate failed");
    }
    static const char *force = getenv("BS_DEBUG_JPEG_BACKEND");      // test hook: "simple", "gpu_hybrid", "hardware"
    const struct { nvjpegBackend_t be; const char *name; } tries[2] = {{NVJPEG_BACKEND_HARDWARE, "hardware"},
                                                                        {NVJPEG_BACKEND_GPU_HYBRID, "gpu_hybrid"}};
    for (int k = 0; k < 2 && !c->batched_ok; ++k) {
        if (force && strcmp(force, tries[k].name) != 0) continue;
        if (nvjpegCreateEx(tries[k].be, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &c->blib) != NVJPEG_STATUS_SUCCESS) {
            c->blib = nullptr;
            continue;
        }
        bool ok2 = true;
        for (int q = 0; q < kPipelineDepth && ok2; ++q)
            ok2 = nvjpegJpegStateCreate(c->blib, &c->bdec[q]) == NVJPEG_STATUS_SUCCESS &&
                  nvjpegDecodeBatchedInitialize(c->blib, c->bdec[q], h->n_cams, std::min(h->n_cams, 8), NVJPEG_OUTPUT_BGRI) == NVJPEG_STATUS_SUCCESS;
        if (ok2) {
            c->batched_ok = true;
            c->backend = tries[k].name;
        } else {
            for (int q = 0; q < kPipelineDepth; ++q) {
                if (c->bdec[q]) nvjpegJpegStateDestroy(c->bdec[q]);
                c->bdec[q] = nullptr;
            }
            nvjpegDestroy(c->blib);
            c->blib = nullptr;
        }
    }
    cudaGetLastError();
    c->dec.assign((size_t)kPipelineDepth * h->n_cams, nullptr);
    bool ok = true;
    for (auto &st : c->dec) ok = ok && nvjpegJpegStateCreate(c->lib, &st) == NVJPEG_STATUS_SUCCESS;
    for (int k = 0; k < kPipelineDepth && ok; ++k)
        ok = nvjpegEncoderStateCreate(c->lib, &c->enc[k], h->s_d2h) == NVJPEG_STATUS_SUCCESS &&
             nvjpegEncoderParamsCreate(c->lib, &c->enc_params[k], h->s_d2h) == NVJPEG_STATUS_SUCCESS;
    h->jpeg = c;
    *out = c;
    if (!ok) return fail(h, BS_ERR_CUDA, "nvJPEG state creation failed");
    jpeg_lanes_create(h, c);
    return BS_OK;
}

void jpeg_ctx_destroy(JpegCtx *c)
{
    if (!c) return;
    jpeg_lanes_destroy(c);
    for (auto st : c->dec)
        if (st) nvjpegJpegStateDestroy(st);
    for (int k = 0; k < kPipelineDepth; ++k)
        if (c->bdec[k]) nvjpegJpegStateDestroy(c->bdec[k]);
    if (c->blib) nvjpegDestroy(c->blib);
    for (int k = 0; k < kPipelineDepth; ++k) {
        if (c->enc_params[k]) nvjpegEncoderParamsDestroy(c->enc_params[k]);
        if (c->enc[k]) nvjpegEncoderStateDestroy(c->enc[k]);
    }
    if (c->lib) nvjpegDestroy(c->lib);
    delete c;
}

// GPU-decoder set-up of one camera's frame (calling thread): parse the headers, size the lane's buffers.  False when
// the file is not one the GPU decoder takes (no restart interval, progressive, other sampling ...) or memory is short.
bool devdec_prepare(bs_handle *h, JpegLane &l, int cam, const uint8_t *jpeg, size_t bytes)
{
    DevDec &d = l.dev;
    if (!d.h_hdr && cudaHostAlloc(reinterpret_cast<void **>(&d.h_hdr), sizeof(jd::Header), cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        d.h_hdr = nullptr;
        return false;
    }
    jd::Header &H = *d.h_hdr;
    if (!jd::parse_header(jpeg, bytes, H) || H.w != h->sw[cam] || H.h != h->sh[cam]) return false;
    if (H.hs == 2 && (H.w < 3 || H.h < 3)) return false;                  // the up-sampling's edge cases need two chroma columns
    const long total = (long)H.mcus_x * H.mcus_y;
    const int n_iv = H.dri > 0 ? (int)((total + H.dri - 1) / H.dri) : 1;
    size_t cb[3], pb[3];
    for (int c = 0; c < 3; ++c) {
        const size_t blocks = (size_t)jd::blocks_w(H, c) * jd::blocks_h(H, c);
        cb[c] = blocks * 64 * sizeof(int16_t);
        pb[c] = blocks * 64;
    }
    cudaError_t e = cudaSuccess;
    if (!d.d_hdr) e = cudaMalloc(reinterpret_cast<void **>(&d.d_hdr), sizeof(jd::Header));
    if (e == cudaSuccess && !d.d_flag) {
        e = cudaMalloc(reinterpret_cast<void **>(&d.d_flag), 2 * sizeof(int));     // status, blocks written (per-interval counters live behind the chunk records)
        if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&d.h_flag), sizeof(int), cudaHostAllocDefault);
        if (e == cudaSuccess) *d.h_flag = 0;
    }
    if (e == cudaSuccess && d.scan_cap < H.scan_len + 16) {
        BS_CUDA_QUIET(cudaStreamSynchronize(l.dec_stream));                  // the old buffers may still be in use
        cudaFreeHost(d.h_scan);
        cudaFree(d.d_scan);
        d.h_scan = d.d_scan = nullptr;
        d.scan_cap = std::max<size_t>(2 * H.scan_len + 16, 1u << 20);
        e = cudaHostAlloc(reinterpret_cast<void **>(&d.h_scan), d.scan_cap, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d.d_scan), d.scan_cap);
        if (e != cudaSuccess) d.scan_cap = 0;
    }
    if (e == cudaSuccess && d.offs_cap < n_iv + 2) {
        BS_CUDA_QUIET(cudaStreamSynchronize(l.dec_stream));
        cudaFreeHost(d.h_offs);
        cudaFree(d.d_offs);
        d.h_offs = d.d_offs = nullptr;
        d.offs_cap = n_iv + 2;
        e = cudaHostAlloc(reinterpret_cast<void **>(&d.h_offs), sizeof(uint32_t) * (size_t)d.offs_cap, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d.d_offs), sizeof(uint32_t) * (size_t)d.offs_cap);
        if (e != cudaSuccess) d.offs_cap = 0;
    }
    if (e == cudaSuccess && (d.coef_bytes[0] != cb[0] || d.coef_bytes[1] != cb[1])) {
        BS_CUDA_QUIET(cudaStreamSynchronize(l.dec_stream));
        cudaFree(d.d_coef);
        cudaFree(d.d_planes);
        d.d_coef = nullptr;
        d.d_planes = nullptr;
        h->device_bytes -= d.device_bytes;
        d.device_bytes = 0;
        e = cudaMalloc(reinterpret_cast<void **>(&d.d_coef), cb[0] + cb[1] + cb[2]);
        if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d.d_planes), pb[0] + pb[1] + pb[2]);
        for (int c = 0; c < 3; ++c) {
            d.coef_bytes[c] = e == cudaSuccess ? cb[c] : 0;
            d.plane_bytes[c] = e == cudaSuccess ? pb[c] : 0;
        }
        if (e == cudaSuccess) {
            d.device_bytes = cb[0] + cb[1] + cb[2] + pb[0] + pb[1] + pb[2];
            h->device_bytes += d.device_bytes;
        }
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    d.ready = true;
    return true;
}

// Queue the GPU decode of one camera's frame (headers, scan data and interval offsets are in the lane's pinned
// buffers): three copies, a memset and three kernels on the upload stream, BGR into the slot's device frame.
int devdec_issue(bs_handle *h, JpegLane &l, uint8_t *d_frame, size_t pitch, cudaStream_t st)
{
    DevDec &d = l.dev;
    const jd::Header &H = *d.h_hdr;
    BS_CUDA(h, cudaMemcpyAsync(d.d_hdr, d.h_hdr, sizeof(jd::Header), cudaMemcpyHostToDevice, st));
    BS_CUDA(h, cudaMemcpyAsync(d.d_scan, d.h_scan, (size_t)d.clean_bytes + 8, cudaMemcpyHostToDevice, st));
    BS_CUDA(h, cudaMemcpyAsync(d.d_offs, d.h_offs, sizeof(uint32_t) * (size_t)(d.n_intervals + 1), cudaMemcpyHostToDevice, st));
    BS_CUDA(h, cudaMemsetAsync(d.d_coef, 0, d.coef_bytes[0] + d.coef_bytes[1] + d.coef_bytes[2], st));
    int16_t *c0 = d.d_coef, *c1 = c0 + d.coef_bytes[0] / 2, *c2 = c1 + d.coef_bytes[1] / 2;
    uint8_t *p0 = d.d_planes, *p1 = p0 + d.plane_bytes[0], *p2 = p1 + d.plane_bytes[1];
    // a guess round, 24 rounds from the neighbours' states (settled chunks skip their round, so late rounds cost a
    // launch; 200 random pictures of five kinds needed at most 21 with the chunk size below), prefix sums, the write
    // pass for the intervals that settled, the one-thread pass for those that did not, the check
    const long total_mcus = (long)H.mcus_x * H.mcus_y;
    const int dri = H.dri > 0 ? H.dri : (int)total_mcus;
    uint32_t longest = 0;
    for (int i = 0; i < d.n_intervals; ++i) longest = std::max(longest, d.h_offs[i + 1] - d.h_offs[i]);
    // a chunk should hold enough BLOCKS to re-synchronise inside itself: 32 average blocks, at least 256 bytes
    // (noise at high quality has 100-byte blocks, flat areas 1-byte blocks)
    static const int env_chunk = getenv("BS_DEBUG_JPEG_CHUNK") ? std::max(64, atoi(getenv("BS_DEBUG_JPEG_CHUNK"))) : 0;   // experiment hook
    const double avg_block = (double)d.clean_bytes / ((double)total_mcus * (H.hs * H.vs + 2));
    const uint32_t chunk_bytes = env_chunk ? (uint32_t)env_chunk : (uint32_t)std::min(16384.0, std::max(256.0, 32.0 * avg_block));
    const int T = (int)std::max<uint32_t>(1u, (longest + chunk_bytes - 1u) / chunk_bytes);
    const size_t need = (size_t)T * (size_t)d.n_intervals + ((size_t)d.n_intervals * sizeof(int) + sizeof(ChunkRec) - 1) / sizeof(ChunkRec);
    if (d.recs_cap < need) {
        BS_CUDA_QUIET(cudaStreamSynchronize(st));
        cudaFree(d.d_recs);
        d.d_recs = nullptr;
        d.recs_cap = 0;
        BS_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&d.d_recs), sizeof(ChunkRec) * need * 2));
        d.recs_cap = need * 2;
    }
    int *unsettled = reinterpret_cast<int *>(d.d_recs + (size_t)T * (size_t)d.n_intervals);      // per interval: chunks that moved in the last round
    constexpr int kRounds = 24;
    BS_CUDA(h, cudaMemsetAsync(d.d_flag, 0, 2 * sizeof(int), st));
    const dim3 grid((T + kSyncThreads - 1) / kSyncThreads, d.n_intervals);
    for (int round = 0; round <= kRounds; ++round) {
        if (round == kRounds) BS_CUDA(h, cudaMemsetAsync(unsettled, 0, sizeof(int) * (size_t)d.n_intervals, st));    // only the last round's changes count
        jpeg_sync_round_kernel<<<grid, kSyncThreads, 0, st>>>(d.d_hdr, d.d_scan, d.d_offs, T, round, d.d_recs, unsettled);
    }
    jpeg_scan_kernel<<<d.n_intervals, 256, 0, st>>>(d.d_recs, T);
    jpeg_sync_write_kernel<<<grid, kSyncThreads, 0, st>>>(d.d_hdr, d.d_scan, d.d_offs, T, kRounds, dri, d.d_recs, c0, c1, c2, d.d_flag + 1, unsettled);
    jpeg_sync_serial_kernel<<<(d.n_intervals + 31) / 32, 32, 0, st>>>(d.d_hdr, d.d_scan, d.d_offs, d.n_intervals, dri, c0, c1, c2, d.d_flag + 1, unsettled);
    jpeg_sync_check_kernel<<<1, 1, 0, st>>>(d.d_hdr, d.d_flag);
    const int n_blocks = (int)((d.plane_bytes[0] + d.plane_bytes[1] + d.plane_bytes[2]) / 64);
    jpeg_idct_kernel<<<(n_blocks + 127) / 128, 128, 0, st>>>(d.d_hdr, c0, c1, c2, p0, p1, p2);
    jpeg_color_kernel<<<dim3(((H.w + 3) / 4 + 255) / 256, H.h), 256, 0, st>>>(d.d_hdr, p0, p1, p2, d_frame, pitch);
    BS_CUDA(h, cudaGetLastError());
    BS_CUDA(h, cudaMemcpyAsync(d.h_flag, d.d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    h->kernel_launches += 9;
    return BS_OK;
}

// bs_submit_jpeg's checks: every bitstream is a JPEG of its camera's frame size.
int check_jpeg_frames(bs_handle *h, JpegCtx *c, const uint8_t *const *jpeg, const size_t *jpeg_bytes)
{
    for (int i = 0; i < h->n_cams; ++i) {
        int nc = 0, ws[NVJPEG_MAX_COMPONENT], hs[NVJPEG_MAX_COMPONENT];
        nvjpegChromaSubsampling_t ss;
        if (nvjpegGetImageInfo(c->lib, jpeg[i], jpeg_bytes[i], &nc, &ss, ws, hs) != NVJPEG_STATUS_SUCCESS)
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: frame %d is not a JPEG nvJPEG can read", i);
        if (ws[0] != h->sw[i] || hs[0] != h->sh[i])
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: frame %d is %dx%d, the handle expects %dx%d", i, ws[0], hs[0], h->sw[i],
                        h->sh[i]);
    }
    return BS_OK;
}

// Queue the GPU side of one frame whose camera images arrive as JPEG bitstreams: decode into the slot's device
// frames on the upload stream, kernels on the compute stream, then the panorama leaves raw (out), as a JPEG (jout),
// or both.  `host_status` != nullptr: the host stages already ran on the worker threads (one status per camera).
int issue_frame_jpeg(bs_handle *h, Slot &s, int slot_index, const uint8_t *const *jpeg, const size_t *jpeg_bytes, uint8_t *out,
                     size_t out_pitch, bs_jpeg_out *jout, const int *host_status)
{
    JpegCtx *c = h->jpeg;
    int rc = BS_OK;
    h->h2d_bytes_last = 0;
    std::vector<nvjpegImage_t> imgs(h->n_cams);
    for (int i = 0; i < h->n_cams; ++i) {
        memset(&imgs[i], 0, sizeof imgs[i]);
        imgs[i].channel[0] = s.d_frame[i];
        imgs[i].pitch[0] = h->d_frame_pitch[i];
        h->h2d_bytes_last += jpeg_bytes[i];
    }
    bool decoded = false;
    if (host_status && c->pending[slot_index].use_dev) {
        // entropy decoding on the GPU (every frame has a restart interval): the workers located the intervals
        decoded = true;
        for (int i = 0; i < h->n_cams; ++i) decoded = decoded && host_status[i] == 0;
        for (int i = 0; i < h->n_cams && decoded; ++i) {
            // the chain needs nothing from the upload stream (the slot's buffers were released by bs_wait); the
            // upload stream goes on when every camera's chain is done
            JpegLane &l = c->lanes[(size_t)slot_index * h->n_cams + i];
            rc = devdec_issue(h, l, s.d_frame[i], h->d_frame_pitch[i], l.dec_stream);
            if (rc != BS_OK) return rc;
            BS_CUDA(h, cudaEventRecord(l.dec_done, l.dec_stream));
            BS_CUDA(h, cudaStreamWaitEvent(h->s_h2d, l.dec_done, 0));
        }
        if (!decoded) {
            cudaStreamSynchronize(h->s_h2d);
            cudaGetLastError();
        }
    } else if (host_status) {
        // device stages in camera order on the upload stream; a bitstream the decoupled decoder could not take
        // (progressive ...) sends the frame to the decoders below
        JpegLane *lanes = &c->lanes[(size_t)slot_index * h->n_cams];
        decoded = true;
        for (int i = 0; i < h->n_cams; ++i) decoded = decoded && host_status[i] == (int)NVJPEG_STATUS_SUCCESS;
        for (int i = 0; i < h->n_cams && decoded; ++i)
            decoded = nvjpegDecodeJpegTransferToDevice(c->lib, c->hdec, lanes[i].state, lanes[i].stream, h->s_h2d) == NVJPEG_STATUS_SUCCESS &&
                      nvjpegDecodeJpegDevice(c->lib, c->hdec, lanes[i].state, &imgs[i], h->s_h2d) == NVJPEG_STATUS_SUCCESS;
        if (!decoded) {
            cudaStreamSynchronize(h->s_h2d);       // a partly issued frame: let it drain before the fall-back rewrites the frames
            cudaGetLastError();
        }
    }
    if (!decoded && c->batched_ok) {
        // all cameras of the frame in one batch; a bitstream this backend cannot take (progressive, odd sampling)
        // sends the frame to the per-image decoder below
        decoded = nvjpegDecodeBatched(c->blib, c->bdec[slot_index], jpeg, jpeg_bytes, imgs.data(), h->s_h2d) == NVJPEG_STATUS_SUCCESS;
        if (!decoded) cudaGetLastError();
    }
    for (int i = 0; i < h->n_cams && !decoded; ++i) {
        const nvjpegStatus_t st = nvjpegDecode(c->lib, c->dec[(size_t)slot_index * h->n_cams + i], jpeg[i], jpeg_bytes[i], NVJPEG_OUTPUT_BGRI,
                                              &imgs[i], h->s_h2d);
        if (st != NVJPEG_STATUS_SUCCESS) return fail(h, BS_ERR_CUDA, "bs_submit_jpeg: nvjpegDecode of frame %d failed (%d)", i, (int)st);
    }
    const bool keep = (h->flags & BS_FLAG_KEEP_CANVAS) != 0;
    if (keep && out)
        BS_CUDA(h, cudaMemcpy2DAsync(s.d_canvas, h->d_canvas_pitch, out, out_pitch, (size_t)h->W * 3, h->H, cudaMemcpyHostToDevice, h->s_h2d));
    BS_CUDA(h, cudaEventRecord(s.ev_h2d, h->s_h2d));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_compute, s.ev_h2d, 0));
    std::vector<const void *> d_frames(s.d_frame.begin(), s.d_frame.end()), warp_frames(h->n_cams);
    std::vector<size_t> warp_pitch(h->n_cams);
    bool lens_ran = false;
    rc = undistort_frames(h, d_frames.data(), h->d_frame_pitch.data(), s.d_undist, warp_frames.data(), warp_pitch.data(), h->s_compute,
                          false, &lens_ran);
    if (rc != BS_OK) return rc;
    // the decoder's kernels ran on another stream, ordered by the event: the frames are settled
    rc = launch_frame(h, warp_frames.data(), warp_pitch.data(), s.d_canvas, h->d_canvas_pitch, keep, h->s_compute, lens_ran);
    if (rc != BS_OK) return rc;
    BS_CUDA(h, cudaEventRecord(s.ev_kernel, h->s_compute));
    BS_CUDA(h, cudaStreamWaitEvent(h->s_d2h, s.ev_kernel, 0));
    if (out)
        BS_CUDA(h, cudaMemcpy2DAsync(out, out_pitch, s.d_canvas, h->d_canvas_pitch, (size_t)h->W * 3, h->H, cudaMemcpyDeviceToHost, h->s_d2h));
    if (jout) {
        const int q = jout->quality > 0 ? jout->quality : 95;        // cv::imwrite's default JPEG quality, and its 4:2:0 sampling
        if (nvjpegEncoderParamsSetQuality(c->enc_params[slot_index], q, h->s_d2h) != NVJPEG_STATUS_SUCCESS ||
            nvjpegEncoderParamsSetSamplingFactors(c->enc_params[slot_index], NVJPEG_CSS_420, h->s_d2h) != NVJPEG_STATUS_SUCCESS)
            return fail(h, BS_ERR_CUDA, "bs_submit_jpeg: encoder parameters rejected (quality %d)", q);
        nvjpegImage_t img;
        memset(&img, 0, sizeof img);
        img.channel[0] = s.d_canvas;
        img.pitch[0] = h->d_canvas_pitch;
        const nvjpegStatus_t st = nvjpegEncodeImage(c->lib, c->enc[slot_index], c->enc_params[slot_index], &img, NVJPEG_INPUT_BGRI, h->W, h->H,
                                                   h->s_d2h);
        if (st != NVJPEG_STATUS_SUCCESS) return fail(h, BS_ERR_CUDA, "bs_submit_jpeg: nvjpegEncodeImage failed (%d)", (int)st);
        jout->size = 0;
    }
    s.jout = jout;
    BS_CUDA(h, cudaEventRecord(s.ev_done, h->s_d2h));
    return BS_OK;
}

// bs_wait's part for a frame that leaves as a JPEG: fetch the bitstream (only it crosses the link).
int retrieve_jpeg(bs_handle *h, Slot &s, int slot_index)
{
    if (h->jpeg && !h->jpeg->lanes.empty())
        for (int i = 0; i < h->n_cams; ++i) {
            DevDec &d = h->jpeg->lanes[(size_t)slot_index * h->n_cams + i].dev;
            if (d.h_flag && *d.h_flag) {
                *d.h_flag = 0;
                s.jout = nullptr;
                return fail(h, BS_ERR_CUDA, "bs_wait: the entropy-coded data of frame %d could not be decoded (damaged bitstream)", i);
            }
        }
    bs_jpeg_out *jout = s.jout;
    s.jout = nullptr;
    if (!jout) return BS_OK;
    JpegCtx *c = h->jpeg;
    size_t len = 0;
    if (nvjpegEncodeRetrieveBitstream(c->lib, c->enc[slot_index], nullptr, &len, h->s_d2h) != NVJPEG_STATUS_SUCCESS)
        return fail(h, BS_ERR_CUDA, "bs_wait: nvJPEG bitstream length unavailable");
    BS_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    if (len > jout->capacity) {
        jout->size = len;
        return fail(h, BS_ERR_INVALID_ARG, "bs_wait: the JPEG needs %zu bytes, the buffer holds %zu", len, jout->capacity);
    }
    if (nvjpegEncodeRetrieveBitstream(c->lib, c->enc[slot_index], jout->data, &len, h->s_d2h) != NVJPEG_STATUS_SUCCESS)
        return fail(h, BS_ERR_CUDA, "bs_wait: nvJPEG bitstream retrieval failed");
    BS_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    jout->size = len;
    return BS_OK;
}

// Issue the GPU side of every frame whose host stages were left running by bs_submit_jpeg, oldest first, up to
// ticket `upto` (frames go to the streams in ticket order).  A frame that fails is dropped and its error kept in
// its slot for the bs_wait of that ticket.
int flush_deferred_jpeg(bs_handle *h, int64_t upto, bool only_ready)
{
    if (!h->jpeg || h->ring.empty()) return BS_OK;
    JpegCtx *c = h->jpeg;
    for (int64_t t = std::max<int64_t>(0, h->next_ticket - kPipelineDepth); t <= upto && t < h->next_ticket; ++t) {
        const int slot_index = (int)(t % kPipelineDepth);
        Slot &s = h->ring[slot_index];
        if (s.ticket != t || !s.deferred) continue;
        if (only_ready && !c->workers->slot_done(slot_index)) break;      // frames go to the streams in ticket order
        c->workers->wait_slot(slot_index);
        s.deferred = false;
        JpegPending &p = c->pending[slot_index];
        const int rc = issue_frame_jpeg(h, s, slot_index, p.jpeg.data(), p.bytes.data(), p.out, p.out_pitch, p.jout, p.status.data());
        if (rc != BS_OK) {
            cudaStreamSynchronize(h->s_h2d);
            cudaStreamSynchronize(h->s_compute);
            cudaStreamSynchronize(h->s_d2h);
            cudaGetLastError();
            s.busy = false;
            s.jout = nullptr;
            s.deferred_rc = rc;
            snprintf(s.deferred_err, sizeof s.deferred_err, "%s", h->err);
        }
    }
    return BS_OK;
}

}  // namespace

extern "C" int bs_submit_jpeg(bs_handle *h, const uint8_t *const *jpeg, const size_t *jpeg_bytes, uint8_t *out, size_t out_pitch,
                              bs_jpeg_out *jout, int64_t *ticket)
{
    if (!h) return fail(nullptr, BS_ERR_INVALID_ARG, "bs_submit_jpeg: NULL handle");
    if (!jpeg || !jpeg_bytes || (!out && !jout)) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: NULL argument (give `out`, `jout` or both)");
    for (int i = 0; i < h->n_cams; ++i)
        if (!jpeg[i] || jpeg_bytes[i] < 4) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: frame %d is empty", i);
    if (out && out_pitch < (size_t)h->W * 3) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: out_pitch=%zu < %d*3", out_pitch, h->W);
    if (jout && (!jout->data || jout->capacity == 0)) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: jout has no buffer");
    if ((h->flags & BS_FLAG_KEEP_CANVAS) && !out) return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: BS_FLAG_KEEP_CANVAS needs the prior canvas in `out`");
    BS_CUDA(h, cudaSetDevice(h->device));
    int rc = ensure_ring(h);
    if (rc != BS_OK) return rc;
    JpegCtx *c = nullptr;
    rc = jpeg_ctx(h, &c);
    if (rc != BS_OK) return rc;
    rc = check_jpeg_frames(h, c, jpeg, jpeg_bytes);
    if (rc != BS_OK) return rc;
    flush_deferred_jpeg(h, h->next_ticket - 1, true);       // earlier frames whose host stages are done go to the GPU now
    const int64_t t = h->next_ticket;
    const int slot_index = (int)(t % kPipelineDepth);
    Slot &s = h->ring[slot_index];
    if (s.busy || s.deferred_rc != BS_OK) {
        rc = bs_wait(h, s.ticket);
        if (rc != BS_OK) return rc;
    }
    if (c->threaded_ok) {
        // The entropy stages start on the worker threads now; the GPU side is issued when they are done -- by the
        // next call on this handle that needs the frame (bs_wait, a later submit, bs_sync), so that the host stages
        // of up to bs_pipeline_depth frames overlap.  The bitstreams must stay valid until bs_wait, as `out` must.
        JpegPending &p = c->pending[slot_index];
        static const bool no_dev = getenv("BS_DEBUG_JPEG_NO_GPU_ENTROPY") != nullptr;     // test hook: always nvJPEG
        p.use_dev = !no_dev;
        for (int i = 0; i < h->n_cams && p.use_dev; ++i)
            p.use_dev = devdec_prepare(h, c->lanes[(size_t)slot_index * h->n_cams + i], i, jpeg[i], jpeg_bytes[i]);
        c->last_use_dev = p.use_dev;
        p.jpeg.assign(jpeg, jpeg + h->n_cams);
        p.bytes.assign(jpeg_bytes, jpeg_bytes + h->n_cams);
        p.status.assign(h->n_cams, 0);
        p.out = out;
        p.out_pitch = out_pitch;
        p.jout = jout;
        c->workers->post(slot_index, h->n_cams);
        s.deferred = true;
    } else {
        rc = issue_frame_jpeg(h, s, slot_index, jpeg, jpeg_bytes, out, out_pitch, jout, nullptr);
        if (rc != BS_OK) {
            cudaStreamSynchronize(h->s_h2d);
            cudaStreamSynchronize(h->s_compute);
            cudaStreamSynchronize(h->s_d2h);
            cudaGetLastError();
            s.busy = false;
            s.ticket = -1;
            s.jout = nullptr;
            return rc;
        }
    }
    s.ticket = t;
    s.busy = true;
    h->next_ticket++;
    if (ticket) *ticket = t;
    return BS_OK;
}

extern "C" const char *bs_jpeg_backend(const bs_handle *h)
{
    if (!h || !h->jpeg) return "";
    return h->jpeg->last_use_dev ? "b200 (entropy decoding on the GPU in self-synchronising chunks; libjpeg arithmetic)" : h->jpeg->backend;
}
