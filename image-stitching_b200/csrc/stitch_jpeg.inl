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
struct JpegLane {                              // one camera of one pipeline slot
    nvjpegJpegState_t state = nullptr;
    nvjpegBufferPinned_t pinned = nullptr;
    nvjpegBufferDevice_t device = nullptr;
    nvjpegJpegStream_t stream = nullptr;
};

// A frame whose host stages are running: everything bs_submit_jpeg was given, kept until the GPU part is issued.
struct JpegPending {
    std::vector<const uint8_t *> jpeg;
    std::vector<size_t> bytes;
    std::vector<int> status;                   // per camera: nvjpegStatus_t of the host stage
    uint8_t *out = nullptr;
    size_t out_pitch = 0;
    bs_jpeg_out *jout = nullptr;
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
    nvjpegJpegState_t bdec[kPipelineDepth] = {nullptr, nullptr, nullptr};
    bool batched_ok = false;
    nvjpegEncoderState_t enc[kPipelineDepth] = {nullptr, nullptr, nullptr};
    nvjpegEncoderParams_t enc_params[kPipelineDepth] = {nullptr, nullptr, nullptr};
    const char *backend = "default";
    // host entropy stage on worker threads (see JpegLane)
    nvjpegJpegDecoder_t hdec = nullptr;
    nvjpegDecodeParams_t hparams = nullptr;
    std::vector<JpegLane> lanes;               // kPipelineDepth x n_cams
    JpegWorkers *workers = nullptr;
    JpegPending pending[kPipelineDepth];
    bool threaded_ok = false;
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
    if (h->n_cams < 2 || (off && off[0] == '0')) return;
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
    if (host_status) {
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
int flush_deferred_jpeg(bs_handle *h, int64_t upto)
{
    if (!h->jpeg || h->ring.empty()) return BS_OK;
    JpegCtx *c = h->jpeg;
    for (int64_t t = std::max<int64_t>(0, h->next_ticket - kPipelineDepth); t <= upto && t < h->next_ticket; ++t) {
        const int slot_index = (int)(t % kPipelineDepth);
        Slot &s = h->ring[slot_index];
        if (s.ticket != t || !s.deferred) continue;
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
    return (h && h->jpeg) ? h->jpeg->backend : "";
}
