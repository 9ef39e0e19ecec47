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

struct JpegWorkers {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv_job, cv_done;
    std::function<void(int)> job;              // job(i) for i in [0, n_jobs)
    int n_jobs = 0, next = 0, done = 0;
    uint64_t generation = 0;
    bool stop = false;

    void start(int n, int device)
    {
        for (int k = 0; k < n; ++k)
            th.emplace_back([this, device] {
                cudaSetDevice(device);
                uint64_t seen = 0;
                std::unique_lock<std::mutex> lk(m);
                for (;;) {
                    cv_job.wait(lk, [&] { return stop || (generation != seen && next < n_jobs); });
                    if (stop) return;
                    while (next < n_jobs) {
                        const int i = next++;
                        lk.unlock();
                        job(i);
                        lk.lock();
                        if (++done == n_jobs) cv_done.notify_all();
                    }
                    seen = generation;
                }
            });
    }
    void run(int n, std::function<void(int)> f)    // returns when every job(i) has finished
    {
        std::unique_lock<std::mutex> lk(m);
        job = std::move(f);
        n_jobs = n;
        next = done = 0;
        ++generation;
        cv_job.notify_all();
        cv_done.wait(lk, [&] { return done == n_jobs; });
        n_jobs = 0;
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
        if (ok) c->workers->start(std::min(h->n_cams, 8), h->device);
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

// Queue one frame whose camera images arrive as JPEG bitstreams: decode into the slot's device frames on the
// upload stream, kernels on the compute stream, then the panorama leaves raw (out), as a JPEG (jout), or both.
int issue_frame_jpeg(bs_handle *h, Slot &s, int slot_index, const uint8_t *const *jpeg, const size_t *jpeg_bytes, uint8_t *out,
                     size_t out_pitch, bs_jpeg_out *jout)
{
    JpegCtx *c = nullptr;
    int rc = jpeg_ctx(h, &c);
    if (rc != BS_OK) return rc;
    h->h2d_bytes_last = 0;
    std::vector<nvjpegImage_t> imgs(h->n_cams);
    for (int i = 0; i < h->n_cams; ++i) {
        int nc = 0, ws[NVJPEG_MAX_COMPONENT], hs[NVJPEG_MAX_COMPONENT];
        nvjpegChromaSubsampling_t ss;
        if (nvjpegGetImageInfo(c->lib, jpeg[i], jpeg_bytes[i], &nc, &ss, ws, hs) != NVJPEG_STATUS_SUCCESS)
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: frame %d is not a JPEG nvJPEG can read", i);
        if (ws[0] != h->sw[i] || hs[0] != h->sh[i])
            return fail(h, BS_ERR_INVALID_ARG, "bs_submit_jpeg: frame %d is %dx%d, the handle expects %dx%d", i, ws[0], hs[0], h->sw[i],
                        h->sh[i]);
        memset(&imgs[i], 0, sizeof imgs[i]);
        imgs[i].channel[0] = s.d_frame[i];
        imgs[i].pitch[0] = h->d_frame_pitch[i];
        h->h2d_bytes_last += jpeg_bytes[i];
    }
    bool decoded = false;
    if (c->threaded_ok) {
        // host stages of all cameras at once on the worker threads, device stages in camera order on the upload
        // stream; a bitstream this decoder cannot take (progressive ...) sends the frame to the decoders below
        std::vector<int> status(h->n_cams, 0);
        JpegLane *lanes = &c->lanes[(size_t)slot_index * h->n_cams];
        c->workers->run(h->n_cams, [&](int i) {
            nvjpegStatus_t st = nvjpegJpegStreamParse(c->lib, jpeg[i], jpeg_bytes[i], 0, 0, lanes[i].stream);
            if (st == NVJPEG_STATUS_SUCCESS) st = nvjpegDecodeJpegHost(c->lib, c->hdec, lanes[i].state, c->hparams, lanes[i].stream);
            status[i] = (int)st;
        });
        decoded = true;
        for (int i = 0; i < h->n_cams; ++i) decoded = decoded && status[i] == (int)NVJPEG_STATUS_SUCCESS;
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
    const int64_t t = h->next_ticket;
    const int slot_index = (int)(t % kPipelineDepth);
    Slot &s = h->ring[slot_index];
    if (s.busy) {
        rc = bs_wait(h, s.ticket);
        if (rc != BS_OK) return rc;
    }
    rc = issue_frame_jpeg(h, s, slot_index, jpeg, jpeg_bytes, out, out_pitch, jout);
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
