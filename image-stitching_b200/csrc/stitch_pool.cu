// stitch_pool.cu -- frame-parallel dispatch over the GPUs of one box, behind the C ABI.
//
// The reference's video loop (/root/reference/tools/stitch/main.cpp:143-155) is one thread producing one
// panorama after the other; every output frame is independent (the canvas is re-created at :144) and the
// calibration is read-only, so frames shard over GPUs with no exchange at all: frame f goes to GPU f mod G
// (SURVEY.md section 8e).  A pool owns one bs_handle and one host thread per GPU; each thread drives its own
// handle (handles are single-threaded) and runs next to its GPU; callers submit frames in order from one
// thread and wait for them in order, so the panoramas come back in frame order.
//
// No collective, no peer traffic: every GPU reads its own frames and writes its own canvas.
#include "../../include/b200_stitch.h"

#include <cuda_runtime.h>

#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

namespace {

struct Job {
    int64_t frame = -1;
    std::vector<const uint8_t *> frames;
    std::vector<size_t> pitch;
    uint8_t *out = nullptr;
    size_t out_pitch = 0;
};

struct Worker {
    bs_handle *h = nullptr;
    int device = 0;
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::deque<Job> queue;            // submitted, not yet handed to the handle
    std::deque<std::pair<int64_t, int64_t>> inflight;   // (frame, ticket) in the handle's pipeline, oldest first
    int64_t done_upto = -1;           // every frame of this worker up to this one has been waited for
    int64_t last_submitted = -1;
    int64_t wanted = -1;              // highest frame somebody is waiting for
    int64_t failed_frame = -1;        // a frame whose submit failed while older frames were still in flight
    int rc = BS_OK;                   // first error of this worker (sticky)
    char err[512] = "";
    bool stop = false;
};

}  // namespace

struct bs_pool {
    int n = 0, n_cams = 0, depth = 0;
    std::vector<Worker *> w;
    int64_t next_frame = 0;
    char err[512] = "";
};

namespace {

int pool_fail(bs_pool *p, int code, const char *msg)
{
    if (p) snprintf(p->err, sizeof p->err, "%s", msg);
    return code;
}

// One per GPU: hand queued frames to the handle while its pipeline has room; wait for the oldest frame in
// flight when the pipeline is full or somebody asked for it (so that up to `depth` frames overlap their copies).
void worker_main(Worker *w, int depth)
{
    cudaSetDevice(w->device);
    std::unique_lock<std::mutex> lk(w->m);
    for (;;) {
        auto can_submit = [&] { return !w->queue.empty() && (int)w->inflight.size() < depth && w->rc == BS_OK; };
        auto must_drop = [&] { return w->rc != BS_OK && !w->queue.empty(); };
        auto must_wait = [&] {
            return !w->inflight.empty() &&
                   ((int)w->inflight.size() >= depth || w->wanted >= w->inflight.front().first || w->stop || w->rc != BS_OK);
        };
        w->cv.wait(lk, [&] { return w->stop || can_submit() || must_drop() || must_wait(); });
        if (can_submit()) {
            Job j = std::move(w->queue.front());
            w->queue.pop_front();
            lk.unlock();
            int64_t t = -1;
            const int rc = bs_submit(w->h, j.frames.data(), j.pitch.data(), j.out, j.out_pitch, &t);
            lk.lock();
            if (rc != BS_OK) {
                w->rc = rc;
                snprintf(w->err, sizeof w->err, "frame %lld on device %d: %s", (long long)j.frame, w->device, bs_last_error(w->h));
                if (w->inflight.empty()) w->done_upto = j.frame;     // nobody must wait for it forever
                else w->failed_frame = j.frame;
                w->cv.notify_all();
            } else {
                w->inflight.emplace_back(j.frame, t);
            }
        } else if (must_wait()) {
            const std::pair<int64_t, int64_t> f = w->inflight.front();
            lk.unlock();
            const int rc = bs_wait(w->h, f.second);
            lk.lock();
            w->inflight.pop_front();
            if (rc != BS_OK && w->rc == BS_OK) {
                w->rc = rc;
                snprintf(w->err, sizeof w->err, "frame %lld on device %d: %s", (long long)f.first, w->device, bs_last_error(w->h));
            }
            w->done_upto = (w->inflight.empty() && w->failed_frame > f.first) ? w->failed_frame : f.first;
            w->cv.notify_all();
        } else if (must_drop()) {      // after an error the remaining frames are dropped
            w->done_upto = w->queue.back().frame;
            w->queue.clear();
            w->cv.notify_all();
        } else if (w->stop) {
            return;
        }
    }
}

}  // namespace

extern "C" {

int bs_pool_create(int n_devices, const int *device_ids, int n_cams, const int *src_w, const int *src_h, int canvas_w,
                   int canvas_h, const double *H, unsigned flags, bs_pool **out)
{
    if (!out) return BS_ERR_INVALID_ARG;
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return BS_ERR_INVALID_ARG;
    bs_pool *p = new (std::nothrow) bs_pool;
    if (!p) return BS_ERR_NOMEM;
    p->n = n_devices;
    p->n_cams = n_cams;
    for (int i = 0; i < n_devices; ++i) {
        Worker *w = new (std::nothrow) Worker;
        if (!w) {
            bs_pool_destroy(p);
            return BS_ERR_NOMEM;
        }
        p->w.push_back(w);
        w->device = device_ids ? device_ids[i] : i;
        const int rc = bs_create(w->device, n_cams, src_w, src_h, canvas_w, canvas_h, H, flags, &w->h);
        if (rc != BS_OK) {
            bs_pool_destroy(p);       // bs_last_error(NULL) keeps the text of the failed bs_create
            return rc;
        }
    }
    p->depth = bs_pipeline_depth(p->w[0]->h);
    for (Worker *w : p->w) w->th = std::thread(worker_main, w, p->depth);
    *out = p;
    return BS_OK;
}

int bs_pool_destroy(bs_pool *p)
{
    if (!p) return BS_OK;
    for (Worker *w : p->w) {
        if (w->th.joinable()) {
            {
                std::lock_guard<std::mutex> lk(w->m);
                w->stop = true;
            }
            w->cv.notify_all();
            w->th.join();
        }
        if (w->h) bs_destroy(w->h);
        delete w;
    }
    delete p;
    return BS_OK;
}

int bs_pool_size(const bs_pool *p) { return p ? p->n : 0; }

int bs_pool_depth(const bs_pool *p) { return p ? p->n * p->depth : 0; }

const char *bs_pool_last_error(const bs_pool *p) { return p ? p->err : bs_last_error(nullptr); }

bs_handle *bs_pool_handle(bs_pool *p, int i) { return (p && i >= 0 && i < p->n) ? p->w[i]->h : nullptr; }

int bs_pool_set_undistort(bs_pool *p, int cam, const double K[9], const double *dist, int n_dist)
{
    if (!p) return BS_ERR_INVALID_ARG;
    if (p->next_frame != 0) return pool_fail(p, BS_ERR_INVALID_ARG, "bs_pool_set_undistort: set the lens models before the first frame");
    for (Worker *w : p->w) {
        const int rc = bs_set_undistort(w->h, cam, K, dist, n_dist);
        if (rc != BS_OK) return pool_fail(p, rc, bs_last_error(w->h));
    }
    return BS_OK;
}

int bs_pool_submit(bs_pool *p, const uint8_t *const *frames, const size_t *pitch, uint8_t *out, size_t out_pitch,
                   int64_t *frame_index)
{
    if (!p || !frames || !pitch || !out) return pool_fail(p, BS_ERR_INVALID_ARG, "bs_pool_submit: NULL argument");
    const int64_t f = p->next_frame;
    Worker *w = p->w[(size_t)(f % p->n)];
    Job j;
    j.frame = f;
    j.frames.assign(frames, frames + p->n_cams);
    j.pitch.assign(pitch, pitch + p->n_cams);
    j.out = out;
    j.out_pitch = out_pitch;
    {
        std::unique_lock<std::mutex> lk(w->m);
        if (w->rc != BS_OK) return pool_fail(p, w->rc, w->err);
        // at most `depth` frames per GPU are outstanding: the caller's buffers are the pipeline's storage
        w->cv.wait(lk, [&] { return (int64_t)w->queue.size() + (int64_t)w->inflight.size() < p->depth || w->rc != BS_OK; });
        if (w->rc != BS_OK) return pool_fail(p, w->rc, w->err);
        w->queue.push_back(std::move(j));
        w->last_submitted = f;
    }
    w->cv.notify_all();
    p->next_frame++;
    if (frame_index) *frame_index = f;
    return BS_OK;
}

int bs_pool_wait(bs_pool *p, int64_t frame_index)
{
    if (!p) return BS_ERR_INVALID_ARG;
    if (frame_index < 0 || frame_index >= p->next_frame) return pool_fail(p, BS_ERR_BAD_TICKET, "bs_pool_wait: frame was never submitted");
    Worker *w = p->w[(size_t)(frame_index % p->n)];
    std::unique_lock<std::mutex> lk(w->m);
    if (frame_index > w->wanted) w->wanted = frame_index;
    w->cv.notify_all();
    w->cv.wait(lk, [&] { return w->done_upto >= frame_index; });
    if (w->rc != BS_OK) return pool_fail(p, w->rc, w->err);
    return BS_OK;
}

int bs_pool_sync(bs_pool *p)
{
    if (!p) return BS_ERR_INVALID_ARG;
    int rc = BS_OK;
    for (Worker *w : p->w) {
        std::unique_lock<std::mutex> lk(w->m);
        if (w->last_submitted > w->wanted) w->wanted = w->last_submitted;
        w->cv.notify_all();
        w->cv.wait(lk, [&] { return w->done_upto >= w->last_submitted; });
        if (w->rc != BS_OK && rc == BS_OK) rc = pool_fail(p, w->rc, w->err);
    }
    return rc;
}

}  // extern "C"
