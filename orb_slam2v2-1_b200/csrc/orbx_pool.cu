// orbx_pool.cu -- the multi-GPU driver of the C ABI (include/orbx.h, orbx_pool_*): SURVEY.md section 8(e).
//
// Frames are independent (reference src/ORBextractor.cc:1043-1105 touches only its own image), so a batch is cut into chunks of
// consecutive frames and chunk c goes to device c mod G -- with chunk = 1 that is frame i -> GPU i mod G.  Every device has `lanes`
// host threads, each with its own extractor handle (= its own stream, device workspace and page-locked staging buffer), so that the
// host-to-device copy of one chunk, the kernels of another and the device-to-host copy of a third overlap on a device: the handles'
// staging buffers are the per-device pinned ring.  Results land in the caller's arrays at the frame's own index (per-frame output
// slots), so the output bytes do not depend on G, the chunk size or the lane count.  No collective, no inter-GPU traffic.
// Batches are queued (submit / wait): a lane that finds no chunk left in the oldest batch starts on the next one, so a stream of
// batches keeps the copy engines busy across batch boundaries.
#include "orbx_common.cuh"
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace orbx;

namespace {

// One queued batch.  The lanes of a device claim that device's chunks (d, d + G, d + 2G, ...) with an atomic counter; a lane that finds
// no chunk left in the oldest batch moves on to the next one, so batches follow each other without a barrier.
struct Batch {
    std::vector<const uint8_t*> imgs;    // the caller's pointer array, copied: it need not outlive the submit call
    int B = 0, w = 0, h = 0, cap = 0, chunk = 1, nchunks = 0;
    size_t stride = 0;
    orbx_kp* kps = nullptr;
    uint8_t* desc = nullptr;
    int* n = nullptr;
    unsigned long long id = 0;
    std::atomic<int> nextOfDevice[64];   // per device: how many of ITS chunks have been claimed
    int remaining = 0;                   // chunks not yet completed (under the pool mutex)
    int status = ORBX_OK;
    std::string error;
    bool done = false;
};

struct Worker {
    int device = 0, dIdx = 0;
    orbx_extractor* ex = nullptr;
    std::thread th;
};

}  // namespace

struct orbx_pool {
    std::vector<int> devices;
    int lanes = 1, chunk = 1;
    std::vector<Worker> workers;
    std::mutex mu;
    std::condition_variable cvWork, cvDone;
    bool quit = false;
    std::deque<std::shared_ptr<Batch>> queue;                          // batches with unclaimed or running chunks, oldest first
    std::map<unsigned long long, std::shared_ptr<Batch>> tickets;      // submitted, not yet waited for
    unsigned long long nextId = 1;
    // shape planning: every lane runs once (a broadcast, not a queue entry)
    unsigned long long planGeneration = 0;
    int planW = 0, planH = 0, planPending = 0, planStatus = ORBX_OK;
    std::string planError;
    int planCap = 0;
};

namespace {

// plan the shape on this handle (allocates its workspace once) by extracting one blank frame
int plan_lane(Worker& wk, int w, int h)
{
    std::vector<uint8_t> blank((size_t)w * h, 0);
    std::vector<orbx_kp> k(1);
    std::vector<uint8_t> d(32);
    int n = 0;
    const int rc = orbx_extract(wk.ex, blank.data(), w, h, (size_t)w, k.data(), d.data(), 1, &n);
    return rc == ORBX_ERR_CAPACITY ? ORBX_OK : rc;
}

// the next chunk of this lane's device in the oldest batch that still has one; false when there is none (called under the mutex)
bool claim(orbx_pool* p, const Worker& wk, std::shared_ptr<Batch>& out, int& chunkIndex)
{
    const int G = (int)p->devices.size();
    for (const std::shared_ptr<Batch>& b : p->queue) {
        if (wk.dIdx + G * b->nextOfDevice[wk.dIdx].load(std::memory_order_relaxed) >= b->nchunks) continue;
        const int c = wk.dIdx + G * b->nextOfDevice[wk.dIdx].fetch_add(1, std::memory_order_relaxed);
        if (c >= b->nchunks) continue;
        out = b;
        chunkIndex = c;
        return true;
    }
    return false;
}

void worker_main(orbx_pool* p, int index)
{
    Worker& wk = p->workers[index];
    cudaSetDevice(wk.device);
    unsigned long long seenPlan = 0;
    std::unique_lock<std::mutex> lk(p->mu);
    for (;;) {
        std::shared_ptr<Batch> b;
        int c = 0;
        bool plan = false;
        p->cvWork.wait(lk, [&] {
            if (p->quit) return true;
            if (p->planGeneration != seenPlan) { plan = true; return true; }
            return claim(p, wk, b, c);
        });
        if (p->quit) return;
        if (plan) {
            seenPlan = p->planGeneration;
            const int w = p->planW, h = p->planH;
            lk.unlock();
            const int rc = plan_lane(wk, w, h);
            lk.lock();
            if (rc != ORBX_OK && p->planStatus == ORBX_OK) {
                p->planStatus = rc;
                p->planError = std::string("device ") + std::to_string(wk.device) + ": " + orbx_last_error();
            }
            if (--p->planPending == 0) p->cvDone.notify_all();
            continue;
        }
        lk.unlock();
        // which lane computes a chunk does not matter: the results go to the frames' own slots
        const int c0 = c * b->chunk, cb = std::min(b->chunk, b->B - c0);
        const int rc = orbx_extract_batch(wk.ex, b->imgs.data() + c0, cb, b->w, b->h, b->stride, b->kps ? b->kps + (size_t)c0 * b->cap : nullptr,
                                          b->desc ? b->desc + (size_t)c0 * b->cap * 32 : nullptr, b->cap, b->n ? b->n + c0 : nullptr);
        std::string err;
        if (rc != ORBX_OK) err = std::string("device ") + std::to_string(wk.device) + ": " + orbx_last_error();
        lk.lock();
        if (rc != ORBX_OK && b->status == ORBX_OK) { b->status = rc; b->error = err; }
        if (--b->remaining == 0) {
            b->done = true;
            while (!p->queue.empty() && p->queue.front()->done) p->queue.pop_front();
            p->cvDone.notify_all();
        }
    }
}

}  // namespace

extern "C" {

int orbx_pool_create(const int* devices, int ndevices, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                     int chunk, int lanes, orbx_pool** out)
{
    if (!out) { set_error("orbx_pool_create: out is null"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (ndevices < 0 || chunk < 1 || lanes < 1 || lanes > 16) { set_error("orbx_pool_create: need chunk >= 1 and 1 <= lanes <= 16"); return ORBX_ERR_ARG; }
    int present = 0;
    ORBX_CUDA(cudaGetDeviceCount(&present));
    std::vector<int> devs;
    if (ndevices == 0 || !devices)
        for (int d = 0; d < present; d++) devs.push_back(d);       // every visible GPU
    else
        devs.assign(devices, devices + ndevices);
    if (devs.empty()) { set_error("orbx_pool_create: no CUDA device"); return ORBX_ERR_CUDA; }
    if (devs.size() > 64) { set_error("orbx_pool_create: at most 64 devices"); return ORBX_ERR_ARG; }
    for (size_t i = 0; i < devs.size(); i++) {
        if (devs[i] < 0 || devs[i] >= present) { set_error("orbx_pool_create: device %d not present (%d CUDA devices)", devs[i], present); return ORBX_ERR_CUDA; }
        for (size_t j = 0; j < i; j++)
            if (devs[j] == devs[i]) { set_error("orbx_pool_create: device %d listed twice", devs[i]); return ORBX_ERR_ARG; }
    }
    orbx_pool* p = new orbx_pool;
    p->devices = devs;
    p->lanes = lanes;
    p->chunk = chunk;
    p->workers.resize(devs.size() * (size_t)lanes);
    int prev = 0;
    cudaGetDevice(&prev);
    for (size_t d = 0; d < devs.size(); d++)
        for (int k = 0; k < lanes; k++) {
            Worker& wk = p->workers[d * lanes + k];
            wk.device = devs[d];
            wk.dIdx = (int)d;
            const int rc = orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, devs[d], chunk, &wk.ex);
            if (rc != ORBX_OK) {
                for (Worker& w2 : p->workers) orbx_destroy(w2.ex);
                delete p;
                cudaSetDevice(prev);
                return rc;
            }
        }
    cudaSetDevice(prev);
    for (size_t i = 0; i < p->workers.size(); i++) p->workers[i].th = std::thread(worker_main, p, (int)i);
    *out = p;
    return ORBX_OK;
}

void orbx_pool_destroy(orbx_pool* p)
{
    if (!p) return;
    {
        std::unique_lock<std::mutex> lk(p->mu);
        p->cvDone.wait(lk, [&] { return p->queue.empty() && p->planPending == 0; });      // batches still in flight run to their end first
        p->quit = true;
    }
    p->cvWork.notify_all();
    for (Worker& wk : p->workers)
        if (wk.th.joinable()) wk.th.join();
    for (Worker& wk : p->workers) orbx_destroy(wk.ex);
    delete p;
}

int orbx_pool_devices(const orbx_pool* p) { return p ? (int)p->devices.size() : 0; }

int orbx_pool_set_fma_mode(orbx_pool* p, int fma_mode)
{
    if (!p) { set_error("null pool"); return ORBX_ERR_ARG; }
    for (Worker& wk : p->workers) {
        const int rc = orbx_set_fma_mode(wk.ex, fma_mode);
        if (rc) return rc;
    }
    return ORBX_OK;
}

int orbx_pool_plan(orbx_pool* p, int w, int h, int* max_keypoints)
{
    if (!p || w <= 0 || h <= 0) { set_error("orbx_pool_plan: bad argument"); return ORBX_ERR_ARG; }
    std::unique_lock<std::mutex> lk(p->mu);
    p->cvDone.wait(lk, [&] { return p->queue.empty() && p->planPending == 0; });      // not while batches are in flight
    p->planW = w; p->planH = h;
    p->planStatus = ORBX_OK;
    p->planError.clear();
    p->planPending = (int)p->workers.size();
    p->planGeneration++;
    p->cvWork.notify_all();
    p->cvDone.wait(lk, [&] { return p->planPending == 0; });
    if (p->planStatus != ORBX_OK) { set_error("orbx_pool: %s", p->planError.c_str()); return p->planStatus; }
    p->planCap = orbx_max_keypoints(p->workers[0].ex);
    if (max_keypoints) *max_keypoints = p->planCap;
    return ORBX_OK;
}

int orbx_pool_submit_batch(orbx_pool* p, const uint8_t* const* imgs, int B, int w, int h, size_t stride, orbx_kp* kps, uint8_t* desc, int cap,
                           int* n, unsigned long long* ticket)
{
    if (!p || !imgs || B < 0 || cap < 0 || !ticket) { set_error("orbx_pool_submit_batch: bad argument"); return ORBX_ERR_ARG; }
    *ticket = 0;
    if (B == 0) return ORBX_OK;
    if (w <= 0 || h <= 0) {            // empty images: the reference returns without touching its outputs (ORBextractor.cc:1046-1047)
        if (n) for (int b = 0; b < B; b++) n[b] = 0;
        return ORBX_OK;
    }
    std::shared_ptr<Batch> b = std::make_shared<Batch>();
    b->imgs.assign(imgs, imgs + B);
    b->B = B; b->w = w; b->h = h; b->stride = stride; b->kps = kps; b->desc = desc; b->cap = cap; b->n = n; b->chunk = p->chunk;
    b->nchunks = (B + p->chunk - 1) / p->chunk;
    b->remaining = b->nchunks;
    for (int d = 0; d < 64; d++) b->nextOfDevice[d].store(0, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> lk(p->mu);
        b->id = p->nextId++;
        p->queue.push_back(b);
        p->tickets[b->id] = b;
        *ticket = b->id;
    }
    p->cvWork.notify_all();
    return ORBX_OK;
}

int orbx_pool_wait(orbx_pool* p, unsigned long long ticket)
{
    if (!p) { set_error("null pool"); return ORBX_ERR_ARG; }
    if (ticket == 0) return ORBX_OK;                                   // an empty batch
    std::unique_lock<std::mutex> lk(p->mu);
    std::map<unsigned long long, std::shared_ptr<Batch>>::iterator it = p->tickets.find(ticket);
    if (it == p->tickets.end()) { set_error("orbx_pool_wait: unknown ticket (already waited for?)"); return ORBX_ERR_ARG; }
    std::shared_ptr<Batch> b = it->second;
    p->cvDone.wait(lk, [&] { return b->done; });
    p->tickets.erase(ticket);
    if (b->status != ORBX_OK) set_error("orbx_pool: %s", b->error.c_str());
    return b->status;
}

int orbx_pool_extract_batch(orbx_pool* p, const uint8_t* const* imgs, int B, int w, int h, size_t stride, orbx_kp* kps, uint8_t* desc, int cap,
                            int* n)
{
    unsigned long long ticket = 0;
    const int rc = orbx_pool_submit_batch(p, imgs, B, w, h, stride, kps, desc, cap, n, &ticket);
    return rc != ORBX_OK ? rc : orbx_pool_wait(p, ticket);
}

}  // extern "C"
