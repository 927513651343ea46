// orbx_pool.cu -- the multi-GPU driver of the C ABI (include/orbx.h, orbx_pool_*): SURVEY.md section 8(e).
//
// Frames are independent (reference src/ORBextractor.cc:1043-1105 touches only its own image), so a batch is cut into chunks of
// consecutive frames and chunk c goes to device c mod G -- with chunk = 1 that is frame i -> GPU i mod G.  Every device has `lanes`
// host threads, each with its own extractor handle (= its own stream, device workspace and page-locked staging buffer), so that the
// host-to-device copy of one chunk, the kernels of another and the device-to-host copy of a third overlap on a device: the handles'
// staging buffers are the per-device pinned ring.  Results land in the caller's arrays at the frame's own index (per-frame output
// slots), so the output bytes do not depend on G, the chunk size or the lane count.  No collective, no inter-GPU traffic.
#include "orbx_common.cuh"
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace orbx;

namespace {

struct Job {
    const uint8_t* const* imgs = nullptr;
    int B = 0, w = 0, h = 0, cap = 0, chunk = 1;
    size_t stride = 0;
    orbx_kp* kps = nullptr;
    uint8_t* desc = nullptr;
    int* n = nullptr;
    bool planOnly = false;
};

struct Worker {
    int device = 0, slot = 0;            // slot: index among the workers of ALL devices that serve the same chunk residue
    orbx_extractor* ex = nullptr;
    std::thread th;
};

}  // namespace

struct orbx_pool {
    std::vector<int> devices;
    int lanes = 1, chunk = 1;
    std::vector<Worker> workers;
    std::mutex mu;
    std::condition_variable cvStart, cvDone;
    unsigned long long generation = 0;   // bumped per job; workers run each generation once
    int pending = 0;
    bool quit = false;
    Job job;
    int status = ORBX_OK;
    std::string error;
    int planCap = 0;
    std::atomic<int> nextOfDevice[64];   // per device: how many of ITS chunks (d, d + G, d + 2G, ...) have been claimed in the current job
};

namespace {

// chunks of device d: c = d, d + G, d + 2G, ...
int run_job(orbx_pool* p, Worker& wk, const Job& J)
{
    const int G = (int)p->devices.size();
    int dIdx = 0;
    for (int i = 0; i < G; i++)
        if (p->devices[i] == wk.device) { dIdx = i; break; }
    if (J.planOnly) {
        // plan the shape on this handle (allocates its workspace once) by extracting one blank frame
        std::vector<uint8_t> blank((size_t)J.w * J.h, 0);
        std::vector<orbx_kp> k(1);
        std::vector<uint8_t> d(32);
        int n = 0;
        const int rc = orbx_extract(wk.ex, blank.data(), J.w, J.h, (size_t)J.w, k.data(), d.data(), 1, &n);
        return rc == ORBX_ERR_CAPACITY ? ORBX_OK : rc;
    }
    const int nchunks = (J.B + J.chunk - 1) / J.chunk;
    // the lanes of a device claim that device's chunks as they become free (which lane computes a chunk does not matter: the
    // results go to the frames' own slots)
    for (;;) {
        const int c = dIdx + G * p->nextOfDevice[dIdx].fetch_add(1, std::memory_order_relaxed);
        if (c >= nchunks) break;
        const int c0 = c * J.chunk, cb = std::min(J.chunk, J.B - c0);
        const int rc = orbx_extract_batch(wk.ex, J.imgs + c0, cb, J.w, J.h, J.stride, J.kps ? J.kps + (size_t)c0 * J.cap : nullptr,
                                          J.desc ? J.desc + (size_t)c0 * J.cap * 32 : nullptr, J.cap, J.n ? J.n + c0 : nullptr);
        if (rc != ORBX_OK) return rc;
    }
    return ORBX_OK;
}

void worker_main(orbx_pool* p, int index)
{
    Worker& wk = p->workers[index];
    cudaSetDevice(wk.device);
    unsigned long long seen = 0;
    for (;;) {
        Job J;
        {
            std::unique_lock<std::mutex> lk(p->mu);
            p->cvStart.wait(lk, [&] { return p->quit || p->generation != seen; });
            if (p->quit) return;
            seen = p->generation;
            J = p->job;
        }
        const int rc = run_job(p, wk, J);
        {
            std::lock_guard<std::mutex> lk(p->mu);
            if (rc != ORBX_OK && p->status == ORBX_OK) {
                p->status = rc;
                p->error = std::string("device ") + std::to_string(wk.device) + ": " + orbx_last_error();
            }
            if (--p->pending == 0) p->cvDone.notify_all();
        }
    }
}

int dispatch(orbx_pool* p, const Job& J)
{
    std::unique_lock<std::mutex> lk(p->mu);
    p->job = J;
    p->status = ORBX_OK;
    p->error.clear();
    p->pending = (int)p->workers.size();
    for (size_t d = 0; d < p->devices.size(); d++) p->nextOfDevice[d].store(0, std::memory_order_relaxed);
    p->generation++;
    p->cvStart.notify_all();
    p->cvDone.wait(lk, [&] { return p->pending == 0; });
    if (p->status != ORBX_OK) set_error("orbx_pool: %s", p->error.c_str());
    return p->status;
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
            wk.slot = k;
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
        std::lock_guard<std::mutex> lk(p->mu);
        p->quit = true;
    }
    p->cvStart.notify_all();
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
    Job J;
    J.w = w; J.h = h; J.planOnly = true;
    const int rc = dispatch(p, J);
    if (rc) return rc;
    p->planCap = orbx_max_keypoints(p->workers[0].ex);
    if (max_keypoints) *max_keypoints = p->planCap;
    return ORBX_OK;
}

int orbx_pool_extract_batch(orbx_pool* p, const uint8_t* const* imgs, int B, int w, int h, size_t stride, orbx_kp* kps, uint8_t* desc, int cap,
                            int* n)
{
    if (!p || !imgs || B < 0 || cap < 0) { set_error("orbx_pool_extract_batch: bad argument"); return ORBX_ERR_ARG; }
    if (B == 0) return ORBX_OK;
    if (w <= 0 || h <= 0) {            // empty images: the reference returns without touching its outputs (ORBextractor.cc:1046-1047)
        if (n) for (int b = 0; b < B; b++) n[b] = 0;
        return ORBX_OK;
    }
    Job J;
    J.imgs = imgs; J.B = B; J.w = w; J.h = h; J.stride = stride; J.kps = kps; J.desc = desc; J.cap = cap; J.n = n; J.chunk = p->chunk;
    return dispatch(p, J);
}

}  // extern "C"
