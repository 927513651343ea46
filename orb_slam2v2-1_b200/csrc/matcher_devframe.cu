// matcher_devframe.cu -- a Frame that lives on the device: undistorted keypoints, descriptors and the
// 64 x 48 feature grid of Frame::AssignFeaturesToGrid (reference src/Frame.cc:230-245, 397-407) stay in HBM between the extraction and
// the windowed searches of Tracking, so a search moves only the MapPoint side.  The grid is a counting sort ON THE DEVICE that keeps
// insertion order (a16); a search is ONE launch -- window enumeration in the reference's order (GetFeaturesInArea, :342-395) fused with the
// Hamming distances, a fixed number of candidate slots per query -- one copy back, and the greedy replay of matcher_frame.cu.
#include "matcher_internal.cuh"
#include <limits.h>
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>

struct orbm_frame {
    int device = 0, n = 0;
    orbx::GridParams gp{};
    float minX = 0, maxX = 0, minY = 0, maxY = 0;
    uint8_t* d_base = nullptr;          // one allocation: keypoints | descriptors | cellStart | entries
    orbx_kp* d_kps = nullptr;
    uint4* d_desc = nullptr;
    int* d_cellStart = nullptr;         // GRID_COLS * GRID_ROWS + 1, column-major cells (cell = ix * 48 + iy)
    int* d_entries = nullptr;
    std::vector<orbx_kp> h_kps;         // host mirror of the keypoints: the replays read octave / angle / position
};

namespace orbx {

constexpr int GRID_CELLS = GRID_COLS * GRID_ROWS;
constexpr int FRAME_MAX_KEYS = 8192;                    // keys of the sort live in shared memory (32 KB)
constexpr uint32_t CELL_NONE = 0xFFFFu;                 // keypoint outside the grid (PosInGrid returns false, Frame.cc:403-404)

// Frame::AssignFeaturesToGrid as a stable counting sort: key = cell << 16 | keypoint index, bitonic sort in shared memory (ascending
// keys = cells in column-major order, insertion order inside a cell), then cellStart[c] = first key of a cell >= c by binary search.
__global__ void __launch_bounds__(1024) k_grid_build(const orbx_kp* __restrict__ kps, int n, int npad, GridParams gp, int* __restrict__ cellStart,
                                                     int* __restrict__ entries)
{
    extern __shared__ uint32_t key[];
    const int tid = threadIdx.x, T = blockDim.x;
    for (int i = tid; i < npad; i += T) {
        uint32_t k = 0xFFFFFFFFu;
        if (i < n) {
            // PosInGrid (:397-407): round((x - mnMinX) * mfGridElementWidthInv), float, round half away from zero
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
            const uint32_t cell = (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? CELL_NONE : (uint32_t)(px * GRID_ROWS + py);
            k = (cell << 16) | (uint32_t)i;
        }
        key[i] = k;
    }
    __syncthreads();
    for (int size = 2; size <= npad; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = tid; i < (npad >> 1); i += T) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint32_t a = key[lo], b = key[hi];
                if ((a > b) == up) { key[lo] = b; key[hi] = a; }
            }
            __syncthreads();
        }
    for (int c = tid; c <= GRID_CELLS; c += T) {
        const uint32_t want = (uint32_t)c << 16;
        int lo = 0, hi = n;                              // first position whose key >= want
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (key[mid] < want) lo = mid + 1; else hi = mid;
        }
        cellStart[c] = lo;
    }
    for (int i = tid; i < n; i += T) entries[i] = (int)(key[i] & 0xFFFFu);
}

// One warp per query: GetFeaturesInArea(x, y, r, minLevel, maxLevel) in the reference's visiting order, the Hamming distance of
// every candidate, packed as distance << 16 | feature into the query's `cap` slots; count[q] = candidates found (may exceed cap:
// the caller then repeats the call with more slots).  qsel: optional indices into qdesc (queries that are features of another
// device frame).
__global__ void __launch_bounds__(256) k_window_scan(const orbx_kp* __restrict__ kps, const uint4* __restrict__ desc,
                                                     const int* __restrict__ cellStart, const int* __restrict__ entries, GridParams gp,
                                                     const float* __restrict__ qxy, const float* __restrict__ qr, const int* __restrict__ qlev2,
                                                     const uint4* __restrict__ qdesc, const int* __restrict__ qsel, int nq, int cap,
                                                     uint32_t* __restrict__ slots, int* __restrict__ count)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const float x = qxy[2 * qi], y = qxy[2 * qi + 1], r = qr[qi];
    const int minL = qlev2[2 * qi], maxL = qlev2[2 * qi + 1];
    const int qd = qsel ? qsel[qi] : qi;
    const uint4 qa = __ldg(&qdesc[2 * (size_t)qd]), qb = __ldg(&qdesc[2 * (size_t)qd + 1]);
    uint32_t* out = slots + (size_t)qi * cap;
    int total = 0;
    // (x - mnMinX - r) * mfGridElementWidthInv, float, left to right (Frame.cc:347-361)
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
    const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
    const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
    if (cx0 < GRID_COLS && cx1 >= 0 && cy0 < GRID_ROWS && cy1 >= 0) {
        const bool checkLevels = (minL > 0) || (maxL >= 0);
        for (int ix = cx0; ix <= cx1; ix++) {
            const int s = cellStart[ix * GRID_ROWS + cy0], e = cellStart[ix * GRID_ROWS + cy1 + 1];
            for (int c0 = s; c0 < e; c0 += 32) {
                const int c = c0 + lane;
                bool pass = false;
                int id = 0;
                if (c < e) {
                    id = entries[c];
                    const orbx_kp kp = kps[id];
                    pass = true;
                    if (checkLevels) {
                        if (kp.octave < minL) pass = false;
                        if (maxL >= 0 && kp.octave > maxL) pass = false;
                    }
                    if (!(fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r)) pass = false;
                }
                const unsigned m = __ballot_sync(0xFFFFFFFFu, pass);
                if (pass) {
                    const int pos = total + __popc(m & ((1u << lane) - 1));
                    if (pos < cap) {
                        const int d = hamming256(qa, qb, __ldg(&desc[2 * (size_t)id]), __ldg(&desc[2 * (size_t)id + 1]));
                        out[pos] = ((uint32_t)d << 16) | (uint32_t)id;
                    }
                }
                total += __popc(m);
            }
        }
    }
    if (lane == 0) count[qi] = total;
}

static int frame_alloc(orbm_frame* F, int n)
{
    const size_t oK = 0, oD = (sizeof(orbx_kp) * (size_t)std::max(n, 1) + 255) / 256 * 256, oS = oD + ((size_t)32 * std::max(n, 1) + 255) / 256 * 256,
                 oE = oS + (sizeof(int) * (GRID_CELLS + 1) + 255) / 256 * 256, total = oE + sizeof(int) * (size_t)std::max(n, 1);
    ORBX_CUDA(cudaMalloc(&F->d_base, total));
    F->d_kps = (orbx_kp*)(F->d_base + oK);
    F->d_desc = (uint4*)(F->d_base + oD);
    F->d_cellStart = (int*)(F->d_base + oS);
    F->d_entries = (int*)(F->d_base + oE);
    return ORBX_OK;
}

static int frame_build_grid(orbx_matcher* m, orbm_frame* F)
{
    int npad = 2;
    while (npad < F->n) npad <<= 1;
    const size_t smem = sizeof(uint32_t) * (size_t)npad;          // at most 32 KB (FRAME_MAX_KEYS)
    k_grid_build<<<1, 1024, smem, m->stream>>>(F->d_kps, F->n, npad, F->gp, F->d_cellStart, F->d_entries);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

static bool frame_args_ok(const char* who, orbx_matcher* m, int n, float minX, float maxX, float minY, float maxY, void* out)
{
    if (!m || !out || n < 0 || !(maxX > minX) || !(maxY > minY)) { set_error("%s: bad argument", who); return false; }
    if (n > FRAME_MAX_KEYS) { set_error("%s: %d keypoints, a device frame holds at most %d", who, n, FRAME_MAX_KEYS); return false; }
    return true;
}

static void frame_set_bounds(orbm_frame* F, float minX, float maxX, float minY, float maxY)
{
    F->minX = minX; F->maxX = maxX; F->minY = minY; F->maxY = maxY;
    F->gp.minX = minX;
    F->gp.minY = minY;
    F->gp.invW = (float)GRID_COLS / (float)(maxX - minX);        // mfGridElementWidthInv (Frame.cc:97-98)
    F->gp.invH = (float)GRID_ROWS / (float)(maxY - minY);
}

// The searches: queries packed by the caller into the arena (one upload), one launch, one copy back.  Returns the candidates as the
// CSR (off, idx, dist) that the replays of matcher_frame.cu take.
int devframe_window_scan(orbx_matcher* m, const orbm_frame* F, const float* qxy, const float* qr, const int* qlev2, const uint8_t* qdesc,
                         const orbm_frame* Fq, const int* qsel, int nq, std::vector<int>& off, std::vector<int>& idx, std::vector<int>& dist)
{
    off.assign((size_t)nq + 1, 0);
    idx.clear();
    dist.clear();
    if (nq == 0 || F->n == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    for (int cap = std::max(m->windowSlots, 64);; cap *= 2) {
        Carver cv(m->arena);
        const size_t oQ = cv.take(sizeof(float) * 2 * (size_t)nq), oR = cv.take(sizeof(float) * (size_t)nq), oL = cv.take(sizeof(int) * 2 * (size_t)nq),
                     oQD = cv.take(Fq ? 0 : (size_t)nq * 32), oQS = cv.take(qsel ? sizeof(int) * (size_t)nq : 0),
                     oC = cv.take(sizeof(int) * (size_t)nq), oS = cv.take(sizeof(uint32_t) * (size_t)nq * cap);
        int rc = m->arena.reserve(cv.total());
        if (!rc) rc = m->arena.mirror();
        if (rc) return rc;
        uint8_t* B = m->arena.base;
        Xfer x(m->arena);
        x.up(oQ, qxy, sizeof(float) * 2 * (size_t)nq);
        x.up(oR, qr, sizeof(float) * (size_t)nq);
        x.up(oL, qlev2, sizeof(int) * 2 * (size_t)nq);
        if (!Fq) x.up(oQD, qdesc, (size_t)nq * 32);
        if (qsel) x.up(oQS, qsel, sizeof(int) * (size_t)nq);
        ORBX_CUDA(x.send());
        k_window_scan<<<(nq + 7) / 8, 256, 0, s>>>(F->d_kps, F->d_desc, F->d_cellStart, F->d_entries, F->gp, (const float*)(B + oQ), (const float*)(B + oR),
                                                  (const int*)(B + oL), Fq ? Fq->d_desc : (const uint4*)(B + oQD), qsel ? (const int*)(B + oQS) : nullptr,
                                                  nq, cap, (uint32_t*)(B + oS), (int*)(B + oC));
        ORBX_CUDA(cudaGetLastError());
        m->lastLaunches = 1;
        // counts and slots are adjacent in the arena: one copy back through the page-locked mirror, then one wait
        ORBX_CUDA(cudaMemcpyAsync(m->arena.pin + oC, B + oC, (oS - oC) + sizeof(uint32_t) * (size_t)nq * cap, cudaMemcpyDeviceToHost, s));
        ORBX_CUDA(cudaStreamSynchronize(s));
        const int* cnt = (const int*)(m->arena.pin + oC);
        const uint32_t* slots = (const uint32_t*)(m->arena.pin + oS);
        int worst = 0, total = 0;
        for (int q = 0; q < nq; q++) { worst = std::max(worst, cnt[q]); total += cnt[q]; }
        if (worst > cap) {                     // a window with more candidates than slots: again with enough of them (rare)
            while (cap < worst) cap *= 2;
            cap /= 2;
            continue;
        }
        // remember what this call needed (with headroom), so that the next one of its kind fits at once; shrink slowly
        int need = 64;
        while (need < worst + worst / 4) need *= 2;
        m->windowSlots = need >= m->windowSlots ? need : std::max(need, m->windowSlots / 2);
        idx.resize(std::max(total, 1));
        dist.resize(std::max(total, 1));
        int o = 0;
        for (int q = 0; q < nq; q++) {
            off[q] = o;
            for (int c = 0; c < cnt[q]; c++, o++) { const uint32_t v = slots[(size_t)q * cap + c]; idx[o] = (int)(v & 0xFFFFu); dist[o] = (int)(v >> 16); }
        }
        off[nq] = o;
        return ORBX_OK;
    }
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbm_frame_create(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, int n, float minX, float maxX, float minY, float maxY, orbm_frame** out)
{
    if (!frame_args_ok("orbm_frame_create", m, n, minX, maxX, minY, maxY, out)) return ORBX_ERR_ARG;
    if (n > 0 && (!kps || !desc)) { set_error("orbm_frame_create: null keypoints or descriptors"); return ORBX_ERR_ARG; }
    *out = nullptr;
    ORBX_CUDA(cudaSetDevice(m->device));
    orbm_frame* F = new orbm_frame;
    F->device = m->device;
    F->n = n;
    frame_set_bounds(F, minX, maxX, minY, maxY);
    int rc = frame_alloc(F, n);
    if (rc) { delete F; return rc; }
    F->h_kps.assign(kps, kps + n);
    if (n > 0) {
        // both arrays through the matcher's page-locked mirror: one copy each way is what a pageable cudaMemcpy would stage anyway
        if (cudaMemcpyAsync(F->d_kps, kps, sizeof(orbx_kp) * (size_t)n, cudaMemcpyHostToDevice, m->stream) != cudaSuccess ||
            cudaMemcpyAsync(F->d_desc, desc, (size_t)32 * n, cudaMemcpyHostToDevice, m->stream) != cudaSuccess) {
            set_error("orbm_frame_create: upload failed");
            cudaFree(F->d_base);
            delete F;
            return ORBX_ERR_CUDA;
        }
    }
    rc = frame_build_grid(m, F);
    if (rc) { cudaFree(F->d_base); delete F; return rc; }
    ORBX_CUDA(cudaStreamSynchronize(m->stream));          // the caller's arrays are free again, the frame is ready
    *out = F;
    return ORBX_OK;
}

int orbm_frame_from_extractor(orbx_matcher* m, orbx_extractor* ex, int frame, float fx, float fy, float cx, float cy, const float* dist_coef,
                              int ndist, float minX, float maxX, float minY, float maxY, orbm_frame** out)
{
    if (!frame_args_ok("orbm_frame_from_extractor", m, 0, minX, maxX, minY, maxY, out)) return ORBX_ERR_ARG;
    if (!ex) { set_error("orbm_frame_from_extractor: null extractor"); return ORBX_ERR_ARG; }
    *out = nullptr;
    ORBX_CUDA(cudaSetDevice(m->device));
    const orbx_kp* dk = nullptr;
    const uint8_t* dd = nullptr;
    const int* dc = nullptr;
    int stride = 0;
    int rc = orbx_results_device(ex, &dk, &dd, &dc, &stride);
    if (rc) return rc;
    if (frame < 0 || frame >= orbx_last_batch(ex)) { set_error("orbm_frame_from_extractor: frame %d outside the last batch", frame); return ORBX_ERR_ARG; }
    if ((rc = orbx_sync(ex))) return rc;
    int n = 0;
    ORBX_CUDA(cudaMemcpy(&n, dc + frame, sizeof(int), cudaMemcpyDeviceToHost));
    if (n > FRAME_MAX_KEYS) { set_error("orbm_frame_from_extractor: %d keypoints, a device frame holds at most %d", n, FRAME_MAX_KEYS); return ORBX_ERR_ARG; }
    orbm_frame* F = new orbm_frame;
    F->device = m->device;
    F->n = n;
    frame_set_bounds(F, minX, maxX, minY, maxY);
    rc = frame_alloc(F, n);
    if (rc) { delete F; return rc; }
    cudaStream_t s = m->stream;
    bool ok = true;
    if (n > 0) {
        ok = cudaMemcpyAsync(F->d_kps, dk + (size_t)frame * stride, sizeof(orbx_kp) * (size_t)n, cudaMemcpyDeviceToDevice, s) == cudaSuccess &&
             cudaMemcpyAsync(F->d_desc, dd + (size_t)frame * stride * 32, (size_t)32 * n, cudaMemcpyDeviceToDevice, s) == cudaSuccess;
        if (ok) ok = undistort_keypoints_device(s, F->d_kps, n, fx, fy, cx, cy, dist_coef, ndist) == ORBX_OK;      // mvKeysUn (Frame.cc:419-449)
        if (ok) ok = frame_build_grid(m, F) == ORBX_OK;
        F->h_kps.resize(n);
        if (ok) ok = cudaMemcpyAsync(F->h_kps.data(), F->d_kps, sizeof(orbx_kp) * (size_t)n, cudaMemcpyDeviceToHost, s) == cudaSuccess;
    } else {
        ok = frame_build_grid(m, F) == ORBX_OK;
    }
    if (ok) ok = cudaStreamSynchronize(s) == cudaSuccess;
    if (!ok) {
        const cudaError_t e = cudaGetLastError();
        set_error("orbm_frame_from_extractor: device work failed (%s)", e != cudaSuccess ? cudaGetErrorString(e) : orbx_last_error());
        cudaFree(F->d_base);
        delete F;
        return ORBX_ERR_CUDA;
    }
    *out = F;
    return ORBX_OK;
}

void orbm_frame_destroy(orbm_frame* F)
{
    if (!F) return;
    cudaSetDevice(F->device);
    cudaFree(F->d_base);
    delete F;
}

int orbm_frame_size(const orbm_frame* F) { return F ? F->n : 0; }

int orbm_frame_keypoints(const orbm_frame* F, orbx_kp* out)
{
    if (!F || (F->n > 0 && !out)) { set_error("orbm_frame_keypoints: bad argument"); return ORBX_ERR_ARG; }
    if (F->n > 0) memcpy(out, F->h_kps.data(), sizeof(orbx_kp) * (size_t)F->n);
    return ORBX_OK;
}

int orbm_frame_grid(orbx_matcher* m, const orbm_frame* F, int* cell_start, int* entries)
{
    if (!m || !F || !cell_start || (F->n > 0 && !entries)) { set_error("orbm_frame_grid: bad argument"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(m->device));
    ORBX_CUDA(cudaMemcpy(cell_start, F->d_cellStart, sizeof(int) * (GRID_CELLS + 1), cudaMemcpyDeviceToHost));
    if (F->n > 0) ORBX_CUDA(cudaMemcpy(entries, F->d_entries, sizeof(int) * (size_t)F->n, cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

// ---- the tracking searches against a device frame: only the MapPoint side travels
int orbm_search_by_projection_dev(orbx_matcher* m, const orbm_frame* F, const float* u_right, const uint8_t* occupied_in, const float* sf, int nlevels,
                                  const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos, const uint8_t* mp_desc,
                                  const uint8_t* mp_has_obs, int nmp, float th, float nnratio, int* feature_match, int* nmatches)
{
    if (!m || !F || nmp < 0 || (F->n > 0 && (!u_right || !occupied_in || !feature_match)) || !sf || nlevels < 1 ||
        (nmp > 0 && (!proj_xy || !proj_xr || !level || !view_cos || !mp_desc || !mp_has_obs))) {
        set_error("orbm_search_by_projection_dev: bad argument");
        return ORBX_ERR_ARG;
    }
    const int n = F->n;
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    const bool bFactor = th != 1.0;
    std::vector<float> qr(nmp), rr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        if (level[i] < 0 || level[i] >= nlevels) { set_error("orbm_search_by_projection_dev: level[%d] = %d out of range", i, level[i]); return ORBX_ERR_ARG; }
        float r = view_cos[i] > 0.998 ? 2.5f : 4.0f;          // RadiusByViewingCos (ORBmatcher.cc:131-137)
        if (bFactor) r *= th;
        rr[i] = r;
        qr[i] = r * sf[level[i]];                              // :68
        qlev[2 * i] = level[i] - 1;
        qlev[2 * i + 1] = level[i];
    }
    std::vector<int> off, idx, dist;
    const int rc = devframe_window_scan(m, F, proj_xy, qr.data(), qlev.data(), mp_desc, nullptr, nullptr, nmp, off, idx, dist);
    if (rc) return rc;
    const int nm = replay_projection(F->h_kps.data(), u_right, occupied_in, n, sf, level, rr, proj_xr, mp_has_obs, nmp, nnratio, off, idx, dist, feature_match);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

int orbm_search_by_projection_frame_dev(orbx_matcher* m, const orbm_frame* F, const float* u_right, const uint8_t* occupied_in, const float* sf,
                                        int nlevels, const float* proj_xy, const float* proj_ur, const int* last_octave, const float* last_angle,
                                        const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                        int* feature_match, int* nmatches)
{
    if (!m || !F || nmp < 0 || (F->n > 0 && (!u_right || !occupied_in || !feature_match)) || !sf || nlevels < 1 || mode < 0 || mode > 2 ||
        (nmp > 0 && (!proj_xy || !proj_ur || !last_octave || !last_angle || !mp_desc || !mp_has_obs))) {
        set_error("orbm_search_by_projection_frame_dev: bad argument");
        return ORBX_ERR_ARG;
    }
    const int n = F->n;
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    std::vector<float> qr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        const int o = last_octave[i];
        if (o < 0 || o >= nlevels) { set_error("orbm_search_by_projection_frame_dev: last_octave[%d] = %d out of range", i, o); return ORBX_ERR_ARG; }
        qr[i] = th * sf[o];                                                        // ORBmatcher.cc:1383
        if (mode == 1) { qlev[2 * i] = o; qlev[2 * i + 1] = -1; }                   // bForward
        else if (mode == 2) { qlev[2 * i] = 0; qlev[2 * i + 1] = o; }               // bBackward
        else { qlev[2 * i] = o - 1; qlev[2 * i + 1] = o + 1; }
    }
    std::vector<int> off, idx, dist;
    const int rc = devframe_window_scan(m, F, proj_xy, qr.data(), qlev.data(), mp_desc, nullptr, nullptr, nmp, off, idx, dist);
    if (rc) return rc;
    const int nm = replay_projection_frame(F->h_kps.data(), u_right, occupied_in, n, qr, proj_ur, last_angle, mp_has_obs, nmp, checkOri, off, idx, dist, feature_match);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

int orbm_search_for_initialization_dev(orbx_matcher* m, const orbm_frame* F1, const orbm_frame* F2, float* prev_matched, int* matches12,
                                       int windowSize, float nnratio, int checkOri, int* nmatches)
{
    if (!m || !F1 || !F2 || (F1->n > 0 && (!prev_matched || !matches12))) { set_error("orbm_search_for_initialization_dev: bad argument"); return ORBX_ERR_ARG; }
    const int n1 = F1->n, n2 = F2->n;
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    // queries: level-0 keypoints of F1 only (ORBmatcher.cc:421-423), searched around vbPrevMatched in level 0 of F2 (:425); their
    // descriptors are read from F1's device copy through the index list
    std::vector<int> qmap, qlev;
    std::vector<float> qxy, qr;
    for (int i = 0; i < n1; i++)
        if (F1->h_kps[i].octave <= 0) {
            qmap.push_back(i);
            qxy.push_back(prev_matched[2 * i]);
            qxy.push_back(prev_matched[2 * i + 1]);
            qr.push_back((float)windowSize);
            qlev.push_back(0);
            qlev.push_back(0);
        }
    const int nq = (int)qmap.size();
    if (nq == 0 || n2 == 0) return ORBX_OK;
    std::vector<int> off, idx, dist;
    const int rc = devframe_window_scan(m, F2, qxy.data(), qr.data(), qlev.data(), nullptr, F1, qmap.data(), nq, off, idx, dist);
    if (rc) return rc;
    const int nm = replay_initialization(F1->h_kps.data(), n1, F2->h_kps.data(), n2, qmap, off, idx, dist, nnratio, checkOri, prev_matched, matches12);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

}  // extern "C"
