// matcher_frame.cu -- frame-level searches of the C ABI: the Frame grid + GetFeaturesInArea (reference src/Frame.cc:230-245,
// 342-407), ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) and Frame::ComputeStereoMatches
// (src/Frame.cc:481-655).  Candidate enumeration, Hamming distances and the SAD refinement run on the device; the
// order-dependent greedy bookkeeping of the reference (vMatchedDistance / vnMatches21 steals, rotation histogram, the
// median cut) is replayed sequentially on the host from the device results, which keeps indices bit-exact.
#include "matcher_internal.cuh"
#include <limits.h>
#include <math.h>
#include <string.h>
#include <algorithm>
#include <utility>
#include <vector>

namespace orbx {


// ------------------------------------------------------------------------------------------------ GetFeaturesInArea
// One warp per query.  The CSR grid is stored column-major (cell = ix * 48 + iy), i.e. in the reference's own loop
// order (ix outer, iy inner, insertion order inside a cell), so the cells of one grid column that the window touches
// are ONE contiguous entry range: a warp filters it 32 entries at a time and compacts with a ballot.
template <bool WRITE>
__global__ void __launch_bounds__(256) k_grid_query(const orbx_kp* __restrict__ kps, const int* __restrict__ cellStart,
                                                    const int* __restrict__ entries, GridParams gp, const float* __restrict__ qxy,
                                                    const float* __restrict__ qr, int nq, int minLevel, int maxLevel,
                                                    const int* __restrict__ qlevel, int qlevelStride, int* __restrict__ off, int* __restrict__ idx_out)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const float x = qxy[2 * qi], y = qxy[2 * qi + 1], r = qr[qi];
    // qlevel: per-query levels; qlevelStride 1 = (l, l), 2 = (minLevel, maxLevel) pairs; null = the scalar arguments
    const int minL = qlevel ? qlevel[qlevelStride * qi] : minLevel, maxL = qlevel ? qlevel[qlevelStride * qi + qlevelStride - 1] : maxLevel;
    int total = 0;
    const int base = WRITE ? off[qi] : 0;
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
                if (WRITE && pass) idx_out[base + total + __popc(m & ((1u << lane) - 1))] = id;
                total += __popc(m);
            }
        }
    }
    if (!WRITE && lane == 0) off[qi] = total;
}

// exclusive scan of a[0..n) in place, a[n] = total; one CTA
__global__ void __launch_bounds__(1024) k_scan_offsets(int* a, int n)
{
    __shared__ int wsum[33];
    const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + T - 1) / T;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; i++) s += a[i];
    int inc = s;
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int v = wsum[lane];
        int vi = v;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, vi, o);
            if (lane >= o) vi += t;
        }
        wsum[lane] = vi - v;
        if (lane == 31) wsum[32] = vi;
    }
    __syncthreads();
    int off = wsum[wid] + inc - s;
    for (int i = b; i < e; i++) { const int t = a[i]; a[i] = off; off += t; }
    if (tid == 0) a[n] = wsum[32];
}

__global__ void __launch_bounds__(256) k_cand_dist(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t,
                                                   const int* __restrict__ off, const int* __restrict__ idx, int* __restrict__ dist)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const uint4 qa = __ldg(&q[2 * qi]), qb = __ldg(&q[2 * qi + 1]);
    for (int c = off[qi] + lane; c < off[qi + 1]; c += 32) {
        const int j = idx[c];
        dist[c] = hamming256(qa, qb, __ldg(&t[2 * (size_t)j]), __ldg(&t[2 * (size_t)j + 1]));
    }
}

// Frame::AssignFeaturesToGrid / PosInGrid (Frame.cc:230-245, 397-407) as a stable counting sort into CSR
void build_grid_host(const orbx_kp* kps, int n, float minX, float maxX, float minY, float maxY, GridParams& gp,
                            std::vector<int>& start, std::vector<int>& entries)
{
    gp.minX = minX;
    gp.minY = minY;
    gp.invW = (float)GRID_COLS / (float)(maxX - minX);
    gp.invH = (float)GRID_ROWS / (float)(maxY - minY);
    std::vector<int> cell(n);
    start.assign(GRID_COLS * GRID_ROWS + 1, 0);
    for (int i = 0; i < n; i++) {
        const int px = (int)roundf((kps[i].x - minX) * gp.invW), py = (int)roundf((kps[i].y - minY) * gp.invH);
        cell[i] = (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
        if (cell[i] >= 0) start[cell[i] + 1]++;
    }
    for (int c = 0; c < GRID_COLS * GRID_ROWS; c++) start[c + 1] += start[c];
    entries.assign(std::max(start.back(), 1), 0);
    std::vector<int> fill(start.begin(), start.end() - 1);
    for (int i = 0; i < n; i++)
        if (cell[i] >= 0) entries[fill[cell[i]]++] = i;
}

static inline size_t idx_bytes(int total) { return (sizeof(int) * (size_t)std::max(total, 1) + 255) / 256 * 256; }

// Device-side candidate enumeration.  On return the CSR lives on the device: *d_off in arena (nq+1 ints),
// *d_idx in arena2 (total ints); *total holds the candidate count.  d_kps: device keypoints of the searched frame.
static int grid_candidates_dev(orbx_matcher* m, const orbx_kp* d_kps, const int* d_cellStart, const int* d_entries, const GridParams& gp,
                               const float* d_qxy, const float* d_qr, int nq, int minLevel, int maxLevel, const int* d_qlevel,
                               int* d_off, int nArrays, int** d_idx, int* total, int qlevelStride = 1)
{
    const int blocks = (nq + 7) / 8;
    k_grid_query<false><<<blocks, 256, 0, m->stream>>>(d_kps, d_cellStart, d_entries, gp, d_qxy, d_qr, nq, minLevel, maxLevel, d_qlevel, qlevelStride, d_off, nullptr);
    k_scan_offsets<<<1, 1024, 0, m->stream>>>(d_off, nq);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(total, d_off + nq, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    int rc = m->arena2.reserve(idx_bytes(*total) * (size_t)nArrays);     // nArrays int arrays of `total` entries, d_idx first
    if (rc) return rc;
    *d_idx = (int*)m->arena2.base;
    k_grid_query<true><<<blocks, 256, 0, m->stream>>>(d_kps, d_cellStart, d_entries, gp, d_qxy, d_qr, nq, minLevel, maxLevel, d_qlevel, qlevelStride, d_off, *d_idx);
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 3;
    return ORBX_OK;
}


}  // namespace orbx

using namespace orbx;

extern "C" {

int orbm_grid_candidates(orbx_matcher* m, const orbx_kp* kps, int n, float minX, float maxX, float minY, float maxY,
                         const float* qxy, int nq, const float* qr, int minLevel, int maxLevel, const int* q_level,
                         int* cand_off, int* cand_idx, int cap)
{
    if (!m || n < 0 || nq < 0 || (n > 0 && !kps) || (nq > 0 && (!qxy || !qr)) || !cand_off || !(maxX > minX) || !(maxY > minY) || cap < 0) {
        set_error("orbm_grid_candidates: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nq == 0) { cand_off[0] = 0; return ORBX_OK; }
    ORBX_CUDA(cudaSetDevice(m->device));
    GridParams gp;
    std::vector<int> start, entries;
    build_grid_host(kps, n, minX, maxX, minY, maxY, gp, start, entries);
    Carver cv(m->arena);
    const size_t oK = cv.take(sizeof(orbx_kp) * (size_t)std::max(n, 1)), oS = cv.take(sizeof(int) * start.size()),
                 oE = cv.take(sizeof(int) * entries.size()), oQ = cv.take(sizeof(float) * 2 * (size_t)nq),
                 oR = cv.take(sizeof(float) * (size_t)nq), oL = cv.take(sizeof(int) * (size_t)nq), oO = cv.take(sizeof(int) * ((size_t)nq + 1));
    int rc = m->arena.reserve(cv.total());
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    if (n > 0) ORBX_CUDA(cudaMemcpyAsync(B + oK, kps, sizeof(orbx_kp) * (size_t)n, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oS, start.data(), sizeof(int) * start.size(), cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oE, entries.data(), sizeof(int) * entries.size(), cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oQ, qxy, sizeof(float) * 2 * (size_t)nq, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oR, qr, sizeof(float) * (size_t)nq, cudaMemcpyHostToDevice, s));
    if (q_level) ORBX_CUDA(cudaMemcpyAsync(B + oL, q_level, sizeof(int) * (size_t)nq, cudaMemcpyHostToDevice, s));
    int* d_idx = nullptr;
    int total = 0;
    rc = grid_candidates_dev(m, (const orbx_kp*)(B + oK), (const int*)(B + oS), (const int*)(B + oE), gp, (const float*)(B + oQ),
                             (const float*)(B + oR), nq, minLevel, maxLevel, q_level ? (const int*)(B + oL) : nullptr, (int*)(B + oO), 1,
                             &d_idx, &total);
    if (rc) return rc;
    ORBX_CUDA(cudaMemcpyAsync(cand_off, B + oO, sizeof(int) * ((size_t)nq + 1), cudaMemcpyDeviceToHost, s));
    const int ncopy = std::min(total, cap);
    if (ncopy > 0 && cand_idx) ORBX_CUDA(cudaMemcpyAsync(cand_idx, d_idx, sizeof(int) * (size_t)ncopy, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    if (total > cap) { set_error("orbm_grid_candidates: %d candidates, capacity %d", total, cap); return ORBX_ERR_CAPACITY; }
    return ORBX_OK;
}

}  // extern "C"

// ================================================================================================ SearchForInitialization
namespace orbx {

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1603-1644) on bin sizes
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// Sequential replay of the greedy loop of SearchForInitialization (src/ORBmatcher.cc:427-483, :485-517): the vMatchedDistance skip
// and the steals depend on visit order.  off / idx / dist: candidates of query q (feature qmap[q] of frame 1) in GetFeaturesInArea order.
int replay_initialization(const orbx_kp* k1, int n1, const orbx_kp* k2, int n2, const std::vector<int>& qmap, const std::vector<int>& off,
                          const std::vector<int>& idx, const std::vector<int>& dist, float nnratio, int checkOri, float* prev_matched,
                          int* matches12)
{
    const int nq = (int)qmap.size();
    int nm = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> matchedDist(n2, INT_MAX), m21(n2, -1);
    for (int q = 0; q < nq; q++) {
        const int i1 = qmap[q];
        if (off[q + 1] == off[q]) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int c = off[q]; c < off[q + 1]; c++) {
            const int i2 = idx[c], d = dist[c];
            if (matchedDist[i2] <= d) continue;
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestIdx2 = i2; }
            else if (d < bestDist2) bestDist2 = d;
        }
        if (bestDist <= TH_LOW && bestDist < (float)bestDist2 * nnratio) {
            if (m21[bestIdx2] >= 0) { matches12[m21[bestIdx2]] = -1; nm--; }
            matches12[i1] = bestIdx2;
            m21[bestIdx2] = i1;
            matchedDist[bestIdx2] = bestDist;
            nm++;
            if (checkOri) {
                float rot = k1[i1].angle - k2[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);       // factor = 1/HISTO_LENGTH as in the reference (:413), kept as is
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(i1);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int i1 : rotHist[i])
                if (matches12[i1] >= 0) { matches12[i1] = -1; nm--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)          // update vbPrevMatched (:515-517)
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = k2[matches12[i1]].x; prev_matched[2 * i1 + 1] = k2[matches12[i1]].y; }
    return nm;
}

// Sequential replay of SearchByProjection(Frame, MapPoints) (src/ORBmatcher.cc:50-126): occupancy changes as points are assigned.
int replay_projection(const orbx_kp* kps, const float* u_right, const uint8_t* occupied_in, int n, const float* sf, const int* level,
                      const std::vector<float>& rr, const float* proj_xr, const uint8_t* mp_has_obs, int nmp, float nnratio,
                      const std::vector<int>& off, const std::vector<int>& idx, const std::vector<int>& dist, int* feature_match)
{
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    int nm = 0;
    for (int i = 0; i < nmp; i++) {
        if (off[i + 1] == off[i]) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int c = off[i]; c < off[i + 1]; c++) {
            const int j = idx[c];
            if (occupied[j]) continue;
            if (u_right[j] > 0) {
                const float er = fabsf(proj_xr[i] - u_right[j]);
                if (er > rr[i] * sf[level[i]]) continue;
            }
            const int d = dist[c];
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestLevel2 = bestLevel; bestLevel = kps[j].octave; bestIdx = j; }
            else if (d < bestDist2) { bestLevel2 = kps[j].octave; bestDist2 = d; }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            feature_match[bestIdx] = i;
            occupied[bestIdx] = mp_has_obs[i];
            nm++;
        }
    }
    return nm;
}

// Sequential replay of SearchByProjection(CurrentFrame, LastFrame) (src/ORBmatcher.cc:1385-1468).
int replay_projection_frame(const orbx_kp* kps, const float* u_right, const uint8_t* occupied_in, int n, const std::vector<float>& qr,
                            const float* proj_ur, const float* last_angle, const uint8_t* mp_has_obs, int nmp, int checkOri,
                            const std::vector<int>& off, const std::vector<int>& idx, const std::vector<int>& dist, int* feature_match)
{
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nm = 0;
    for (int i = 0; i < nmp; i++) {
        if (off[i + 1] == off[i]) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int c = off[i]; c < off[i + 1]; c++) {
            const int i2 = idx[c];
            if (occupied[i2]) continue;
            if (u_right[i2] > 0) {
                const float er = fabsf(proj_ur[i] - u_right[i2]);
                if (er > qr[i]) continue;
            }
            if (dist[c] < bestDist) { bestDist = dist[c]; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            feature_match[bestIdx2] = i;
            occupied[bestIdx2] = mp_has_obs[i];
            nm++;
            if (checkOri) {
                float rot = last_angle[i] - kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int j : rotHist[i]) { feature_match[j] = -1; nm--; }
    }
    return nm;
}

}  // namespace orbx

extern "C" int orbm_search_for_initialization(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1,
                                              const orbx_kp* k2, const uint8_t* d2, int n2,
                                              float minX, float maxX, float minY, float maxY,
                                              float* prev_matched, int* matches12, int windowSize, float nnratio, int checkOri,
                                              int* nmatches)
{
    if (!m || n1 < 0 || n2 < 0 || (n1 > 0 && (!k1 || !d1 || !prev_matched || !matches12)) || (n2 > 0 && (!k2 || !d2)) ||
        !(maxX > minX) || !(maxY > minY)) {
        set_error("orbm_search_for_initialization: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    // queries: level-0 keypoints of F1 only (:421-423), searched around vbPrevMatched in level 0 of F2 (:425)
    std::vector<int> qmap;
    std::vector<float> qxy, qr;
    std::vector<uint8_t> qdesc;
    for (int i = 0; i < n1; i++)
        if (k1[i].octave <= 0) {
            qmap.push_back(i);
            qxy.push_back(prev_matched[2 * i]);
            qxy.push_back(prev_matched[2 * i + 1]);
            qr.push_back((float)windowSize);
        }
    const int nq = (int)qmap.size();
    if (nq == 0 || n2 == 0) return ORBX_OK;
    qdesc.resize((size_t)nq * 32);
    for (int q = 0; q < nq; q++) memcpy(&qdesc[(size_t)q * 32], d1 + (size_t)qmap[q] * 32, 32);

    ORBX_CUDA(cudaSetDevice(m->device));
    GridParams gp;
    std::vector<int> start, entries;
    build_grid_host(k2, n2, minX, maxX, minY, maxY, gp, start, entries);
    Carver cv(m->arena);
    const size_t oK = cv.take(sizeof(orbx_kp) * (size_t)n2), oS = cv.take(sizeof(int) * start.size()), oE = cv.take(sizeof(int) * entries.size()),
                 oQ = cv.take(sizeof(float) * 2 * (size_t)nq), oR = cv.take(sizeof(float) * (size_t)nq), oO = cv.take(sizeof(int) * ((size_t)nq + 1)),
                 oQD = cv.take((size_t)nq * 32), oTD = cv.take((size_t)n2 * 32);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    {
        Xfer x(m->arena);             // the seven input arrays travel as one copy through the page-locked mirror
        x.up(oK, k2, sizeof(orbx_kp) * (size_t)n2);
        x.up(oS, start.data(), sizeof(int) * start.size());
        x.up(oE, entries.data(), sizeof(int) * entries.size());
        x.up(oQ, qxy.data(), sizeof(float) * 2 * (size_t)nq);
        x.up(oR, qr.data(), sizeof(float) * (size_t)nq);
        x.up(oQD, qdesc.data(), (size_t)nq * 32);
        x.up(oTD, d2, (size_t)n2 * 32);
        ORBX_CUDA(x.send());
    }
    int* d_idx = nullptr;
    int total = 0;
    // GetFeaturesInArea(x, y, windowSize, level1, level1) with level1 == 0
    rc = grid_candidates_dev(m, (const orbx_kp*)(B + oK), (const int*)(B + oS), (const int*)(B + oE), gp, (const float*)(B + oQ),
                             (const float*)(B + oR), nq, 0, 0, nullptr, (int*)(B + oO), 2, &d_idx, &total);
    if (rc) return rc;
    int* d_dist = (int*)(m->arena2.base + idx_bytes(total));      // second array of arena2
    k_cand_dist<<<(nq + 7) / 8, 256, 0, s>>>((const uint4*)(B + oQD), nq, (const uint4*)(B + oTD), (const int*)(B + oO), d_idx, d_dist);
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 4;
    std::vector<int> off(nq + 1), idx(std::max(total, 1)), dist(std::max(total, 1));
    ORBX_CUDA(cudaMemcpyAsync(off.data(), B + oO, sizeof(int) * ((size_t)nq + 1), cudaMemcpyDeviceToHost, s));
    if (total > 0) {
        ORBX_CUDA(cudaMemcpyAsync(idx.data(), d_idx, sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost, s));
        ORBX_CUDA(cudaMemcpyAsync(dist.data(), d_dist, sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost, s));
    }
    ORBX_CUDA(cudaStreamSynchronize(s));

    const int nm = replay_initialization(k1, n1, k2, n2, qmap, off, idx, dist, nnratio, checkOri, prev_matched, matches12);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

// ================================================================================================ windowed searches
namespace orbx {

// Candidates (GetFeaturesInArea with a per-query level range) and their Hamming distances for nq queries against one
// frame, on the device; the CSR comes back to the host for the sequential replay of the reference's greedy loops.
int windowed_distances(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, int n, float minX, float maxX, float minY, float maxY,
                              const float* qxy, const float* qr, const int* qlev2, const uint8_t* qdesc, int nq,
                              std::vector<int>& off, std::vector<int>& idx, std::vector<int>& dist)
{
    off.assign((size_t)nq + 1, 0);
    idx.clear();
    dist.clear();
    if (nq == 0 || n == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    GridParams gp;
    std::vector<int> start, entries;
    build_grid_host(kps, n, minX, maxX, minY, maxY, gp, start, entries);
    Carver cv(m->arena);
    const size_t oK = cv.take(sizeof(orbx_kp) * (size_t)n), oS = cv.take(sizeof(int) * start.size()), oE = cv.take(sizeof(int) * entries.size()),
                 oQ = cv.take(sizeof(float) * 2 * (size_t)nq), oR = cv.take(sizeof(float) * (size_t)nq), oL = cv.take(sizeof(int) * 2 * (size_t)nq),
                 oO = cv.take(sizeof(int) * ((size_t)nq + 1)), oQD = cv.take((size_t)nq * 32), oTD = cv.take((size_t)n * 32);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    Xfer x(m->arena);
    x.up(oK, kps, sizeof(orbx_kp) * (size_t)n);
    x.up(oS, start.data(), sizeof(int) * start.size());
    x.up(oE, entries.data(), sizeof(int) * entries.size());
    x.up(oQ, qxy, sizeof(float) * 2 * (size_t)nq);
    x.up(oR, qr, sizeof(float) * (size_t)nq);
    x.up(oL, qlev2, sizeof(int) * 2 * (size_t)nq);
    x.up(oQD, qdesc, (size_t)nq * 32);
    x.up(oTD, desc, (size_t)n * 32);
    ORBX_CUDA(x.send());
    int* d_idx = nullptr;
    int total = 0;
    rc = grid_candidates_dev(m, (const orbx_kp*)(B + oK), (const int*)(B + oS), (const int*)(B + oE), gp, (const float*)(B + oQ),
                             (const float*)(B + oR), nq, 0, 0, (const int*)(B + oL), (int*)(B + oO), 2, &d_idx, &total, 2);
    if (rc) return rc;
    int* d_dist = (int*)(m->arena2.base + idx_bytes(total));      // second array of arena2
    k_cand_dist<<<(nq + 7) / 8, 256, 0, s>>>((const uint4*)(B + oQD), nq, (const uint4*)(B + oTD), (const int*)(B + oO), d_idx, d_dist);
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 4;
    idx.resize(std::max(total, 1));
    dist.resize(std::max(total, 1));
    // offsets through the arena's pinned mirror, the two candidate arrays (contiguous in arena2) through arena2's: two copies
    rc = m->arena2.mirror();
    if (rc) return rc;
    if (total > 0)
        ORBX_CUDA(cudaMemcpyAsync(m->arena2.pin, m->arena2.base, idx_bytes(total) + sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost, s));
    x.down(off.data(), oO, sizeof(int) * ((size_t)nq + 1));
    ORBX_CUDA(x.fetch());
    if (total > 0) {
        memcpy(idx.data(), m->arena2.pin, sizeof(int) * (size_t)total);
        memcpy(dist.data(), m->arena2.pin + idx_bytes(total), sizeof(int) * (size_t)total);
    }
    return ORBX_OK;
}

// distances of host CSR candidate lists (queries q against rows of t) on the device
int csr_distances(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, const std::vector<int>& off,
                         const std::vector<int>& idx, std::vector<int>& dist)
{
    const int total = off[nq];
    dist.assign(std::max(total, 1), 0);
    if (total == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oQ = cv.take((size_t)nq * 32), oT = cv.take((size_t)nt * 32), oO = cv.take(sizeof(int) * ((size_t)nq + 1)),
                 oI = cv.take(sizeof(int) * (size_t)total), oD = cv.take(sizeof(int) * (size_t)total);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    Xfer x(m->arena);
    x.up(oQ, q, (size_t)nq * 32);
    x.up(oT, t, (size_t)nt * 32);
    x.up(oO, off.data(), sizeof(int) * ((size_t)nq + 1));
    x.up(oI, idx.data(), sizeof(int) * (size_t)total);
    ORBX_CUDA(x.send());
    k_cand_dist<<<(nq + 7) / 8, 256, 0, s>>>((const uint4*)(B + oQ), nq, (const uint4*)(B + oT), (const int*)(B + oO), (const int*)(B + oI), (int*)(B + oD));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    x.down(dist.data(), oD, sizeof(int) * (size_t)total);
    ORBX_CUDA(x.fetch());
    return ORBX_OK;
}

}  // namespace orbx

extern "C" int orbm_search_by_projection(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied_in,
                                         int n, float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                         const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos,
                                         const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, float nnratio,
                                         int* feature_match, int* nmatches)
{
    if (!m || n < 0 || nmp < 0 || (n > 0 && (!kps || !desc || !u_right || !occupied_in || !feature_match)) || !sf || nlevels < 1 ||
        (nmp > 0 && (!proj_xy || !proj_xr || !level || !view_cos || !mp_desc || !mp_has_obs)) || !(maxX > minX) || !(maxY > minY)) {
        set_error("orbm_search_by_projection: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    const bool bFactor = th != 1.0;
    std::vector<float> qr(nmp), rr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        if (level[i] < 0 || level[i] >= nlevels) { set_error("orbm_search_by_projection: level[%d] = %d out of range", i, level[i]); return ORBX_ERR_ARG; }
        float r = view_cos[i] > 0.998 ? 2.5f : 4.0f;          // RadiusByViewingCos (:131-137)
        if (bFactor) r *= th;
        rr[i] = r;
        qr[i] = r * sf[level[i]];                              // :68
        qlev[2 * i] = level[i] - 1;
        qlev[2 * i + 1] = level[i];
    }
    std::vector<int> off, idx, dist;
    int rc = windowed_distances(m, kps, desc, n, minX, maxX, minY, maxY, proj_xy, qr.data(), qlev.data(), mp_desc, nmp, off, idx, dist);
    if (rc) return rc;
    const int nm = replay_projection(kps, u_right, occupied_in, n, sf, level, rr, proj_xr, mp_has_obs, nmp, nnratio, off, idx, dist, feature_match);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

extern "C" int orbm_search_by_projection_frame(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right,
                                               const uint8_t* occupied_in, int n, float minX, float maxX, float minY, float maxY,
                                               const float* sf, int nlevels, const float* proj_xy, const float* proj_ur,
                                               const int* last_octave, const float* last_angle, const uint8_t* mp_desc,
                                               const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                               int* feature_match, int* nmatches)
{
    if (!m || n < 0 || nmp < 0 || (n > 0 && (!kps || !desc || !u_right || !occupied_in || !feature_match)) || !sf || nlevels < 1 ||
        (nmp > 0 && (!proj_xy || !proj_ur || !last_octave || !last_angle || !mp_desc || !mp_has_obs)) || !(maxX > minX) || !(maxY > minY) ||
        mode < 0 || mode > 2) {
        set_error("orbm_search_by_projection_frame: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    std::vector<float> qr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        const int o = last_octave[i];
        if (o < 0 || o >= nlevels) { set_error("orbm_search_by_projection_frame: last_octave[%d] = %d out of range", i, o); return ORBX_ERR_ARG; }
        qr[i] = th * sf[o];                                                        // :1383
        if (mode == 1) { qlev[2 * i] = o; qlev[2 * i + 1] = -1; }                   // bForward: GetFeaturesInArea(u, v, radius, nLastOctave)
        else if (mode == 2) { qlev[2 * i] = 0; qlev[2 * i + 1] = o; }               // bBackward
        else { qlev[2 * i] = o - 1; qlev[2 * i + 1] = o + 1; }
    }
    std::vector<int> off, idx, dist;
    int rc = windowed_distances(m, kps, desc, n, minX, maxX, minY, maxY, proj_xy, qr.data(), qlev.data(), mp_desc, nmp, off, idx, dist);
    if (rc) return rc;
    const int nm = replay_projection_frame(kps, u_right, occupied_in, n, qr, proj_ur, last_angle, mp_has_obs, nmp, checkOri, off, idx, dist, feature_match);
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

namespace orbx {

// SearchByBoW inner scan (ORBmatcher.cc:205-226, :571-596) for one feature of the first key frame over the bucket of its
// vocabulary node in the second (a range [q.y, q.z) of idx2, visited in stored order): the BOW_TOP smallest keys
// distance << 22 | position, ascending.  "Already matched" is the only state the reference's loop reads, so the host replay
// walks these few keys instead of every distance; when the taken features exhaust them it rescans that one bucket itself.
constexpr int BOW_TOP = 8;      // with 4 keys about one query in nine needs the host rescan (synthetic key frames), with 8 one in eighty

__global__ void __launch_bounds__(256) k_bow_top(const uint4* __restrict__ d1, const int4* __restrict__ query, int nq,
                                                 const uint4* __restrict__ d2, const uint8_t* __restrict__ valid2,
                                                 const int* __restrict__ idx2, uint4* __restrict__ top)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const int4 q = query[qi];
    const uint4 qa = __ldg(&d1[2 * (size_t)q.x]), qb = __ldg(&d1[2 * (size_t)q.x + 1]);
    uint32_t t[BOW_TOP];             // this lane's smallest keys, ascending
#pragma unroll
    for (int r = 0; r < BOW_TOP; r++) t[r] = KEY_NONE;
    for (int c = q.y + lane; c < q.z; c += 32) {
        const int f2 = idx2[c];
        if (valid2 && !valid2[f2]) continue;
        const int d = hamming256(qa, qb, __ldg(&d2[2 * (size_t)f2]), __ldg(&d2[2 * (size_t)f2 + 1]));
        uint32_t key = ((uint32_t)d << KEY_SHIFT) | (uint32_t)(c - q.y);
        if (key < t[BOW_TOP - 1]) {
#pragma unroll
            for (int r = 0; r < BOW_TOP; r++) {          // insertion: key sinks to its place, the largest falls out
                const uint32_t lo = min(t[r], key), hi = max(t[r], key);
                t[r] = lo;
                key = hi;
            }
        }
    }
    uint32_t out[BOW_TOP];
#pragma unroll
    for (int r = 0; r < BOW_TOP; r++) {
        const uint32_t g = __reduce_min_sync(0xFFFFFFFFu, t[0]);
        out[r] = g;
        if (t[0] == g) {
#pragma unroll
            for (int k = 0; k + 1 < BOW_TOP; k++) t[k] = t[k + 1];
            t[BOW_TOP - 1] = KEY_NONE;
        }
    }
    if (lane == 0) {
        top[2 * (size_t)qi] = make_uint4(out[0], out[1], out[2], out[3]);
        top[2 * (size_t)qi + 1] = make_uint4(out[4], out[5], out[6], out[7]);
    }
}

static inline int hamming256_host(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32);
    memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) + __builtin_popcountll(x[2] ^ y[2]) +
           __builtin_popcountll(x[3] ^ y[3]);
}

}  // namespace orbx

extern "C" int orbm_search_by_bow(orbx_matcher* m, int variant, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* valid1,
                                  const int* node1, const int* off1, const int* idx1, int nn1,
                                  const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* valid2,
                                  const int* node2, const int* off2, const int* idx2, int nn2,
                                  float nnratio, int checkOri, int* matches, int* nmatches)
{
    if (!m || (variant != 0 && variant != 1) || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || !matches ||
        (n1 > 0 && (!k1 || !d1 || !valid1)) || (n2 > 0 && (!k2 || !d2)) || (variant == 1 && n2 > 0 && !valid2) ||
        (nn1 > 0 && (!node1 || !off1 || !idx1)) || (nn2 > 0 && (!node2 || !off2 || !idx2))) {
        set_error("orbm_search_by_bow: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    const int nout = variant == 0 ? n2 : n1;
    for (int i = 0; i < nout; i++) matches[i] = -1;
    if (n2 >= (int)KEY_POS_MASK) { set_error("orbm_search_by_bow: too many features"); return ORBX_ERR_ARG; }
    const int nidx2 = nn2 > 0 ? off2[nn2] : 0;
    for (int p2 = 0; p2 < nidx2; p2++)
        if (idx2[p2] < 0 || idx2[p2] >= n2) { set_error("orbm_search_by_bow: feature index out of range"); return ORBX_ERR_ARG; }
    // merge-join of the two feature vectors on the node id (:180-264): every valid feature of a common node of the first
    // key frame becomes a query whose candidates are the features of the same node in the second, in stored order -- a range
    // of idx2, not a copy
    std::vector<int4> query;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (node1[a] == node2[b]) {
            for (int p1 = off1[a]; p1 < off1[a + 1]; p1++) {
                const int f1 = idx1[p1];
                if (f1 < 0 || f1 >= n1) { set_error("orbm_search_by_bow: feature index out of range"); return ORBX_ERR_ARG; }
                if (!valid1[f1]) continue;
                query.push_back(make_int4(f1, off2[b], off2[b + 1], 0));
            }
            a++; b++;
        } else if (node1[a] < node2[b]) {
            a = (int)(std::lower_bound(node1, node1 + nn1, node2[b]) - node1);
        } else {
            b = (int)(std::lower_bound(node2, node2 + nn2, node1[a]) - node2);
        }
    }
    const int nq = (int)query.size();
    if (nq == 0) return ORBX_OK;
    std::vector<uint32_t> top((size_t)nq * BOW_TOP);
    {
        ORBX_CUDA(cudaSetDevice(m->device));
        Carver cv(m->arena);
        const size_t oD1 = cv.take((size_t)n1 * 32), oD2 = cv.take((size_t)std::max(n2, 1) * 32), oV2 = cv.take((size_t)std::max(n2, 1)),
                     oQ = cv.take(sizeof(int4) * (size_t)nq), oI = cv.take(sizeof(int) * (size_t)std::max(nidx2, 1)),
                     oT = cv.take(sizeof(uint32_t) * (size_t)nq * BOW_TOP);
        int rc = m->arena.reserve(cv.total());
        if (!rc) rc = m->arena.mirror();
        if (rc) return rc;
        uint8_t* B = m->arena.base;
        Xfer x(m->arena);
        x.up(oD1, d1, (size_t)n1 * 32);
        x.up(oD2, d2, (size_t)n2 * 32);
        if (variant == 1) x.up(oV2, valid2, (size_t)n2);
        x.up(oQ, query.data(), sizeof(int4) * (size_t)nq);
        x.up(oI, idx2, sizeof(int) * (size_t)nidx2);
        ORBX_CUDA(x.send());
        k_bow_top<<<(nq + 7) / 8, 256, 0, m->stream>>>((const uint4*)(B + oD1), (const int4*)(B + oQ), nq, (const uint4*)(B + oD2),
                                                       variant == 1 ? (const uint8_t*)(B + oV2) : nullptr, (const int*)(B + oI), (uint4*)(B + oT));
        ORBX_CUDA(cudaGetLastError());
        m->lastLaunches = 1;
        x.down(top.data(), oT, sizeof(uint32_t) * (size_t)nq * BOW_TOP);
        ORBX_CUDA(x.fetch());
    }
    // sequential replay: features of the second frame taken by earlier queries are skipped (:209, :576)
    std::vector<uint8_t> taken2(n2, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nm = 0;
    for (int q = 0; q < nq; q++) {
        const int f1 = query[q].x, qs = query[q].y, qe = query[q].z;
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        const uint32_t* keys = &top[(size_t)q * BOW_TOP];
        int found = 0;
        bool complete = false;          // the keys seen so far cover every candidate that can still matter
        for (int r = 0; r < BOW_TOP && found < 2; r++) {
            const int d = (int)(keys[r] >> KEY_SHIFT);
            if (d >= 256) { complete = true; break; }                       // the rest never passes `dist < bestDist`
            const int f2 = idx2[qs + (int)(keys[r] & KEY_POS_MASK)];
            if (taken2[f2]) continue;
            if (found == 0) { bestDist1 = d; bestIdx2 = f2; }
            else bestDist2 = d;
            found++;
        }
        const bool lowSoFar = variant == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;
        if (found < 2 && !complete && qe - qs > BOW_TOP && (found == 0 || lowSoFar)) {
            // the taken features used up the returned keys: rescan this bucket here, exactly as the reference does
            bestDist1 = 256; bestIdx2 = -1; bestDist2 = 256;
            for (int c = qs; c < qe; c++) {
                const int f2 = idx2[c];
                if (taken2[f2]) continue;
                if (variant == 1 && !valid2[f2]) continue;
                const int d = hamming256_host(d1 + (size_t)f1 * 32, d2 + (size_t)f2 * 32);
                if (d < bestDist1) { bestDist2 = bestDist1; bestDist1 = d; bestIdx2 = f2; }
                else if (d < bestDist2) bestDist2 = d;
            }
        }
        const bool low = variant == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;      // :228 vs :598
        if (low && (float)bestDist1 < nnratio * (float)bestDist2) {
            taken2[bestIdx2] = 1;
            if (variant == 0) matches[bestIdx2] = f1; else matches[f1] = bestIdx2;
            if (checkOri) {
                float rot = k1[f1].angle - k2[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(variant == 0 ? bestIdx2 : f1);
            }
            nm++;
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { matches[j] = -1; nm--; }
        }
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

// ================================================================================================ ComputeStereoMatches
namespace orbx {

struct StereoView {
    const uint8_t* L[MAX_LEVELS];
    const uint8_t* R[MAX_LEVELS];
    int pitchL[MAX_LEVELS], pitchR[MAX_LEVELS], colsR[MAX_LEVELS];
    float sf[MAX_LEVELS], isf[MAX_LEVELS];
};

// Sub-pixel refinement of one left keypoint against its best right candidate (Frame.cc:566-637): 11-shift 11x11 SAD on the
// unblurred level of the left keypoint, parabola fit, disparity test.  Whole warp; lane 0 writes the three outputs
// (pre-initialised to -1 by the caller).
struct StereoLevel {      // what the refinement reads of the left keypoint's pyramid level
    const uint8_t* IL;
    const uint8_t* IR;
    int pL, pR, colsR;
    float sf, isf;
};

__device__ __forceinline__ void stereo_refine(int lane, const orbx_kp& kpL, float uL, float uR0, const StereoLevel& V,
                                              float bf, float minD, float maxD, float* uRightOut, float* depthOut, int* sadOut)
{
    const float scaleFactor = V.isf;
    const float scaleduL = roundf(__fmul_rn(kpL.x, scaleFactor));
    const float scaledvL = roundf(__fmul_rn(kpL.y, scaleFactor));
    const float scaleduR0 = roundf(__fmul_rn(uR0, scaleFactor));
    const int w = 5, Ls = 5;
    const float iniu = scaleduR0 + Ls - w, endu = scaleduR0 + Ls + w + 1;        // small integers: exact
    if (iniu < 0 || endu >= (float)V.colsR) return;
    const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
    const uint8_t* IL = V.IL;
    const uint8_t* IR = V.IR;
    const int pL = V.pL, pR = V.pR;
    const int cL = IL[(size_t)cv * pL + cu];
    int cR = 0;                                                 // centre of the right window at shift lane - 5
    if (lane < 11) cR = IR[(size_t)cv * pR + cr + lane - Ls];
    int sad[11], cRk[11];
#pragma unroll
    for (int k = 0; k < 11; k++) { sad[k] = 0; cRk[k] = __shfl_sync(0xFFFFFFFFu, cR, k); }
    for (int p = lane; p < 121; p += 32) {
        const int dy = p / 11 - w, dx = p % 11 - w;
        const int a = (int)IL[(size_t)(cv + dy) * pL + cu + dx] - cL;
        const uint8_t* rrow = IR + (size_t)(cv + dy) * pR + cr + dx - Ls;
        // |a - (r - cR)| = |(a + cR) - r|: one add and one sum-of-absolute-differences instruction per shift
#pragma unroll
        for (int k = 0; k < 11; k++) sad[k] = (int)__sad(a + cRk[k], (int)rrow[k], (unsigned)sad[k]);
    }
    // a lane's partial sums are below 2^12 and a whole SAD is at most 121 * 510 < 2^16: two shifts share a register in the
    // butterfly (6 x 5 shuffles instead of 11 x 5)
    {
        unsigned pk[6];
#pragma unroll
        for (int k = 0; k < 5; k++) pk[k] = (unsigned)sad[2 * k] | ((unsigned)sad[2 * k + 1] << 16);
        pk[5] = (unsigned)sad[10];
#pragma unroll
        for (int k = 0; k < 6; k++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) pk[k] += __shfl_xor_sync(0xFFFFFFFFu, pk[k], o);
#pragma unroll
        for (int k = 0; k < 5; k++) { sad[2 * k] = (int)(pk[k] & 0xFFFFu); sad[2 * k + 1] = (int)(pk[k] >> 16); }
        sad[10] = (int)pk[5];
    }
    if (lane != 0) return;
    int bestDist = INT_MAX, bestincR = 0;
#pragma unroll
    for (int k = 0; k < 11; k++)
        if (sad[k] < bestDist) { bestDist = sad[k]; bestincR = k - Ls; }
    if (bestincR == -Ls || bestincR == Ls) return;
    float d1 = 0, d2 = 0, d3 = 0;
#pragma unroll
    for (int k = 1; k < 10; k++)
        if (k == bestincR + Ls) { d1 = (float)sad[k - 1]; d2 = (float)sad[k]; d3 = (float)sad[k + 1]; }
    const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
    if (deltaR < -1 || deltaR > 1) return;                      // a NaN (flat triple) passes, as in the reference
    float bestuR = __fmul_rn(V.sf, __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
            disparity = 0.01f;
            bestuR = (float)((double)uL - 0.01);                // `bestuR = uL-0.01;` is a double subtraction
        }
        *depthOut = __fdiv_rn(bf, disparity);
        *uRightOut = bestuR;
        *sadOut = bestDist;
    }
}

// One warp per left keypoint: row-band Hamming search (Frame.cc:521-563), then the 11-shift 11x11 SAD on the unblurred
// pyramid level of the left keypoint (:566-607) and the parabola fit (:609-637).  SADs are integers (centre-subtracted
// u8 patches), so lanes can sum them in any order; the float tail runs on individually rounded operations.
__global__ void __launch_bounds__(256) k_stereo(const orbx_kp* __restrict__ kl, const uint4* __restrict__ dl, int nl,
                                                const orbx_kp* __restrict__ kr, const uint4* __restrict__ dr,
                                                const int* __restrict__ rowStart, const int* __restrict__ rowEntries, int nRows,
                                                StereoView V, float bf, float minD, float maxD,
                                                float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ bestSad)
{
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (iL >= nl) return;
    if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; bestSad[iL] = -1; }
    const orbx_kp kpL = kl[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int row = (int)vL;                                   // vRowIndices[vL]
    if (row < 0 || row >= nRows) return;
    const int s = rowStart[row], e = rowStart[row + 1];
    if (s == e) return;
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const uint4 qa = __ldg(&dl[2 * iL]), qb = __ldg(&dl[2 * iL + 1]);
    uint32_t best = KEY_NONE;
    for (int c = s + lane; c < e; c += 32) {
        const int iR = rowEntries[c];
        const orbx_kp kpR = kr[iR];
        if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
        if (kpR.x >= minU && kpR.x <= maxU) {
            const int d = hamming256(qa, qb, __ldg(&dr[2 * (size_t)iR]), __ldg(&dr[2 * (size_t)iR + 1]));
            if (d < TH_HIGH) best = min(best, ((uint32_t)d << KEY_SHIFT) | (uint32_t)(c - s));     // strict <, first in list order
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
    if (best == KEY_NONE || (int)(best >> KEY_SHIFT) >= (TH_HIGH + TH_LOW) / 2) return;
    const int bestIdxR = rowEntries[s + (int)(best & KEY_POS_MASK)];

    const StereoLevel SL = {V.L[levelL], V.R[levelL], V.pitchL[levelL], V.pitchR[levelL], V.colsR[levelL], V.sf[levelL], V.isf[levelL]};
    stereo_refine(lane, kpL, uL, kr[bestIdxR].x, SL, bf, minD, maxD, uRight + iL, depth + iL, bestSad + iL);
}

}  // namespace orbx

extern "C" int orbm_stereo_matches(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, int frame,
                                   const orbx_kp* kl, const uint8_t* dl, int nl, const orbx_kp* kr, const uint8_t* dr, int nr,
                                   float bf, float minD, float maxD, float* uRight, float* depth, int* nmatched)
{
    if (!m || !left || !right || nl < 0 || nr < 0 || (nl > 0 && (!kl || !dl || !uRight || !depth)) || (nr > 0 && (!kr || !dr))) {
        set_error("orbm_stereo_matches: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatched) *nmatched = 0;
    for (int i = 0; i < nl; i++) { uRight[i] = -1.0f; depth[i] = -1.0f; }
    if (nl == 0 || nr == 0) return ORBX_OK;
    const int nlev = orbx_levels(left);
    if (nlev != orbx_levels(right) || nlev > MAX_LEVELS) { set_error("orbm_stereo_matches: extractors differ"); return ORBX_ERR_ARG; }
    int rc = orbx_sync(left);
    if (rc) return rc;
    rc = orbx_sync(right);
    if (rc) return rc;
    StereoView V;
    memset(&V, 0, sizeof V);
    int nRows = 0;
    int lvW[MAX_LEVELS] = {0}, lvH[MAX_LEVELS] = {0};
    rc = orbx_scale_tables(left, V.sf, V.isf, nullptr, nullptr);
    if (rc) return rc;
    for (int l = 0; l < nlev; l++) {
        size_t pl = 0, pr = 0;
        int wl, hl, wr, hr;
        if ((rc = orbx_pyramid_device(left, frame, l, &V.L[l], &pl)) || (rc = orbx_pyramid_device(right, frame, l, &V.R[l], &pr)) ||
            (rc = orbx_level_size(left, l, &wl, &hl)) || (rc = orbx_level_size(right, l, &wr, &hr)))
            return rc;
        if (wl != wr || hl != hr) { set_error("orbm_stereo_matches: left/right pyramids differ in size"); return ORBX_ERR_ARG; }
        V.pitchL[l] = (int)pl;
        V.pitchR[l] = (int)pr;
        V.colsR[l] = wr;
        lvW[l] = wl; lvH[l] = hl;
        if (l == 0) nRows = hl;        // mpORBextractorLeft->mvImagePyramid[0].rows (:488)
    }
    // The reference reads an 11 x 11 window around the left keypoint and 11 x 21 around the right one in mvImagePyramid[octave]
    // (:578-595), which is always inside its 19-px padded buffers.  The device planes carry no border: keypoints of the paired
    // extractors stay 19 px (16 for the +-1 level of the right image) inside a level, anything else is refused here.
    for (int i = 0; i < nl; i++) {
        if (kl[i].octave < 0 || kl[i].octave >= nlev) { set_error("orbm_stereo_matches: left octave out of range"); return ORBX_ERR_ARG; }
        const int l = kl[i].octave;
        const float cu = roundf(kl[i].x * V.isf[l]), cv = roundf(kl[i].y * V.isf[l]);
        if (!(cu >= 5 && cu < lvW[l] - 5 && cv >= 5 && cv < lvH[l] - 5)) {
            set_error("orbm_stereo_matches: left keypoint %d lies within 5 px of the border of its level (not from this extractor?)", i);
            return ORBX_ERR_ARG;
        }
    }
    for (int i = 0; i < nr; i++) {
        if (kr[i].octave < 0 || kr[i].octave >= nlev) { set_error("orbm_stereo_matches: right octave out of range"); return ORBX_ERR_ARG; }
        for (int l = std::max(kr[i].octave - 1, 0); l <= std::min(kr[i].octave + 1, nlev - 1); l++) {
            const float cr = roundf(kr[i].x * V.isf[l]);
            if (!(cr >= 10 && cr < lvW[l] - 10)) {
                set_error("orbm_stereo_matches: right keypoint %d lies within 10 px of the border of level %d (not from this extractor?)", i, l);
                return ORBX_ERR_ARG;
            }
        }
    }
    // row table (:490-508): right keypoint iR is registered in rows floor(y - r) .. ceil(y + r), r = 2 * scale[octave]
    std::vector<int> rowStart(nRows + 1, 0), rowEntries;
    std::vector<std::pair<int, int>> span(nr);
    for (int iR = 0; iR < nr; iR++) {
        if (kr[iR].octave < 0 || kr[iR].octave >= nlev) { set_error("orbm_stereo_matches: right octave out of range"); return ORBX_ERR_ARG; }
        const float kpY = kr[iR].y, r = 2.0f * V.sf[kr[iR].octave];
        const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
        span[iR] = std::make_pair(std::max(minr, 0), std::min(maxr, nRows - 1));   // the reference indexes unchecked
        for (int y = span[iR].first; y <= span[iR].second; y++) rowStart[y + 1]++;
    }
    for (int y = 0; y < nRows; y++) rowStart[y + 1] += rowStart[y];
    rowEntries.assign(std::max(rowStart[nRows], 1), 0);
    {
        std::vector<int> fill(rowStart.begin(), rowStart.end() - 1);
        for (int iR = 0; iR < nr; iR++)
            for (int y = span[iR].first; y <= span[iR].second; y++) rowEntries[fill[y]++] = iR;
    }
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oKL = cv.take(sizeof(orbx_kp) * (size_t)nl), oDL = cv.take((size_t)nl * 32), oKR = cv.take(sizeof(orbx_kp) * (size_t)nr),
                 oDR = cv.take((size_t)nr * 32), oRS = cv.take(sizeof(int) * rowStart.size()), oRE = cv.take(sizeof(int) * rowEntries.size()),
                 oU = cv.take(sizeof(float) * (size_t)nl), oD = cv.take(sizeof(float) * (size_t)nl), oS = cv.take(sizeof(int) * (size_t)nl);
    rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    Xfer x(m->arena);                 // one copy per direction through the page-locked mirror
    x.up(oKL, kl, sizeof(orbx_kp) * (size_t)nl);
    x.up(oDL, dl, (size_t)nl * 32);
    x.up(oKR, kr, sizeof(orbx_kp) * (size_t)nr);
    x.up(oDR, dr, (size_t)nr * 32);
    x.up(oRS, rowStart.data(), sizeof(int) * rowStart.size());
    x.up(oRE, rowEntries.data(), sizeof(int) * rowEntries.size());
    ORBX_CUDA(x.send());
    k_stereo<<<(nl + 7) / 8, 256, 0, s>>>((const orbx_kp*)(B + oKL), (const uint4*)(B + oDL), nl, (const orbx_kp*)(B + oKR),
                                          (const uint4*)(B + oDR), (const int*)(B + oRS), (const int*)(B + oRE), nRows, V, bf, minD, maxD,
                                          (float*)(B + oU), (float*)(B + oD), (int*)(B + oS));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    std::vector<int> sad(nl);
    x.down(uRight, oU, sizeof(float) * (size_t)nl);
    x.down(depth, oD, sizeof(float) * (size_t)nl);
    x.down(sad.data(), oS, sizeof(int) * (size_t)nl);
    ORBX_CUDA(x.fetch());
    // median cut (:641-654): sort (SAD, iL) pairs, drop everything at or above 1.5 * 1.4 * median
    std::vector<std::pair<int, int>> distIdx;
    for (int i = 0; i < nl; i++)
        if (sad[i] >= 0) distIdx.push_back(std::make_pair(sad[i], i));
    if (distIdx.empty()) return ORBX_OK;      // the reference reads vDistIdx[0] of an empty vector here (:642): guarded
    std::sort(distIdx.begin(), distIdx.end());
    const float median = (float)distIdx[distIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)distIdx.size();
    for (int i = (int)distIdx.size() - 1; i >= 0; i--) {
        if (distIdx[i].first < thDist) break;
        uRight[distIdx[i].second] = -1;
        depth[distIdx[i].second] = -1;
        kept--;
    }
    if (nmatched) *nmatched = kept;
    return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------ batched stereo
namespace orbx {

struct StereoBatchView {
    const uint8_t* L[MAX_LEVELS];
    const uint8_t* R[MAX_LEVELS];
    int pitchL[MAX_LEVELS], pitchR[MAX_LEVELS], colsR[MAX_LEVELS];
    unsigned long long frameL[MAX_LEVELS], frameR[MAX_LEVELS];
    float sf[MAX_LEVELS], isf[MAX_LEVELS];
    int nRows, nlev;
};

// right keypoints as (row band, octave, x): rows floor(y - r) .. ceil(y + r), r = 2 * scale[octave] (Frame.cc:490-508)
__global__ void k_stereo_prep(const orbx_kp* __restrict__ kr, const int* __restrict__ cntR, int stride, StereoBatchView V, int4* __restrict__ tab)
{
    const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cntR[f]) return;
    const orbx_kp k = kr[(size_t)f * stride + i];
    const int o = min(max(k.octave, 0), V.nlev - 1);
    const float r = __fmul_rn(2.0f, V.sf[o]);
    const int maxr = min((int)ceilf(__fadd_rn(k.y, r)), V.nRows - 1), minr = max((int)floorf(__fsub_rn(k.y, r)), 0);
    tab[(size_t)f * stride + i] = make_int4(minr, maxr, k.octave, __float_as_int(k.x));
}

// vRowIndices (Frame.cc:490-508) at a granularity of STEREO_BAND rows: for every band of a pair, the right keypoints whose
// row range touches it, in index order.  One CTA per pair: count per band (a warp per band, ballots), scan, ordered fill.
// A left keypoint then scans one band list (a few dozen entries) instead of every right keypoint of the frame.
constexpr int STEREO_BAND = 8, STEREO_LIST_FACTOR = 4;      // a row range spans at most 18 rows = 4 bands of 8

__global__ void __launch_bounds__(256) k_stereo_bands(const int* __restrict__ cntR, int stride, const int4* __restrict__ tab, int nBands,
                                                      int* __restrict__ bandStart, int* __restrict__ bandList)
{
    extern __shared__ int s_cnt[];                               // nBands + 1, then the row ranges of the right keypoints
    const int f = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31, nr = min(cntR[f], stride);
    const int4* tf = tab + (size_t)f * stride;
    int* start = bandStart + (size_t)f * (nBands + 1);
    int* list = bandList + (size_t)f * stride * STEREO_LIST_FACTOR;
    // every band scans every right keypoint twice: keep (first row | last row << 16) in shared memory
    unsigned* s_rng = reinterpret_cast<unsigned*>(s_cnt + nBands + 1);
    for (int c = threadIdx.x; c < nr; c += 256) {
        const int4 t = __ldg(&tf[c]);
        s_rng[c] = (unsigned)t.x | ((unsigned)t.y << 16);
    }
    __syncthreads();
    for (int pass = 0; pass < 2; pass++) {
        for (int b = w; b < nBands; b += 8) {
            const int r0 = b * STEREO_BAND, r1 = r0 + STEREO_BAND - 1;
            int n = pass ? s_cnt[b] : 0;
            for (int c0 = 0; c0 < nr; c0 += 32) {
                const int c = c0 + lane;
                bool hit = false;
                if (c < nr) {
                    const unsigned t = s_rng[c];
                    hit = (int)(t & 0xFFFFu) <= r1 && (int)(t >> 16) >= r0;
                }
                const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
                if (pass && hit) list[n + __popc(m & ((1u << lane) - 1))] = c;
                n += __popc(m);
            }
            if (!pass && lane == 0) s_cnt[b] = n;
        }
        __syncthreads();
        if (!pass) {
            if (threadIdx.x == 0) {
                int acc = 0;
                for (int b = 0; b < nBands; b++) { const int t = s_cnt[b]; s_cnt[b] = acc; acc += t; }
                s_cnt[nBands] = acc;
            }
            __syncthreads();
            for (int b = threadIdx.x; b <= nBands; b += 256) start[b] = s_cnt[b];
        }
    }
}

// One warp per (pair, left keypoint).  Candidates = right keypoints whose row band contains the left row, visited in
// index order like vRowIndices (Frame.cc:521-563); the refinement is stereo_refine.
#ifndef STB_MINB
#define STB_MINB 8      // latency-bound: 64 warps per SM at 32 registers, 144 bytes of spills (measured 3.0 -> 2.6 us per 752x480 pair against 4 CTAs per SM)
#endif
__global__ void __launch_bounds__(256, STB_MINB) k_stereo_batch(const orbx_kp* __restrict__ kl, const uint4* __restrict__ dl, const int* __restrict__ cntL,
                                                      int strideL, const orbx_kp* __restrict__ kr, const uint4* __restrict__ dr,
                                                      const int* __restrict__ cntR, int strideR, const int4* __restrict__ tab,
                                                      const int* __restrict__ bandStart, const int* __restrict__ bandList, int nBands,
                                                      StereoBatchView VB, float bf, float minD, float maxD, int cap,
                                                      float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ bestSad)
{
    const int f = blockIdx.y, iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int nl = cntL[f];
    if (iL >= nl || iL >= cap) return;
    const size_t oo = (size_t)f * cap + iL;
    if (lane == 0) { uRight[oo] = -1.0f; depth[oo] = -1.0f; bestSad[oo] = -1; }
    const orbx_kp kpL = kl[(size_t)f * strideL + iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int row = (int)vL;
    if (row < 0 || row >= VB.nRows) return;
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const uint4* dlf = dl + 2 * ((size_t)f * strideL + iL);
    const uint4 qa = __ldg(dlf), qb = __ldg(dlf + 1);
    const int4* tf = tab + (size_t)f * strideR;
    const uint4* drf = dr + 2 * (size_t)f * strideR;
    uint32_t best = KEY_NONE;
    const int* bs = bandStart + (size_t)f * (nBands + 1) + row / STEREO_BAND;
    const int* bl = bandList + (size_t)f * strideR * STEREO_LIST_FACTOR;
    for (int p = bs[0] + lane, pe = bs[1]; p < pe; p += 32) {
        const int c = __ldg(&bl[p]);
        const int4 t = __ldg(&tf[c]);
        if (row < t.x || row > t.y) continue;
        if (t.z < levelL - 1 || t.z > levelL + 1) continue;
        const float x = __int_as_float(t.w);
        if (x >= minU && x <= maxU) {
            const int d = hamming256(qa, qb, __ldg(&drf[2 * (size_t)c]), __ldg(&drf[2 * (size_t)c + 1]));
            if (d < TH_HIGH) best = min(best, ((uint32_t)d << KEY_SHIFT) | (uint32_t)c);       // strict <, first in index order
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
    if (best == KEY_NONE || (int)(best >> KEY_SHIFT) >= (TH_HIGH + TH_LOW) / 2) return;
    const int bestIdxR = (int)(best & KEY_POS_MASK);
    if (levelL < 0 || levelL >= VB.nlev) return;
    const StereoLevel SL = {VB.L[levelL] + (size_t)f * VB.frameL[levelL], VB.R[levelL] + (size_t)f * VB.frameR[levelL], VB.pitchL[levelL],
                            VB.pitchR[levelL], VB.colsR[levelL], VB.sf[levelL], VB.isf[levelL]};
    stereo_refine(lane, kpL, uL, __int_as_float(__ldg(&tf[bestIdxR]).w), SL, bf, minD, maxD, uRight + oo, depth + oo, bestSad + oo);
}

// Median cut (Frame.cc:641-654) of one pair per CTA: the reference sorts (SAD, index) pairs and takes the SAD at position
// size / 2; that value is a rank selection, done here with two 8-bit histogram passes over the 16-bit SADs
// (SAD <= 121 * 510).  Matches with (float)SAD >= 1.5f * 1.4f * median are removed.
__global__ void __launch_bounds__(256) k_stereo_median(const int* __restrict__ cntL, int cap, const int* __restrict__ bestSad,
                                                       float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ nmatched)
{
    __shared__ int hist[256];
    __shared__ int s_bin, s_rank, s_total;
    const int f = blockIdx.x, tid = threadIdx.x, nl = min(cntL[f], cap);
    const int* sad = bestSad + (size_t)f * cap;
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nl; i += 256) if (sad[i] >= 0) atomicAdd(&hist[(sad[i] >> 8) & 255], 1);
    __syncthreads();
    if (tid == 0) {
        int total = 0;
        for (int b = 0; b < 256; b++) total += hist[b];
        s_total = total;
        int k = total / 2, b = 0;                       // vDistIdx[vDistIdx.size() / 2]
        while (b < 255 && k >= hist[b]) { k -= hist[b]; b++; }
        s_bin = b;
        s_rank = k;
    }
    __syncthreads();
    const int total = s_total, hb = s_bin, k = s_rank;
    if (total == 0) { if (tid == 0) nmatched[f] = 0; return; }     // the reference reads vDistIdx[0] of an empty vector here: guarded
    __syncthreads();
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nl; i += 256) if (sad[i] >= 0 && ((sad[i] >> 8) & 255) == hb) atomicAdd(&hist[sad[i] & 255], 1);
    __syncthreads();
    if (tid == 0) {
        int kk = k, b = 0;
        while (b < 255 && kk >= hist[b]) { kk -= hist[b]; b++; }
        s_bin = (hb << 8) | b;
        s_total = 0;
    }
    __syncthreads();
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)s_bin);
    int kept = 0;
    for (int i = tid; i < nl; i += 256) {
        if (sad[i] < 0) continue;
        if ((float)sad[i] < thDist) kept++;
        else { uRight[(size_t)f * cap + i] = -1.0f; depth[(size_t)f * cap + i] = -1.0f; }
    }
    atomicAdd(&s_total, kept);
    __syncthreads();
    if (tid == 0) nmatched[f] = s_total;
}

}  // namespace orbx

extern "C" int orbm_stereo_matches_batch(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, float bf, float minD, float maxD,
                                         float* uRight, float* depth, int cap, int* nmatched)
{
    if (!m || !left || !right || !uRight || !depth || cap < 1) { set_error("orbm_stereo_matches_batch: bad argument"); return ORBX_ERR_ARG; }
    const int B = orbx_last_batch(left);
    if (B < 1 || orbx_last_batch(right) != B) { set_error("orbm_stereo_matches_batch: the two handles must hold results of equal batch size"); return ORBX_ERR_STATE; }
    const int nlev = orbx_levels(left);
    if (nlev != orbx_levels(right) || nlev > MAX_LEVELS) { set_error("orbm_stereo_matches_batch: extractors differ"); return ORBX_ERR_ARG; }
    int rc = orbx_sync(left);
    if (rc) return rc;
    if ((rc = orbx_sync(right))) return rc;
    StereoBatchView V;
    memset(&V, 0, sizeof V);
    V.nlev = nlev;
    if ((rc = orbx_scale_tables(left, V.sf, V.isf, nullptr, nullptr))) return rc;
    for (int l = 0; l < nlev; l++) {
        size_t pl = 0, pr = 0, fl = 0, fr = 0;
        int wl, hl, wr, hr;
        if ((rc = orbx_pyramid_layout(left, l, &V.L[l], &pl, &fl)) || (rc = orbx_pyramid_layout(right, l, &V.R[l], &pr, &fr)) ||
            (rc = orbx_level_size(left, l, &wl, &hl)) || (rc = orbx_level_size(right, l, &wr, &hr)))
            return rc;
        if (wl != wr || hl != hr) { set_error("orbm_stereo_matches_batch: left/right pyramids differ in size"); return ORBX_ERR_ARG; }
        V.pitchL[l] = (int)pl; V.pitchR[l] = (int)pr; V.frameL[l] = fl; V.frameR[l] = fr; V.colsR[l] = wr;
        if (l == 0) V.nRows = hl;
    }
    const orbx_kp *dkl, *dkr;
    const uint8_t *ddl, *ddr;
    const int *dcl, *dcr;
    int sl = 0, sr = 0;
    if ((rc = orbx_results_device(left, &dkl, &ddl, &dcl, &sl)) || (rc = orbx_results_device(right, &dkr, &ddr, &dcr, &sr))) return rc;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const int nBands = (V.nRows + STEREO_BAND - 1) / STEREO_BAND;
    // results first (uRight | depth | counts are one device-to-host copy into the page-locked mirror), workspace after
    const size_t oU = cv.take(sizeof(float) * (size_t)B * cap), oD = cv.take(sizeof(float) * (size_t)B * cap), oN = cv.take(sizeof(int) * (size_t)B),
                 oEnd = cv.take(0), oT = cv.take(sizeof(int4) * (size_t)B * sr), oS = cv.take(sizeof(int) * (size_t)B * cap),
                 oBS = cv.take(sizeof(int) * (size_t)B * (nBands + 1)), oBL = cv.take(sizeof(int) * (size_t)B * sr * STEREO_LIST_FACTOR);
    if ((rc = m->arena.reserve(cv.total())) || (rc = m->arena.mirror())) return rc;
    uint8_t* A = m->arena.base;
    cudaStream_t s = m->stream;
    k_stereo_prep<<<dim3((sr + 255) / 256, B), 256, 0, s>>>(dkr, dcr, sr, V, (int4*)(A + oT));
    // left keypoints beyond cap would be dropped: cap must cover the per-frame keypoint stride of the handle
    const int nlMax = std::min(sl, cap);
    const size_t bandSmem = sizeof(int) * ((size_t)(nBands + 1) + (size_t)sr);
    if (bandSmem > 48 * 1024) {
        if (bandSmem > 227 * 1024) { set_error("orbm_stereo_matches_batch: too many keypoints per frame for the row-band kernel"); return ORBX_ERR_ARG; }
        ORBX_CUDA(cudaFuncSetAttribute(k_stereo_bands, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bandSmem));
    }
    k_stereo_bands<<<B, 256, bandSmem, s>>>(dcr, sr, (const int4*)(A + oT), nBands, (int*)(A + oBS), (int*)(A + oBL));
    k_stereo_batch<<<dim3((nlMax + 7) / 8, B), 256, 0, s>>>(dkl, (const uint4*)ddl, dcl, sl, dkr, (const uint4*)ddr, dcr, sr, (const int4*)(A + oT),
                                                          (const int*)(A + oBS), (const int*)(A + oBL), nBands, V,
                                                          bf, minD, maxD, cap, (float*)(A + oU), (float*)(A + oD), (int*)(A + oS));
    k_stereo_median<<<B, 256, 0, s>>>(dcl, cap, (const int*)(A + oS), (float*)(A + oU), (float*)(A + oD), (int*)(A + oN));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 4;
    ORBX_CUDA(cudaMemcpyAsync(m->arena.pin + oU, A + oU, oEnd - oU, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(uRight, m->arena.pin + oU, sizeof(float) * (size_t)B * cap);
    memcpy(depth, m->arena.pin + oD, sizeof(float) * (size_t)B * cap);
    if (nmatched) memcpy(nmatched, m->arena.pin + oN, sizeof(int) * (size_t)B);
    return ORBX_OK;
}
