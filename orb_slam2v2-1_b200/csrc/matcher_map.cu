// matcher_map.cu -- the searches of the LocalMapping / LoopClosing / relocalisation callers: ORBmatcher::Fuse x2
// (reference src/ORBmatcher.cc:827-975, :977-1102), SearchBySim3 (:1104-1328), SearchByProjection(KeyFrame*, Scw) (:290-403),
// SearchByProjection(Frame&, KeyFrame*) (:1474-1601) and SearchForTriangulation (:657-825); MapPoint::ComputeDistinctiveDescriptors
// (src/MapPoint.cc:252-312) batched over MapPoints.
//
// Fuse and SearchBySim3 do not update any state the later points read inside the search, so their whole inner loop
// (KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:572-611, level filter, chi-square gate, Hamming argmin) is ONE kernel,
// one warp per projected point.  The other three mark features as taken while they iterate; they get the candidate
// distances from the device and replay the greedy bookkeeping on the host, like the searches in matcher_frame.cu.
#include "matcher_internal.cuh"
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>

namespace orbx {

// One warp per query.  Visits the grid columns of the window in the reference's order (ix outer, iy inner, insertion order
// in a cell); a lane keeps the smallest (distance << 22 | visiting position) of the entries it saw, so the warp minimum is
// the FIRST minimum of the sequential scan (`if (dist < bestDist)`).
__global__ void __launch_bounds__(256) k_area_best(const orbx_kp* __restrict__ kps, const uint4* __restrict__ desc,
                                                   const float* __restrict__ uRight, const uint8_t* __restrict__ skip,
                                                   const int* __restrict__ cellStart, const int* __restrict__ entries, GridParams gp,
                                                   const float* __restrict__ qxy, const float* __restrict__ qur,
                                                   const float* __restrict__ qr, const int* __restrict__ qlevel,
                                                   const uint4* __restrict__ qdesc, int nq, const float* __restrict__ invSigma2,
                                                   int* __restrict__ bestIdx, int* __restrict__ bestDist)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const int level = qlevel[qi];
    uint32_t myKey = KEY_NONE;
    int myId = -1;
    if (level >= 0) {
        const float x = qxy[2 * qi], y = qxy[2 * qi + 1], r = qr[qi], ur = qur ? qur[qi] : 0.0f;
        const uint4 qa = __ldg(&qdesc[2 * qi]), qb = __ldg(&qdesc[2 * qi + 1]);
        // KeyFrame.cc:577-591, float, left to right
        const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
        const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
        const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
        if (cx0 < GRID_COLS && cx1 >= 0 && cy0 < GRID_ROWS && cy1 >= 0) {
            uint32_t posBase = 0;                    // entries visited before this grid column
            for (int ix = cx0; ix <= cx1; ix++) {
                const int s = cellStart[ix * GRID_ROWS + cy0], e = cellStart[ix * GRID_ROWS + cy1 + 1];
                for (int c = s + lane; c < e; c += 32) {
                    const int id = entries[c];
                    const orbx_kp kp = kps[id];
                    const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                    if (!(fabsf(dx) < r && fabsf(dy) < r)) continue;
                    if (skip && skip[id]) continue;
                    if (kp.octave < level - 1 || kp.octave > level) continue;
                    if (invSigma2) {                                             // Fuse, ORBmatcher.cc:898-922
                        const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
                        float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                        const float kr = uRight[id];
                        double gate = 5.99;
                        if (kr >= 0) {
                            const float er = __fsub_rn(ur, kr);
                            e2 = __fadd_rn(e2, __fmul_rn(er, er));
                            gate = 7.8;
                        }
                        if ((double)__fmul_rn(e2, invSigma2[kp.octave]) > gate) continue;
                    }
                    const int d = hamming256(qa, qb, __ldg(&desc[2 * (size_t)id]), __ldg(&desc[2 * (size_t)id + 1]));
                    const uint32_t key = ((uint32_t)d << KEY_SHIFT) | (posBase + (uint32_t)(c - s));
                    if (key < myKey) { myKey = key; myId = id; }
                }
                posBase += (uint32_t)(e - s);
            }
        }
    }
    const uint32_t mn = __reduce_min_sync(0xFFFFFFFFu, myKey);
    if (mn >= (256u << KEY_SHIFT)) {
        if (lane == 0) { bestIdx[qi] = -1; bestDist[qi] = 256; }
    } else if (myKey == mn) {
        bestIdx[qi] = myId;
        bestDist[qi] = (int)(mn >> KEY_SHIFT);
    }
}

// Candidate distances of SearchForTriangulation with the state-independent gates folded in: a candidate that fails the
// stereo flags, the Hamming threshold, the epipole distance (:735-741) or CheckDistEpipolarLine (:140-157) gets 1000.
struct TriParams { float F[9]; float ex, ey; int onlyStereo; };

struct TriQuery {       // per-warp constants of one untracked feature of the first key frame
    uint4 qa, qb;
    float la, lb, lc, den;
    bool stereo1;
};

__device__ __forceinline__ TriQuery tri_query(const orbx_kp* __restrict__ k1, const uint4* __restrict__ d1, const float* __restrict__ uR1,
                                              int f1, const TriParams& T)
{
    TriQuery Q;
    const orbx_kp kp1 = k1[f1];
    Q.stereo1 = uR1[f1] >= 0;
    Q.qa = __ldg(&d1[2 * (size_t)f1]);
    Q.qb = __ldg(&d1[2 * (size_t)f1 + 1]);
    // epipolar line in the second image, l = x1' F12 (:143-145), left to right, no contraction
    Q.la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, T.F[0]), __fmul_rn(kp1.y, T.F[3])), T.F[6]);
    Q.lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, T.F[1]), __fmul_rn(kp1.y, T.F[4])), T.F[7]);
    Q.lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, T.F[2]), __fmul_rn(kp1.y, T.F[5])), T.F[8]);
    Q.den = __fadd_rn(__fmul_rn(Q.la, Q.la), __fmul_rn(Q.lb, Q.lb));
    return Q;
}

// distance of candidate f2, or 1000 when a state-independent gate rejects it
__device__ __forceinline__ int tri_candidate(const TriQuery& Q, int f2, const orbx_kp* __restrict__ k2, const uint4* __restrict__ d2,
                                             const float* __restrict__ uR2, const uint8_t* __restrict__ hasMp2,
                                             const float* __restrict__ sf2, const float* __restrict__ sigma2, const TriParams& T)
{
    const bool stereo2 = uR2[f2] >= 0;
    if (hasMp2[f2] || (T.onlyStereo && !stereo2)) return 1000;
    const int d = hamming256(Q.qa, Q.qb, __ldg(&d2[2 * (size_t)f2]), __ldg(&d2[2 * (size_t)f2 + 1]));
    if (d > TH_LOW) return 1000;
    const orbx_kp kp2 = k2[f2];
    if (!Q.stereo1 && !stereo2) {
        const float distex = __fsub_rn(T.ex, kp2.x), distey = __fsub_rn(T.ey, kp2.y);
        if (__fadd_rn(__fmul_rn(distex, distex), __fmul_rn(distey, distey)) < __fmul_rn(100.0f, sf2[kp2.octave])) return 1000;
    }
    const float num = __fadd_rn(__fadd_rn(__fmul_rn(Q.la, kp2.x), __fmul_rn(Q.lb, kp2.y)), Q.lc);
    if (Q.den == 0) return 1000;
    const float dsqr = __fdiv_rn(__fmul_rn(num, num), Q.den);
    return (double)dsqr < __dmul_rn(3.84, (double)sigma2[kp2.octave]) ? d : 1000;
}

__global__ void __launch_bounds__(256) k_triangulation_dist(const orbx_kp* __restrict__ k1, const uint4* __restrict__ d1,
                                                            const float* __restrict__ uR1, const int* __restrict__ qfeat, int nq,
                                                            const orbx_kp* __restrict__ k2, const uint4* __restrict__ d2,
                                                            const float* __restrict__ uR2, const uint8_t* __restrict__ hasMp2,
                                                            const float* __restrict__ sf2, const float* __restrict__ sigma2,
                                                            TriParams T, const int* __restrict__ off, const int* __restrict__ idx,
                                                            int* __restrict__ dist)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const TriQuery Q = tri_query(k1, d1, uR1, qfeat[qi], T);
    for (int c = off[qi] + lane; c < off[qi + 1]; c += 32) dist[c] = tri_candidate(Q, idx[c], k2, d2, uR2, hasMp2, sf2, sigma2, T);
}

// The same scan, but only the candidates that pass every gate are written back: the host replay (vbMatched2, running best)
// never looks at the others.  A query = one feature of key frame 1 and the range [q.y, q.z) of its vocabulary node's bucket
// in idx2 (no per-query copy of the bucket); survivors go, in candidate order, into the query's slot of TRI_SLOT entries
// (f2 | distance << 24); cnt[q] counts them all, so the host sees an overflow and falls back to k_triangulation_dist.
constexpr int TRI_SLOT = 16;

__global__ void __launch_bounds__(256) k_triangulation_compact(const orbx_kp* __restrict__ k1, const uint4* __restrict__ d1,
                                                               const float* __restrict__ uR1, const int4* __restrict__ query, int nq,
                                                               const orbx_kp* __restrict__ k2, const uint4* __restrict__ d2,
                                                               const float* __restrict__ uR2, const uint8_t* __restrict__ hasMp2,
                                                               const float* __restrict__ sf2, const float* __restrict__ sigma2,
                                                               TriParams T, const int* __restrict__ idx2, int* __restrict__ cnt,
                                                               uint32_t* __restrict__ slots)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const int4 q = query[qi];
    const TriQuery Q = tri_query(k1, d1, uR1, q.x, T);
    int n = 0;
    for (int c0 = q.y; c0 < q.z; c0 += 32) {
        const int c = c0 + lane;
        int f2 = 0, d = 1000;
        if (c < q.z) {
            f2 = idx2[c];
            d = tri_candidate(Q, f2, k2, d2, uR2, hasMp2, sf2, sigma2, T);
        }
        const unsigned m = __ballot_sync(0xFFFFFFFFu, d != 1000);
        if (d != 1000) {
            const int pos = n + __popc(m & ((1u << lane) - 1));
            if (pos < TRI_SLOT) slots[(size_t)qi * TRI_SLOT + pos] = (uint32_t)f2 | ((uint32_t)d << 24);
        }
        n += __popc(m);
    }
    if (lane == 0) cnt[qi] = n;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:281-311), one warp per MapPoint.  Row i of the distance matrix is
// never stored: its values (0..256) go into a 257-bin histogram in shared memory and the median vDists[0.5*(N-1)] is the bin
// where the running count passes k = (N-1)/2.  The winner is the FIRST row with the least median (`median < BestMedian`).
constexpr int DD_WARPS = 8, DD_BINS = 288;          // 9 bins per lane, 257 used
__global__ void __launch_bounds__(DD_WARPS * 32) k_distinctive(const uint4* __restrict__ desc, const int* __restrict__ off, int npoints,
                                                               int* __restrict__ best)
{
    __shared__ int hist[DD_WARPS][DD_BINS];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int* H = hist[w];
    for (int p = blockIdx.x * DD_WARPS + w; p < npoints; p += gridDim.x * DD_WARPS) {
        const int base = off[p], n = off[p + 1] - base;
        if (n <= 0) {
            if (lane == 0) best[p] = -1;
            continue;
        }
        const int k = (n - 1) >> 1;
        uint32_t bestKey = 0xFFFFFFFFu;
        for (int i = 0; i < n; i++) {
#pragma unroll
            for (int b = 0; b < 9; b++) H[9 * lane + b] = 0;
            __syncwarp();
            const uint4 qa = __ldg(&desc[2 * (size_t)(base + i)]), qb = __ldg(&desc[2 * (size_t)(base + i) + 1]);
            for (int j = lane; j < n; j += 32)
                atomicAdd(&H[hamming256(qa, qb, __ldg(&desc[2 * (size_t)(base + j)]), __ldg(&desc[2 * (size_t)(base + j) + 1]))], 1);
            __syncwarp();
            int c[9], mine = 0;
#pragma unroll
            for (int b = 0; b < 9; b++) { c[b] = H[9 * lane + b]; mine += c[b]; }
            int incl = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                if (lane >= d) incl += t;
            }
            int run = incl - mine, median = -1;                 // entries below this lane's first bin
            if (run <= k && k < incl) {
#pragma unroll
                for (int b = 0; b < 9; b++) {
                    run += c[b];
                    if (median < 0 && run > k) median = 9 * lane + b;
                }
            }
            const unsigned who = __ballot_sync(0xFFFFFFFFu, median >= 0);
            median = __shfl_sync(0xFFFFFFFFu, median, __ffs(who) - 1);
            const uint32_t key = ((uint32_t)median << 16) | (uint32_t)i;
            bestKey = min(bestKey, key);
            __syncwarp();
        }
        if (lane == 0) best[p] = (int)(bestKey & 0xFFFFu);
    }
}

// uploads a frame (keypoints, grid, descriptors, optional uRight / skip) and nq queries, runs k_area_best, downloads
static int area_best(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip, int n,
                     float minX, float maxX, float minY, float maxY, const float* qxy, const float* qur, const float* qr,
                     const int* qlevel, const uint8_t* qdesc, int nq, const float* inv_sigma2, int nlevels, int* best_idx, int* best_dist)
{
    if (nq == 0) return ORBX_OK;
    if (n == 0) {
        for (int i = 0; i < nq; i++) { best_idx[i] = -1; best_dist[i] = 256; }
        return ORBX_OK;
    }
    ORBX_CUDA(cudaSetDevice(m->device));
    GridParams gp;
    std::vector<int> start, entries;
    build_grid_host(kps, n, minX, maxX, minY, maxY, gp, start, entries);
    Carver cv(m->arena);
    const size_t oK = cv.take(sizeof(orbx_kp) * (size_t)n), oS = cv.take(sizeof(int) * start.size()), oE = cv.take(sizeof(int) * entries.size()),
                 oTD = cv.take((size_t)n * 32), oU = cv.take(sizeof(float) * (size_t)n), oSk = cv.take((size_t)n),
                 oQ = cv.take(sizeof(float) * 2 * (size_t)nq), oQU = cv.take(sizeof(float) * (size_t)nq), oR = cv.take(sizeof(float) * (size_t)nq),
                 oL = cv.take(sizeof(int) * (size_t)nq), oQD = cv.take((size_t)nq * 32), oSg = cv.take(sizeof(float) * (size_t)std::max(nlevels, 1)),
                 oBI = cv.take(sizeof(int) * (size_t)nq), oBD = cv.take(sizeof(int) * (size_t)nq);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    Xfer x(m->arena);
    x.up(oK, kps, sizeof(orbx_kp) * (size_t)n);
    x.up(oS, start.data(), sizeof(int) * start.size());
    x.up(oE, entries.data(), sizeof(int) * entries.size());
    x.up(oTD, desc, (size_t)n * 32);
    if (u_right) x.up(oU, u_right, sizeof(float) * (size_t)n);
    if (skip) x.up(oSk, skip, (size_t)n);
    x.up(oQ, qxy, sizeof(float) * 2 * (size_t)nq);
    if (qur) x.up(oQU, qur, sizeof(float) * (size_t)nq);
    x.up(oR, qr, sizeof(float) * (size_t)nq);
    x.up(oL, qlevel, sizeof(int) * (size_t)nq);
    x.up(oQD, qdesc, (size_t)nq * 32);
    if (inv_sigma2) x.up(oSg, inv_sigma2, sizeof(float) * (size_t)nlevels);
    ORBX_CUDA(x.send());
    k_area_best<<<(nq + 7) / 8, 256, 0, s>>>((const orbx_kp*)(B + oK), (const uint4*)(B + oTD), u_right ? (const float*)(B + oU) : nullptr,
                                             skip ? (const uint8_t*)(B + oSk) : nullptr, (const int*)(B + oS), (const int*)(B + oE), gp,
                                             (const float*)(B + oQ), qur ? (const float*)(B + oQU) : nullptr, (const float*)(B + oR),
                                             (const int*)(B + oL), (const uint4*)(B + oQD), nq, inv_sigma2 ? (const float*)(B + oSg) : nullptr,
                                             (int*)(B + oBI), (int*)(B + oBD));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    x.down(best_idx, oBI, sizeof(int) * (size_t)nq);
    x.down(best_dist, oBD, sizeof(int) * (size_t)nq);
    ORBX_CUDA(x.fetch());
    return ORBX_OK;
}

static bool levels_ok(const char* who, const int* level, int nmp, int nlevels, bool allowSkip)
{
    for (int i = 0; i < nmp; i++)
        if (level[i] >= nlevels || (level[i] < 0 && !allowSkip)) {
            set_error("%s: level[%d] = %d out of range", who, i, level[i]);
            return false;
        }
    return true;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbm_search_in_area(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip, int n,
                        float minX, float maxX, float minY, float maxY, const float* qxy, const float* qur, const float* qr,
                        const int* qlevel, const uint8_t* qdesc, int nq, const float* inv_level_sigma2, int nlevels,
                        int* best_idx, int* best_dist)
{
    if (!m || n < 0 || nq < 0 || n >= (1 << 21) || (n > 0 && (!kps || !desc)) || (nq > 0 && (!qxy || !qr || !qlevel || !qdesc || !best_idx || !best_dist)) ||
        !(maxX > minX) || !(maxY > minY) || nlevels < 1 || nlevels > MAX_LEVELS || (inv_level_sigma2 && n > 0 && nq > 0 && (!u_right || !qur))) {
        set_error("orbm_search_in_area: bad argument");
        return ORBX_ERR_ARG;
    }
    if (!levels_ok("orbm_search_in_area", qlevel, nq, nlevels, true)) return ORBX_ERR_ARG;
    return area_best(m, kps, desc, u_right, skip, n, minX, maxX, minY, maxY, qxy, qur, qr, qlevel, qdesc, nq, inv_level_sigma2, nlevels,
                     best_idx, best_dist);
}

int orbm_fuse(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, int n,
              float minX, float maxX, float minY, float maxY, const float* sf, const float* inv_level_sigma2, int nlevels,
              const float* proj_xy, const float* proj_ur, const int* level, const uint8_t* mp_desc, int nmp, float th, int stereo_chi2,
              int* best_idx, int* nfused)
{
    if (!m || n < 0 || nmp < 0 || n >= (1 << 21) || (n > 0 && (!kps || !desc)) || !sf || nlevels < 1 || nlevels > MAX_LEVELS ||
        (nmp > 0 && (!proj_xy || !level || !mp_desc || !best_idx)) || !(maxX > minX) || !(maxY > minY) ||
        (stereo_chi2 && (!inv_level_sigma2 || (n > 0 && !u_right) || (nmp > 0 && !proj_ur)))) {
        set_error("orbm_fuse: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nfused) *nfused = 0;
    if (!levels_ok("orbm_fuse", level, nmp, nlevels, false)) return ORBX_ERR_ARG;
    std::vector<float> qr(nmp);
    std::vector<int> bd(std::max(nmp, 1));
    for (int i = 0; i < nmp; i++) qr[i] = th * sf[level[i]];                                // :878 / :1046
    int rc = area_best(m, kps, desc, u_right, nullptr, n, minX, maxX, minY, maxY, proj_xy, stereo_chi2 ? proj_ur : nullptr, qr.data(), level,
                       mp_desc, nmp, stereo_chi2 ? inv_level_sigma2 : nullptr, nlevels, best_idx, bd.data());
    if (rc) return rc;
    int nf = 0;
    for (int i = 0; i < nmp; i++) {
        if (bd[i] <= TH_LOW) nf++;                                                          // :938 / :1081
        else best_idx[i] = -1;
    }
    if (nfused) *nfused = nf;
    return ORBX_OK;
}

int orbm_search_by_sim3(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1, const orbx_kp* k2, const uint8_t* d2, int n2,
                        float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                        const float* proj12_xy, const int* level12, const uint8_t* mp1_desc,
                        const float* proj21_xy, const int* level21, const uint8_t* mp2_desc, float th, int* match12, int* nfound)
{
    if (!m || n1 < 0 || n2 < 0 || n1 >= (1 << 21) || n2 >= (1 << 21) || !sf || nlevels < 1 || nlevels > MAX_LEVELS ||
        (n1 > 0 && (!k1 || !d1 || !proj12_xy || !level12 || !mp1_desc || !match12)) ||
        (n2 > 0 && (!k2 || !d2 || !proj21_xy || !level21 || !mp2_desc)) || !(maxX > minX) || !(maxY > minY)) {
        set_error("orbm_search_by_sim3: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nfound) *nfound = 0;
    if (!levels_ok("orbm_search_by_sim3", level12, n1, nlevels, true) || !levels_ok("orbm_search_by_sim3", level21, n2, nlevels, true))
        return ORBX_ERR_ARG;
    std::vector<float> r12(n1), r21(n2);
    for (int i = 0; i < n1; i++) r12[i] = level12[i] >= 0 ? th * sf[level12[i]] : 0.0f;     // :1176
    for (int i = 0; i < n2; i++) r21[i] = level21[i] >= 0 ? th * sf[level21[i]] : 0.0f;     // :1256
    std::vector<int> m1(std::max(n1, 1)), b1(std::max(n1, 1)), m2(std::max(n2, 1)), b2(std::max(n2, 1));
    int rc = area_best(m, k2, d2, nullptr, nullptr, n2, minX, maxX, minY, maxY, proj12_xy, nullptr, r12.data(), level12, mp1_desc, n1,
                       nullptr, nlevels, m1.data(), b1.data());
    if (rc) return rc;
    rc = area_best(m, k1, d1, nullptr, nullptr, n1, minX, maxX, minY, maxY, proj21_xy, nullptr, r21.data(), level21, mp2_desc, n2,
                   nullptr, nlevels, m2.data(), b2.data());
    if (rc) return rc;
    m->lastLaunches = 2;
    int nf = 0;
    for (int i1 = 0; i1 < n1; i1++) {                                                        // :1305-1322
        match12[i1] = -1;
        const int idx2 = b1[i1] <= TH_HIGH ? m1[i1] : -1;
        if (idx2 < 0) continue;
        const int idx1 = b2[idx2] <= TH_HIGH ? m2[idx2] : -1;
        if (idx1 == i1) { match12[i1] = idx2; nf++; }
    }
    if (nfound) *nfound = nf;
    return ORBX_OK;
}

int orbm_search_by_projection_sim3(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const uint8_t* matched_in, int n,
                                   float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                   const float* proj_xy, const int* level, const uint8_t* mp_desc, int nmp, float th,
                                   int* feature_match, int* nmatches)
{
    if (!m || n < 0 || nmp < 0 || (n > 0 && (!kps || !desc || !matched_in || !feature_match)) || !sf || nlevels < 1 ||
        (nmp > 0 && (!proj_xy || !level || !mp_desc)) || !(maxX > minX) || !(maxY > minY)) {
        set_error("orbm_search_by_projection_sim3: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    if (!levels_ok("orbm_search_by_projection_sim3", level, nmp, nlevels, false)) return ORBX_ERR_ARG;
    std::vector<float> qr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        qr[i] = th * sf[level[i]];                                                           // :360
        qlev[2 * i] = level[i] - 1;                                                          // :380-381
        qlev[2 * i + 1] = level[i];
    }
    std::vector<int> off, idx, dist;
    int rc = windowed_distances(m, kps, desc, n, minX, maxX, minY, maxY, proj_xy, qr.data(), qlev.data(), mp_desc, nmp, off, idx, dist);
    if (rc) return rc;
    std::vector<uint8_t> matched(matched_in, matched_in + n);
    int nm = 0;
    for (int i = 0; i < nmp; i++) {
        int bestDist = 256, bestIdx = -1;
        for (int c = off[i]; c < off[i + 1]; c++) {
            const int j = idx[c];
            if (matched[j]) continue;                                                        // :375-376
            if (dist[c] < bestDist) { bestDist = dist[c]; bestIdx = j; }
        }
        if (bestDist <= TH_LOW) { feature_match[bestIdx] = i; matched[bestIdx] = 1; nm++; } // :393-397
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

int orbm_search_by_projection_kf(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const uint8_t* occupied_in, int n,
                                 float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                 const float* proj_xy, const int* level, const float* kf_angle, const uint8_t* mp_desc, int nmp,
                                 float th, int ORBdist, int checkOri, int* feature_match, int* nmatches)
{
    if (!m || n < 0 || nmp < 0 || (n > 0 && (!kps || !desc || !occupied_in || !feature_match)) || !sf || nlevels < 1 ||
        (nmp > 0 && (!proj_xy || !level || !kf_angle || !mp_desc)) || !(maxX > minX) || !(maxY > minY)) {
        set_error("orbm_search_by_projection_kf: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    if (!levels_ok("orbm_search_by_projection_kf", level, nmp, nlevels, false)) return ORBX_ERR_ARG;
    std::vector<float> qr(nmp);
    std::vector<int> qlev(2 * (size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        qr[i] = th * sf[level[i]];                                                           // :1528
        qlev[2 * i] = level[i] - 1;                                                          // :1530
        qlev[2 * i + 1] = level[i] + 1;
    }
    std::vector<int> off, idx, dist;
    int rc = windowed_distances(m, kps, desc, n, minX, maxX, minY, maxY, proj_xy, qr.data(), qlev.data(), mp_desc, nmp, off, idx, dist);
    if (rc) return rc;
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nm = 0;
    for (int i = 0; i < nmp; i++) {
        int bestDist = 256, bestIdx2 = -1;
        for (int c = off[i]; c < off[i + 1]; c++) {
            const int i2 = idx[c];
            if (occupied[i2]) continue;                                                      // :1543-1544
            if (dist[c] < bestDist) { bestDist = dist[c]; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist && bestIdx2 >= 0) {
            feature_match[bestIdx2] = i;
            occupied[bestIdx2] = 1;
            nm++;
            if (checkOri) {
                float rot = kf_angle[i] - kps[bestIdx2].angle;
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
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

int orbm_search_for_triangulation(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* has_mp1, const float* u_right1,
                                  const int* node1, const int* off1, const int* idx1, int nn1,
                                  const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* has_mp2, const float* u_right2,
                                  const int* node2, const int* off2, const int* idx2, int nn2,
                                  const float* F12, float ex, float ey, const float* sf2, const float* sigma2_2, int nlevels,
                                  int onlyStereo, int checkOri, int* matches12, int* nmatches)
{
    if (!m || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || !F12 || !sf2 || !sigma2_2 || nlevels < 1 || nlevels > MAX_LEVELS ||
        (n1 > 0 && (!k1 || !d1 || !has_mp1 || !u_right1 || !matches12)) || (n2 > 0 && (!k2 || !d2 || !has_mp2 || !u_right2)) ||
        (nn1 > 0 && (!node1 || !off1 || !idx1)) || (nn2 > 0 && (!node2 || !off2 || !idx2))) {
        set_error("orbm_search_for_triangulation: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    for (int j = 0; j < n2; j++)
        if (k2[j].octave < 0 || k2[j].octave >= nlevels) { set_error("orbm_search_for_triangulation: octave out of range"); return ORBX_ERR_ARG; }
    if (n2 >= (1 << 24)) { set_error("orbm_search_for_triangulation: more than 2^24 features"); return ORBX_ERR_ARG; }
    const int nidx2 = nn2 > 0 ? off2[nn2] : 0;
    for (int p2 = 0; p2 < nidx2; p2++)
        if (idx2[p2] < 0 || idx2[p2] >= n2) { set_error("orbm_search_for_triangulation: feature index out of range"); return ORBX_ERR_ARG; }
    // merge-join on the node id (:693-786): each untracked feature of a common node becomes a query over that node's bucket,
    // given as a range of idx2 (the bucket is not copied per query)
    std::vector<int4> query;
    long long total = 0;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (node1[a] == node2[b]) {
            for (int p1 = off1[a]; p1 < off1[a + 1]; p1++) {
                const int f1 = idx1[p1];
                if (f1 < 0 || f1 >= n1) { set_error("orbm_search_for_triangulation: feature index out of range"); return ORBX_ERR_ARG; }
                if (has_mp1[f1]) continue;                                                   // :699-703
                if (onlyStereo && !(u_right1[f1] >= 0)) continue;                            // :705-709
                query.push_back(make_int4(f1, off2[b], off2[b + 1], 0));
                total += off2[b + 1] - off2[b];
            }
            a++; b++;
        } else if (node1[a] < node2[b]) {
            a = (int)(std::lower_bound(node1, node1 + nn1, node2[b]) - node1);
        } else {
            b = (int)(std::lower_bound(node2, node2 + nn2, node1[a]) - node2);
        }
    }
    const int nq = (int)query.size();
    if (nq == 0 || total == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oK1 = cv.take(sizeof(orbx_kp) * (size_t)n1), oD1 = cv.take((size_t)n1 * 32), oU1 = cv.take(sizeof(float) * (size_t)n1),
                 oK2 = cv.take(sizeof(orbx_kp) * (size_t)n2), oD2 = cv.take((size_t)n2 * 32), oU2 = cv.take(sizeof(float) * (size_t)n2),
                 oH2 = cv.take((size_t)n2), oSF = cv.take(sizeof(float) * (size_t)nlevels), oSG = cv.take(sizeof(float) * (size_t)nlevels),
                 oQ = cv.take(sizeof(int4) * (size_t)nq), oI = cv.take(sizeof(int) * (size_t)nidx2),
                 oC = cv.take(sizeof(int) * (size_t)nq), oS = cv.take(sizeof(uint32_t) * (size_t)nq * TRI_SLOT);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    TriParams T;
    memcpy(T.F, F12, sizeof(T.F));
    T.ex = ex;
    T.ey = ey;
    T.onlyStereo = onlyStereo;
    std::vector<int> cnt(nq);
    std::vector<uint32_t> slots((size_t)nq * TRI_SLOT);
    {
        Xfer x(m->arena);
        x.up(oK1, k1, sizeof(orbx_kp) * (size_t)n1);
        x.up(oD1, d1, (size_t)n1 * 32);
        x.up(oU1, u_right1, sizeof(float) * (size_t)n1);
        x.up(oK2, k2, sizeof(orbx_kp) * (size_t)n2);
        x.up(oD2, d2, (size_t)n2 * 32);
        x.up(oU2, u_right2, sizeof(float) * (size_t)n2);
        x.up(oH2, has_mp2, (size_t)n2);
        x.up(oSF, sf2, sizeof(float) * (size_t)nlevels);
        x.up(oSG, sigma2_2, sizeof(float) * (size_t)nlevels);
        x.up(oQ, query.data(), sizeof(int4) * (size_t)nq);
        x.up(oI, idx2, sizeof(int) * (size_t)nidx2);
        ORBX_CUDA(x.send());
        k_triangulation_compact<<<(nq + 7) / 8, 256, 0, s>>>((const orbx_kp*)(B + oK1), (const uint4*)(B + oD1), (const float*)(B + oU1),
                                                             (const int4*)(B + oQ), nq, (const orbx_kp*)(B + oK2), (const uint4*)(B + oD2),
                                                             (const float*)(B + oU2), (const uint8_t*)(B + oH2), (const float*)(B + oSF),
                                                             (const float*)(B + oSG), T, (const int*)(B + oI), (int*)(B + oC), (uint32_t*)(B + oS));
        ORBX_CUDA(cudaGetLastError());
        m->lastLaunches = 1;
        x.down(cnt.data(), oC, sizeof(int) * (size_t)nq);
        x.down(slots.data(), oS, sizeof(uint32_t) * (size_t)nq * TRI_SLOT);
        ORBX_CUDA(x.fetch());
    }
    // candidates that passed the gates, per query, in candidate order: (f2 | distance << 24)
    std::vector<int> eoff(nq + 1, 0);
    std::vector<uint32_t> ent;
    bool overflow = false;
    for (int q = 0; q < nq; q++) overflow |= cnt[q] > TRI_SLOT;
    if (!overflow) {
        for (int q = 0; q < nq; q++) {
            ent.insert(ent.end(), slots.begin() + (size_t)q * TRI_SLOT, slots.begin() + (size_t)q * TRI_SLOT + cnt[q]);
            eoff[q + 1] = (int)ent.size();
        }
    } else {
        // a query with more than TRI_SLOT surviving candidates: every distance comes back (one more launch)
        std::vector<int> off(nq + 1, 0), qfeat(nq);
        for (int q = 0; q < nq; q++) { qfeat[q] = query[q].x; off[q + 1] = off[q] + (query[q].z - query[q].y); }
        std::vector<int> idx((size_t)off[nq]), dist((size_t)off[nq]);
        for (int q = 0; q < nq; q++) std::copy(idx2 + query[q].y, idx2 + query[q].z, idx.begin() + off[q]);
        Carver cv2(m->arena2);
        const size_t oF = cv2.take(sizeof(int) * (size_t)nq), oO = cv2.take(sizeof(int) * ((size_t)nq + 1)),
                     oX = cv2.take(sizeof(int) * idx.size()), oD = cv2.take(sizeof(int) * idx.size());
        rc = m->arena2.reserve(cv2.total());
        if (!rc) rc = m->arena2.mirror();
        if (rc) return rc;
        uint8_t* B2 = m->arena2.base;
        Xfer x(m->arena2);
        x.up(oF, qfeat.data(), sizeof(int) * (size_t)nq);
        x.up(oO, off.data(), sizeof(int) * ((size_t)nq + 1));
        x.up(oX, idx.data(), sizeof(int) * idx.size());
        ORBX_CUDA(x.send());
        k_triangulation_dist<<<(nq + 7) / 8, 256, 0, s>>>((const orbx_kp*)(B + oK1), (const uint4*)(B + oD1), (const float*)(B + oU1),
                                                          (const int*)(B2 + oF), nq, (const orbx_kp*)(B + oK2), (const uint4*)(B + oD2),
                                                          (const float*)(B + oU2), (const uint8_t*)(B + oH2), (const float*)(B + oSF),
                                                          (const float*)(B + oSG), T, (const int*)(B2 + oO), (const int*)(B2 + oX), (int*)(B2 + oD));
        ORBX_CUDA(cudaGetLastError());
        m->lastLaunches = 2;
        x.down(dist.data(), oD, sizeof(int) * dist.size());
        ORBX_CUDA(x.fetch());
        for (int q = 0; q < nq; q++) {
            for (int c = off[q]; c < off[q + 1]; c++)
                if (dist[c] <= TH_LOW) ent.push_back((uint32_t)idx[c] | ((uint32_t)dist[c] << 24));
            eoff[q + 1] = (int)ent.size();
        }
    }
    // sequential replay of the greedy part: vbMatched2 and the running bestDist (ties replace, `dist > bestDist` at :729)
    std::vector<uint8_t> vbMatched2(n2, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nm = 0;
    for (int q = 0; q < nq; q++) {
        const int f1 = query[q].x;
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int c = eoff[q]; c < eoff[q + 1]; c++) {
            const int f2 = (int)(ent[c] & 0xFFFFFFu), d = (int)(ent[c] >> 24);
            if (vbMatched2[f2]) continue;
            if (d > bestDist) continue;
            bestIdx2 = f2;
            bestDist = d;
        }
        if (bestIdx2 >= 0) {
            matches12[f1] = bestIdx2;
            vbMatched2[bestIdx2] = 1;
            nm++;
            if (checkOri) {
                float rot = k1[f1].angle - k2[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(f1);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { vbMatched2[matches12[j]] = 0; matches12[j] = -1; nm--; }
        }
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}


int orbm_distinctive_descriptors(orbx_matcher* m, const uint8_t* desc, const int* off, int npoints, int* best)
{
    if (!m || npoints < 0 || (npoints > 0 && (!off || !best))) {
        set_error("orbm_distinctive_descriptors: bad argument");
        return ORBX_ERR_ARG;
    }
    if (npoints == 0) return ORBX_OK;
    for (int p = 0; p < npoints; p++)
        if (off[p + 1] < off[p] || off[p + 1] - off[p] > 65535 || off[0] < 0) {
            set_error("orbm_distinctive_descriptors: off[%d..%d] = %d, %d (at most 65535 observations per point)", p, p + 1, off[p], off[p + 1]);
            return ORBX_ERR_ARG;
        }
    const int total = off[npoints];
    if (total > 0 && !desc) { set_error("orbm_distinctive_descriptors: bad argument"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oD = cv.take((size_t)std::max(total, 1) * 32), oO = cv.take(sizeof(int) * ((size_t)npoints + 1)), oB = cv.take(sizeof(int) * (size_t)npoints);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    Xfer x(m->arena);
    x.up(oO, off, sizeof(int) * ((size_t)npoints + 1));
    if (total > 0) x.up(oD, desc, (size_t)total * 32);
    ORBX_CUDA(x.send());
    const int blocks = std::min((npoints + DD_WARPS - 1) / DD_WARPS, m->smCount * 8);
    k_distinctive<<<blocks, DD_WARPS * 32, 0, s>>>((const uint4*)(B + oD), (const int*)(B + oO), npoints, (int*)(B + oB));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    x.down(best, oB, sizeof(int) * (size_t)npoints);
    ORBX_CUDA(x.fetch());
    return ORBX_OK;
}

}  // extern "C"
