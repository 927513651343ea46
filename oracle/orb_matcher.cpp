// orb_matcher.cpp -- CPU ORACLE (test infrastructure, not product).
// Restatement of ORBmatcher::DescriptorDistance (ORBmatcher.cc:1649-1665), the best/second-best inner
// loop shared by the SearchBy* routines (e.g. ORBmatcher.cc:205-230, 84-110), the Frame grid
// (Frame.cc:230-245, 342-407), ORBmatcher::SearchForInitialization (ORBmatcher.cc:405-520) with
// ComputeThreeMaxima (:1603-1644), Frame::ComputeStereoMatches (Frame.cc:481-655), SearchByProjection(Frame, MapPoints)
// (ORBmatcher.cc:45-129), SearchByProjection(CurrentFrame, LastFrame) (:1330-1472) and both SearchByBoW (:159-288, :522-655).
// The last four read SLAM objects (MapPoint, KeyFrame, DBoW2::FeatureVector) in the reference; here those reads are the
// plain arrays documented at each function, computed by the caller exactly as the reference computes them.
#include "orb_oracle.h"
#include "orb_internal.h"
#include <climits>
#include <cstring>
#include <algorithm>
#include <utility>

namespace orbo {

// 8 x (xor + SWAR popcount) on 32-bit words.
int descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t wa, wb;
        std::memcpy(&wa, a + 4 * i, 4);
        std::memcpy(&wb, b + 4 * i, 4);
        uint32_t v = wa ^ wb;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
    }
    return dist;
}

static const int GRID_COLS = 64, GRID_ROWS = 48;   // Frame.h:37-38
static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // ORBmatcher.cc:37-39

struct Grid {
    float minX, minY, invW, invH;
    std::vector<int> cell[GRID_COLS][GRID_ROWS];
};

static void three_maxima(std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
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

} // namespace orbo

struct orbo_grid { orbo::Grid g; };

extern "C" {

int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return orbo::descriptor_distance(a, b); }

void orbo_match_bruteforce(const uint8_t* q, int nq, const uint8_t* t, int nt, int* best_idx, int* best_dist, int* second_dist)
{
    for (int i = 0; i < nq; i++) {
        int b1 = 256, b2 = 256, bi = -1;
        for (int j = 0; j < nt; j++) {
            int d = orbo::descriptor_distance(q + 32 * (size_t)i, t + 32 * (size_t)j);
            if (d < b1) { b2 = b1; b1 = d; bi = j; }
            else if (d < b2) { b2 = d; }
        }
        best_idx[i] = bi; best_dist[i] = b1; second_dist[i] = b2;
    }
}

void orbo_match_candidates(const uint8_t* q, int nq, const uint8_t* t, int nt, const int* off, const int* idx,
                           int* best_idx, int* best_dist, int* second_idx, int* second_dist)
{
    (void)nt;
    for (int i = 0; i < nq; i++) {
        int b1 = 256, b2 = 256, bi = -1, si = -1;
        for (int c = off[i]; c < off[i + 1]; c++) {
            int j = idx[c];
            int d = orbo::descriptor_distance(q + 32 * (size_t)i, t + 32 * (size_t)j);
            if (d < b1) { b2 = b1; si = bi; b1 = d; bi = j; }
            else if (d < b2) { b2 = d; si = j; }
        }
        best_idx[i] = bi; best_dist[i] = b1; second_idx[i] = si; second_dist[i] = b2;
    }
}

orbo_grid* orbo_grid_build(const orbo_kp* kps, int n, float minX, float maxX, float minY, float maxY)
{
    orbo_grid* G = new orbo_grid;
    G->g.minX = minX; G->g.minY = minY;
    G->g.invW = static_cast<float>(orbo::GRID_COLS) / static_cast<float>(maxX - minX);   // Frame.cc ctor
    G->g.invH = static_cast<float>(orbo::GRID_ROWS) / static_cast<float>(maxY - minY);
    for (int i = 0; i < n; i++) {
        int px = (int)std::round((kps[i].x - minX) * G->g.invW);     // PosInGrid, Frame.cc:397-407
        int py = (int)std::round((kps[i].y - minY) * G->g.invH);
        if (px < 0 || px >= orbo::GRID_COLS || py < 0 || py >= orbo::GRID_ROWS) continue;
        G->g.cell[px][py].push_back(i);
    }
    return G;
}
void orbo_grid_destroy(orbo_grid* g) { delete g; }

static void grid_query(const orbo::Grid& g, const orbo_kp* kps, float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out)
{
    using namespace orbo;
    out.clear();
    const int nMinCellX = std::max(0, (int)std::floor((x - g.minX - r) * g.invW));
    if (nMinCellX >= GRID_COLS) return;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - g.minX + r) * g.invW));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - g.minY - r) * g.invH));
    if (nMinCellY >= GRID_ROWS) return;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - g.minY + r) * g.invH));
    if (nMaxCellY < 0) return;
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (int id : g.cell[ix][iy]) {
                const orbo_kp& kp = kps[id];
                if (checkLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float dx = kp.x - x, dy = kp.y - y;
                if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(id);
            }
}

int orbo_grid_query(const orbo_grid* G, const orbo_kp* kps, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap)
{
    std::vector<int> v;
    grid_query(G->g, kps, x, y, r, minLevel, maxLevel, v);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
    return (int)v.size();
}

int orbo_search_for_initialization(const orbo_kp* k1, const uint8_t* d1, int n1, const orbo_kp* k2, const uint8_t* d2, int n2,
                                   float minX, float maxX, float minY, float maxY,
                                   float* prev, int* m12, int windowSize, float nnratio, int checkOri)
{
    using namespace orbo;
    int nmatches = 0;
    for (int i = 0; i < n1; i++) m12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> matchedDist(n2, INT_MAX), m21(n2, -1);
    orbo_grid* G = orbo_grid_build(k2, n2, minX, maxX, minY, maxY);
    std::vector<int> cand;
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = k1[i1].octave;
        if (level1 > 0) continue;
        grid_query(G->g, k2, prev[2 * i1], prev[2 * i1 + 1], (float)windowSize, level1, level1, cand);
        if (cand.empty()) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : cand) {
            int dist = descriptor_distance(d1 + 32 * (size_t)i1, d2 + 32 * (size_t)i2);
            if (matchedDist[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (m21[bestIdx2] >= 0) { m12[m21[bestIdx2]] = -1; nmatches--; }
                m12[i1] = bestIdx2; m21[bestIdx2] = i1; matchedDist[bestIdx2] = bestDist; nmatches++;
                if (checkOri) {
                    float rot = k1[i1].angle - k2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    orbo_grid_destroy(G);
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) if (m12[idx1] >= 0) { m12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (m12[i1] >= 0) { prev[2 * i1] = k2[m12[i1]].x; prev[2 * i1 + 1] = k2[m12[i1]].y; }
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th)   (ORBmatcher.cc:45-129)
// Map points are passed after the mbTrackInView / isBad filters (:52-56).  Per point: proj_xy = (mTrackProjX, mTrackProjY),
// proj_xr = mTrackProjXR, level = mnTrackScaleLevel, view_cos = mTrackViewCos, its descriptor, has_obs = Observations() > 0.
// Frame: mvKeysUn, mDescriptors, mvuRight, occupied[j] = F.mvpMapPoints[j] && Observations() > 0 on entry, mvScaleFactors.
// feature_match[j] = index of the point stored into F.mvpMapPoints[j] by this call (-1: untouched).
int orbo_search_by_projection(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied_in, int n,
                              float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                              const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos,
                              const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, float nnratio, int* feature_match)
{
    using namespace orbo;
    (void)nlevels;
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nmp; iMP++) {
        const int nPredictedLevel = level[iMP];
        float r = view_cos[iMP] > 0.998 ? 2.5f : 4.0f;                    // RadiusByViewingCos (:131-137)
        if (bFactor) r *= th;
        grid_query(G->g, kps, proj_xy[2 * iMP], proj_xy[2 * iMP + 1], r * sf[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = mp_desc + 32 * (size_t)iMP;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (occupied[idx]) continue;                                   // :88-90
            if (u_right[idx] > 0) {
                const float er = std::fabs(proj_xr[iMP] - u_right[idx]);
                if (er > r * sf[nPredictedLevel]) continue;
            }
            const int dist = descriptor_distance(MPdescriptor, desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kps[idx].octave; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = kps[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            feature_match[bestIdx] = iMP;                                  // F.mvpMapPoints[bestIdx]=pMP
            occupied[bestIdx] = mp_has_obs[iMP];
            nmatches++;
        }
    }
    orbo_grid_destroy(G);
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)  (:1330-1472)
// Points of the last frame are passed after the pMP / mvbOutlier / invzc / image-bounds filters (:1354-1378), with the
// projection (u, v), ur = u - mbf*invzc (:1411), their octave and angle in the last frame, descriptor and has_obs.
// mode: 0 = neither, 1 = bForward, 2 = bBackward (:1351-1352).
int orbo_search_by_projection_frame(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied_in, int n,
                                    float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                    const float* proj_xy, const float* proj_ur, const int* last_octave, const float* last_angle,
                                    const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                    int* feature_match)
{
    using namespace orbo;
    (void)nlevels;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> vIndices2;
    for (int i = 0; i < nmp; i++) {
        const float u = proj_xy[2 * i], v = proj_xy[2 * i + 1];
        const int nLastOctave = last_octave[i];
        const float radius = th * sf[nLastOctave];
        if (mode == 1) grid_query(G->g, kps, u, v, radius, nLastOctave, -1, vIndices2);
        else if (mode == 2) grid_query(G->g, kps, u, v, radius, 0, nLastOctave, vIndices2);
        else grid_query(G->g, kps, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mp_desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (occupied[i2]) continue;
            if (u_right[i2] > 0) {
                const float er = std::fabs(proj_ur[i] - u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = descriptor_distance(dMP, desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            feature_match[bestIdx2] = i;
            occupied[bestIdx2] = mp_has_obs[i];
            nmatches++;
            if (checkOri) {
                float rot = last_angle[i] - kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    orbo_grid_destroy(G);
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int j : rotHist[i]) { feature_match[j] = -1; nmatches--; }
    }
    return nmatches;
}

// ---- ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (:159-288, variant 0) and SearchByBoW(KeyFrame*, KeyFrame*, ...)
// (:522-655, variant 1).  A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) is passed as CSR: ascending
// node ids, offsets, feature indices.  valid1[i]: feature i of the first key frame has a MapPoint that is not bad;
// valid2[j] (variant 1): same for the second key frame.  variant 0 writes matches[j] (n2 entries) = index of the first
// key frame's feature whose MapPoint went to feature j of the frame; variant 1 writes matches[i] (n1 entries) = matched
// feature of the second key frame.
int orbo_search_by_bow(int variant, const orbo_kp* k1, const uint8_t* d1, int n1, const uint8_t* valid1,
                       const int* node1, const int* off1, const int* idx1, int nn1,
                       const orbo_kp* k2, const uint8_t* d2, int n2, const uint8_t* valid2,
                       const int* node2, const int* off2, const int* idx2, int nn2,
                       float nnratio, int checkOri, int* matches)
{
    using namespace orbo;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    const int nout = variant == 0 ? n2 : n1;
    for (int i = 0; i < nout; i++) matches[i] = -1;
    std::vector<uint8_t> taken2(n2, 0);                 // vpMapPointMatches[realIdxF] != NULL (variant 0) / vbMatched2 (variant 1)
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (node1[a] == node2[b]) {
            for (int p1 = off1[a]; p1 < off1[a + 1]; p1++) {
                const int realIdx1 = idx1[p1];
                if (!valid1[realIdx1]) continue;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int p2 = off2[b]; p2 < off2[b + 1]; p2++) {
                    const int realIdx2 = idx2[p2];
                    if (taken2[realIdx2]) continue;
                    if (variant == 1 && !valid2[realIdx2]) continue;
                    const int dist = descriptor_distance(d1 + 32 * (size_t)realIdx1, d2 + 32 * (size_t)realIdx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = realIdx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                const bool low = variant == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;      // :228 vs :598
                if (low && (float)bestDist1 < nnratio * (float)bestDist2) {
                    taken2[bestIdx2] = 1;
                    if (variant == 0) matches[bestIdx2] = realIdx1; else matches[realIdx1] = bestIdx2;
                    if (checkOri) {
                        float rot = k1[realIdx1].angle - k2[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(variant == 0 ? bestIdx2 : realIdx1);
                    }
                    nmatches++;
                }
            }
            a++; b++;
        } else if (node1[a] < node2[b]) {
            a = (int)(std::lower_bound(node1, node1 + nn1, node2[b]) - node1);
        } else {
            b = (int)(std::lower_bound(node2, node2 + nn2, node1[a]) - node2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { matches[j] = -1; nmatches--; }
        }
    }
    return nmatches;
}

int orbo_stereo_matches(const orbo_kp* kl, const uint8_t* dl, int nl, const orbo_kp* kr, const uint8_t* dr, int nr,
                        const uint8_t* const* pyrL, const uint8_t* const* pyrR, const int* lw, const int* lh,
                        const size_t* strideL, const size_t* strideR,
                        const float* sf, const float* isf, int nlevels, float bf, float minD, float maxD,
                        float* uRight, float* depth)
{
    using namespace orbo;
    (void)nlevels; (void)lh;
    for (int i = 0; i < nl; i++) { uRight[i] = -1.0f; depth[i] = -1.0f; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = lh[0];
    std::vector<std::vector<int>> rows(nRows);
    for (int iR = 0; iR < nr; iR++) {
        const float kpY = kr[iR].y;
        const float r = 2.0f * sf[kr[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) rows[yi].push_back(iR);     // the reference indexes unchecked; in range for real keypoints
    }
    std::vector<std::pair<int, int>> distIdx;
    for (int iL = 0; iL < nl; iL++) {
        const orbo_kp& kpL = kl[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        const std::vector<int>& cand = rows[row];
        if (cand.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        int bestIdxR = 0;
        for (int iR : cand) {
            const orbo_kp& kpR = kr[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dl + 32 * (size_t)iL, dr + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kr[bestIdxR].x;
            const float scaleFactor = isf[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const int lev = kpL.octave;
            const uint8_t* IL = pyrL[lev];
            const uint8_t* IR = pyrR[lev];
            const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
            int bestSad = INT_MAX, bestincR = 0;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= lw[lev]) continue;
            const float cL = (float)IL[(size_t)cv * strideL[lev] + cu];
            for (int incR = -L; incR <= +L; incR++) {
                const float cR = (float)IR[(size_t)cv * strideR[lev] + cr + incR];
                double acc = 0;
                for (int dy = -w; dy <= w; dy++)
                    for (int dx = -w; dx <= w; dx++) {
                        float a = (float)IL[(size_t)(cv + dy) * strideL[lev] + cu + dx] - cL;
                        float b = (float)IR[(size_t)(cv + dy) * strideR[lev] + cr + incR + dx] - cR;
                        acc += std::fabs(a - b);
                    }
                float dist = (float)acc;
                if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = sf[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = bf / disparity;
                uRight[iL] = bestuR;
                distIdx.push_back(std::make_pair(bestSad, iL));
            }
        }
    }
    if (distIdx.empty()) return 0;        // the reference reads vDistIdx[0] of an empty vector here (Frame.cc:642)
    std::sort(distIdx.begin(), distIdx.end());
    const float median = (float)distIdx[distIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)distIdx.size();
    for (int i = (int)distIdx.size() - 1; i >= 0; i--) {
        if (distIdx[i].first < thDist) break;
        uRight[distIdx[i].second] = -1; depth[distIdx[i].second] = -1; kept--;
    }
    return kept;
}


// ================================================================================================================
// The searches of the LocalMapping / LoopClosing threads.  MapPoint / KeyFrame reads are plain arrays again: the caller
// runs the reference's own skip tests and projection (isBad, spAlreadyFound, depth sign, IsInImage, distance range, viewing
// angle, PredictScale) and passes, per surviving point, the projection (u, v[, ur]), the predicted level and the descriptor.

// The inner loop shared by Fuse(KF, MapPoints) (ORBmatcher.cc:886-935), Fuse(KF, Scw) (:1059-1077) and SearchBySim3
// (:1183-1203, :1263-1283): KeyFrame::GetFeaturesInArea(u, v, radius) (KeyFrame.cc:572-611: the Frame grid without a level
// filter), keep octave in [level-1, level], optionally the chi-square reprojection test of Fuse (:898-922), first minimum
// of the Hamming distance in visiting order.  best_idx = -1 / best_dist = 256 when nothing survives.
static void area_best(const orbo::Grid& g, const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip,
                      float u, float v, float ur, float radius, int nPredictedLevel, const uint8_t* dMP,
                      const float* invLevelSigma2, std::vector<int>& vIndices, int& bestIdx, int& bestDist)
{
    using namespace orbo;
    bestDist = 256;
    bestIdx = -1;
    grid_query(g, kps, u, v, radius, -1, -1, vIndices);
    for (int idx : vIndices) {
        if (skip && skip[idx]) continue;
        const orbo_kp& kp = kps[idx];
        const int kpLevel = kp.octave;
        if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
        if (invLevelSigma2) {
            if (u_right[idx] >= 0) {
                const float ex = u - kp.x, ey = v - kp.y, er = ur - u_right[idx];
                const float e2 = ex * ex + ey * ey + er * er;
                if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
            } else {
                const float ex = u - kp.x, ey = v - kp.y;
                const float e2 = ex * ex + ey * ey;
                if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
            }
        }
        const int dist = descriptor_distance(dMP, desc + 32 * (size_t)idx);
        if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
}

void orbo_search_in_area(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip, int n,
                         float minX, float maxX, float minY, float maxY, const float* qxy, const float* qur, const float* qr,
                         const int* qlevel, const uint8_t* qdesc, int nq, const float* inv_level_sigma2, int* best_idx, int* best_dist)
{
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> tmp;
    for (int i = 0; i < nq; i++)
        area_best(G->g, kps, desc, u_right, skip, qxy[2 * i], qxy[2 * i + 1], qur ? qur[i] : 0.0f, qr[i], qlevel[i],
                  qdesc + 32 * (size_t)i, inv_level_sigma2, tmp, best_idx[i], best_dist[i]);
    orbo_grid_destroy(G);
}

// ---- ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) (:827-975; stereo_chi2 = 1) and Fuse(KeyFrame*, Scw, vpPoints, th,
// vpReplacePoint) (:977-1102; stereo_chi2 = 0).  best_idx[i] = the key frame feature point i fuses into (bestDist <= TH_LOW,
// :938 / :1081), -1 otherwise; what happens to that feature's MapPoint (Replace / AddObservation) is map bookkeeping of the caller.
// Returns the number of points with a feature (nFused).
int orbo_fuse(const orbo_kp* kps, const uint8_t* desc, const float* u_right, int n, float minX, float maxX, float minY, float maxY,
              const float* sf, const float* inv_level_sigma2, int nlevels, const float* proj_xy, const float* proj_ur,
              const int* level, const uint8_t* mp_desc, int nmp, float th, int stereo_chi2, int* best_idx)
{
    using namespace orbo;
    (void)nlevels;
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> tmp;
    int nFused = 0;
    for (int i = 0; i < nmp; i++) {
        const float radius = th * sf[level[i]];
        int bestIdx, bestDist;
        area_best(G->g, kps, desc, u_right, nullptr, proj_xy[2 * i], proj_xy[2 * i + 1], stereo_chi2 ? proj_ur[i] : 0.0f, radius, level[i],
                  mp_desc + 32 * (size_t)i, stereo_chi2 ? inv_level_sigma2 : nullptr, tmp, bestIdx, bestDist);
        best_idx[i] = -1;
        if (bestDist <= TH_LOW) { best_idx[i] = bestIdx; nFused++; }
    }
    orbo_grid_destroy(G);
    return nFused;
}

// ---- ORBmatcher::SearchBySim3 (:1104-1328).  Direction 1->2: the MapPoints of key frame 1 that pass :1145-1180 projected
// into key frame 2 (proj12_xy, level12, desc of the MapPoint; level12[i] < 0 = point skipped), direction 2->1 likewise.
// match12[i1] = i2 where both directions agree (:1309-1322), -1 otherwise.  Returns nFound.
int orbo_search_by_sim3(const orbo_kp* k1, const uint8_t* d1, int n1, const orbo_kp* k2, const uint8_t* d2, int n2,
                        float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                        const float* proj12_xy, const int* level12, const uint8_t* mp1_desc,
                        const float* proj21_xy, const int* level21, const uint8_t* mp2_desc, float th, int* match12)
{
    using namespace orbo;
    (void)nlevels;
    std::vector<int> vnMatch1(n1, -1), vnMatch2(n2, -1), tmp;
    orbo_grid* G2 = orbo_grid_build(k2, n2, minX, maxX, minY, maxY);
    for (int i1 = 0; i1 < n1; i1++) {
        if (level12[i1] < 0) continue;
        int bestIdx, bestDist;
        area_best(G2->g, k2, d2, nullptr, nullptr, proj12_xy[2 * i1], proj12_xy[2 * i1 + 1], 0.0f, th * sf[level12[i1]], level12[i1],
                  mp1_desc + 32 * (size_t)i1, nullptr, tmp, bestIdx, bestDist);
        if (bestDist <= TH_HIGH) vnMatch1[i1] = bestIdx;
    }
    orbo_grid_destroy(G2);
    orbo_grid* G1 = orbo_grid_build(k1, n1, minX, maxX, minY, maxY);
    for (int i2 = 0; i2 < n2; i2++) {
        if (level21[i2] < 0) continue;
        int bestIdx, bestDist;
        area_best(G1->g, k1, d1, nullptr, nullptr, proj21_xy[2 * i2], proj21_xy[2 * i2 + 1], 0.0f, th * sf[level21[i2]], level21[i2],
                  mp2_desc + 32 * (size_t)i2, nullptr, tmp, bestIdx, bestDist);
        if (bestDist <= TH_HIGH) vnMatch2[i2] = bestIdx;
    }
    orbo_grid_destroy(G1);
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { match12[i1] = idx2; nFound++; }
    }
    return nFound;
}

// ---- ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (:290-403).  matched_in[j] = vpMatched[j] != NULL
// on entry; the points passed :316-351; a feature that receives a point is skipped by the later points (:375-376, :394).
// feature_match[j] = index of the point stored into vpMatched[j] by this call, -1 = untouched.
int orbo_search_by_projection_sim3(const orbo_kp* kps, const uint8_t* desc, const uint8_t* matched_in, int n,
                                   float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                   const float* proj_xy, const int* level, const uint8_t* mp_desc, int nmp, float th, int* feature_match)
{
    using namespace orbo;
    (void)nlevels;
    std::vector<uint8_t> matched(matched_in, matched_in + n);
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> tmp;
    int nmatches = 0;
    for (int i = 0; i < nmp; i++) {
        int bestIdx, bestDist;
        area_best(G->g, kps, desc, nullptr, matched.data(), proj_xy[2 * i], proj_xy[2 * i + 1], 0.0f, th * sf[level[i]], level[i],
                  mp_desc + 32 * (size_t)i, nullptr, tmp, bestIdx, bestDist);
        if (bestDist <= TH_LOW) { feature_match[bestIdx] = i; matched[bestIdx] = 1; nmatches++; }
    }
    orbo_grid_destroy(G);
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame*, sAlreadyFound, th, ORBdist) (:1474-1601), the
// relocalisation search.  occupied_in[j] = CurrentFrame.mvpMapPoints[j] != NULL; points of the key frame that passed
// :1494-1523, with level = nPredictedLevel and kf_angle = pKF->mvKeysUn[i].angle.
int orbo_search_by_projection_kf(const orbo_kp* kps, const uint8_t* desc, const uint8_t* occupied_in, int n,
                                 float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                 const float* proj_xy, const int* level, const float* kf_angle, const uint8_t* mp_desc, int nmp,
                                 float th, int ORBdist, int checkOri, int* feature_match)
{
    using namespace orbo;
    (void)nlevels;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<uint8_t> occupied(occupied_in, occupied_in + n);
    for (int j = 0; j < n; j++) feature_match[j] = -1;
    orbo_grid* G = orbo_grid_build(kps, n, minX, maxX, minY, maxY);
    std::vector<int> vIndices2;
    for (int i = 0; i < nmp; i++) {
        const int nPredictedLevel = level[i];
        const float radius = th * sf[nPredictedLevel];
        grid_query(G->g, kps, proj_xy[2 * i], proj_xy[2 * i + 1], radius, nPredictedLevel - 1, nPredictedLevel + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mp_desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (occupied[i2]) continue;
            const int dist = descriptor_distance(dMP, desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist) {
            feature_match[bestIdx2] = i;
            occupied[bestIdx2] = 1;
            nmatches++;
            if (checkOri) {
                float rot = kf_angle[i] - kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    orbo_grid_destroy(G);
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int j : rotHist[i]) { feature_match[j] = -1; nmatches--; }
    }
    return nmatches;
}

// ---- ORBmatcher::SearchForTriangulation (:657-825) with CheckDistEpipolarLine (:140-157).  has_mp1/has_mp2[i] = the key
// frame's feature already has a MapPoint (:699-703, :721-725); F12 row-major 3x3; (ex, ey) the epipole in image 2 (:667-672);
// sf2 / sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.  matches12[i1] = i2 or -1 (the vMatchedPairs list in index order).
int orbo_search_for_triangulation(const orbo_kp* k1, const uint8_t* d1, int n1, const uint8_t* has_mp1, const float* u_right1,
                                  const int* node1, const int* off1, const int* idx1, int nn1,
                                  const orbo_kp* k2, const uint8_t* d2, int n2, const uint8_t* has_mp2, const float* u_right2,
                                  const int* node2, const int* off2, const int* idx2, int nn2,
                                  const float* F12, float ex, float ey, const float* sf2, const float* sigma2_2, int nlevels,
                                  int onlyStereo, int checkOri, int* matches12)
{
    using namespace orbo;
    (void)nlevels;
    int nmatches = 0;
    std::vector<uint8_t> vbMatched2(n2, 0);
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (node1[a] == node2[b]) {
            for (int p1 = off1[a]; p1 < off1[a + 1]; p1++) {
                const int i1 = idx1[p1];
                if (has_mp1[i1]) continue;
                const bool bStereo1 = u_right1[i1] >= 0;
                if (onlyStereo && !bStereo1) continue;
                const orbo_kp& kp1 = k1[i1];
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int p2 = off2[b]; p2 < off2[b + 1]; p2++) {
                    const int i2 = idx2[p2];
                    if (vbMatched2[i2] || has_mp2[i2]) continue;
                    const bool bStereo2 = u_right2[i2] >= 0;
                    if (onlyStereo && !bStereo2) continue;
                    const int dist = descriptor_distance(d1 + 32 * (size_t)i1, d2 + 32 * (size_t)i2);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orbo_kp& kp2 = k2[i2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * sf2[kp2.octave]) continue;
                    }
                    // CheckDistEpipolarLine (:140-157)
                    const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                    const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                    const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                    const float num = la * kp2.x + lb * kp2.y + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) continue;
                    const float dsqr = num * num / den;
                    if (dsqr < 3.84 * sigma2_2[kp2.octave]) { bestIdx2 = i2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    matches12[i1] = bestIdx2;
                    vbMatched2[bestIdx2] = 1;
                    nmatches++;
                    if (checkOri) {
                        float rot = kp1.angle - k2[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(i1);
                    }
                }
            }
            a++; b++;
        } else if (node1[a] < node2[b]) {
            a = (int)(std::lower_bound(node1, node1 + nn1, node2[b]) - node1);
        } else {
            b = (int)(std::lower_bound(node2, node2 + nn2, node1[a]) - node2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { vbMatched2[matches12[j]] = 0; matches12[j] = -1; nmatches--; }
        }
    }
    return nmatches;
}


// ---- MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:252-312), the descriptor-level part (:281-311): of the n
// descriptors observing one MapPoint (rows of desc, in std::map<KeyFrame*, size_t> iteration order, bad key frames already
// removed), the index of the one with the least median distance to the rest.  -1 for n = 0 (:266-267, :277-278 return early).
int orbo_distinctive_descriptor(const uint8_t* desc, int n)
{
    using namespace orbo;
    if (n <= 0) return -1;
    const size_t N = (size_t)n;
    std::vector<float> Distances(N * N);
    for (size_t i = 0; i < N; i++) {
        Distances[i * N + i] = 0;
        for (size_t j = i + 1; j < N; j++) {
            const int distij = descriptor_distance(desc + 32 * i, desc + 32 * j);
            Distances[i * N + j] = distij;
            Distances[j * N + i] = distij;
        }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (size_t i = 0; i < N; i++) {
        std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + (i + 1) * N);
        std::sort(vDists.begin(), vDists.end());
        const int median = vDists[0.5 * (N - 1)];
        if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
    }
    return BestIdx;
}

// the same for npoints MapPoints at once: observations of point p are rows off[p] .. off[p+1] of desc
void orbo_distinctive_descriptors(const uint8_t* desc, const int* off, int npoints, int* best)
{
    for (int p = 0; p < npoints; p++) best[p] = orbo_distinctive_descriptor(desc + 32 * (size_t)off[p], off[p + 1] - off[p]);
}


// ---- Frame glue around the extractor (Frame.cc:419-479, :658-679).
// cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) (Frame.cc:438, :464): OpenCV is an un-vendored dependency; this is its
// published algorithm (calib3d undistortPoints, 5 fixed-point iterations, no EPS test with the default TermCriteria(MAX_ITER, 5, 0.01)),
// in double, checked bit for bit against cv2 4.13.0 on tests/golden/undistort_s9000.npz.  dist = (k1, k2, p1, p2[, k3]).
void orbo_undistort_points(const float* pts, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* out)
{
    double k[12] = {0};
    for (int i = 0; i < ndist && i < 12; i++) k[i] = dist[i];
    const double dfx = fx, dfy = fy, dcx = cx, dcy = cy, ifx = 1.0 / dfx, ify = 1.0 / dfy;
    for (int i = 0; i < n; i++) {
        const double u = pts[2 * i], v = pts[2 * i + 1];
        double x = (u - dcx) * ifx, y = (v - dcy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - dcx) * ifx; y = (v - dcy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        out[2 * i] = (float)(dfx * x + dcx);
        out[2 * i + 1] = (float)(dfy * y + dcy);
    }
}

// Frame::UndistortKeyPoints (Frame.cc:419-449): mvKeysUn = mvKeys when k1 == 0, else the same keypoints with pt undistorted.
void orbo_undistort_keypoints(const orbo_kp* kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, orbo_kp* out)
{
    for (int i = 0; i < n; i++) out[i] = kps[i];
    if (dist[0] == 0.0) return;
    std::vector<float> mat(2 * (size_t)n), und(2 * (size_t)n);
    for (int i = 0; i < n; i++) { mat[2 * i] = kps[i].x; mat[2 * i + 1] = kps[i].y; }
    orbo_undistort_points(mat.data(), n, fx, fy, cx, cy, dist, ndist, und.data());
    for (int i = 0; i < n; i++) { out[i].x = und[2 * i]; out[i].y = und[2 * i + 1]; }
}

// Frame::ComputeImageBounds (Frame.cc:451-479) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)
void orbo_image_bounds(int cols, int rows, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* bounds4)
{
    if (dist[0] != 0.0) {
        const float mat[8] = {0.0f, 0.0f, (float)cols, 0.0f, 0.0f, (float)rows, (float)cols, (float)rows};
        float u[8];
        orbo_undistort_points(mat, 4, fx, fy, cx, cy, dist, ndist, u);
        bounds4[0] = std::min(u[0], u[4]);
        bounds4[1] = std::max(u[2], u[6]);
        bounds4[2] = std::min(u[1], u[3]);
        bounds4[3] = std::max(u[5], u[7]);
    } else {
        bounds4[0] = 0.0f; bounds4[1] = (float)cols; bounds4[2] = 0.0f; bounds4[3] = (float)rows;
    }
}

// Frame::ComputeStereoFromRGBD (Frame.cc:658-679): imDepth.at<float>(v, u) with float v, u truncates to int indices.
void orbo_stereo_from_rgbd(const orbo_kp* kps, const orbo_kp* kps_un, int n, const float* im_depth, size_t stride_bytes, float bf,
                           float* u_right, float* depth)
{
    for (int i = 0; i < n; i++) {
        u_right[i] = -1;
        depth[i] = -1;
        const float v = kps[i].y, u = kps[i].x;
        const float d = *(const float*)((const uint8_t*)im_depth + (size_t)(int)v * stride_bytes + sizeof(float) * (size_t)(int)u);
        if (d > 0) {
            depth[i] = d;
            u_right[i] = kps_un[i].x - bf / d;
        }
    }
}

} // extern "C"
