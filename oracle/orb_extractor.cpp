// orb_extractor.cpp -- CPU ORACLE (test infrastructure, not product).
// Restatement of ORBextractor (reference src/ORBextractor.cc): ctor :410-470, ComputePyramid :1107-1132,
// ComputeKeyPointsOctTree :765-853, IC_Angle :77-104, computeOrbDescriptor :108-147, operator() :1043-1105.
#include "orb_oracle.h"
#include "orb_internal.h"
#include <chrono>
#include <cstring>
#include <algorithm>

namespace orbo {

static const int PATCH_SIZE = 31;
static const int HALF_PATCH_SIZE = 15;
static const int EDGE_THRESHOLD = 19;

static const int PATTERN[256 * 4] = {
#include "orb_pattern.inc"
};

struct Level {
    int w = 0, h = 0;
    std::vector<uint8_t> padded;          // (w+38) x (h+38)
    std::vector<uint8_t> blurred;         // w x h
    std::vector<XYR> candidates;          // relative to minBorder
    std::vector<orbo_kp> keys;            // level coordinates, angle filled
    size_t pstride() const { return (size_t)w + 2 * EDGE_THRESHOLD; }
    const uint8_t* interior() const { return padded.data() + (size_t)EDGE_THRESHOLD * pstride() + EDGE_THRESHOLD; }
    uint8_t* interior() { return padded.data() + (size_t)EDGE_THRESHOLD * pstride() + EDGE_THRESHOLD; }
};

struct Extractor {
    int nfeatures; double scaleFactor; int nlevels, iniThFAST, minThFAST;
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> perLevel, umax;
    std::vector<Level> lv;
    int tie_mode = 0, fma_mode = 0;
    double secs[6] = {0, 0, 0, 0, 0, 0};

    Extractor(int nf, float s, int nl, int ini, int mn)
        : nfeatures(nf), scaleFactor(s), nlevels(nl), iniThFAST(ini), minThFAST(mn)
    {
        sf.resize(nl); sig2.resize(nl); isf.resize(nl); isig2.resize(nl);
        sf[0] = 1.0f; sig2[0] = 1.0f;
        for (int i = 1; i < nl; i++) {
            sf[i] = (float)(sf[i - 1] * scaleFactor);      // float * double member, rounded once to float
            sig2[i] = sf[i] * sf[i];
        }
        for (int i = 0; i < nl; i++) { isf[i] = 1.0f / sf[i]; isig2[i] = 1.0f / sig2[i]; }
        lv.resize(nl);

        perLevel.resize(nl);
        float factor = (float)(1.0f / scaleFactor);
        float want = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) {
            perLevel[l] = cv_round(want);
            sum += perLevel[l];
            want *= factor;
        }
        perLevel[nl - 1] = std::max(nfeatures - sum, 0);

        umax.assign(HALF_PATCH_SIZE + 1, 0);
        int vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
        int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
        const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
        for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
        for (int v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }

    void computePyramid(const uint8_t* img, int w, int h, size_t stride)
    {
        for (int l = 0; l < nlevels; l++) {
            Level& L = lv[l];
            float scale = isf[l];
            L.w = cv_round((float)w * scale);
            L.h = cv_round((float)h * scale);
            L.padded.assign(L.pstride() * (size_t)(L.h + 2 * EDGE_THRESHOLD), 0);
            if (l != 0) {
                const Level& P = lv[l - 1];
                resize_linear_u8(P.interior(), P.w, P.h, P.pstride(), L.interior(), L.w, L.h, L.pstride());
                // in-place border from the freshly resized interior (REFLECT_101 + ISOLATED)
                std::vector<uint8_t> tmp((size_t)L.w * L.h);
                for (int y = 0; y < L.h; y++) std::memcpy(&tmp[(size_t)y * L.w], L.interior() + (size_t)y * L.pstride(), L.w);
                border_reflect101(tmp.data(), L.w, L.h, L.w, L.padded.data(), EDGE_THRESHOLD, L.pstride());
            } else {
                border_reflect101(img, w, h, stride, L.padded.data(), EDGE_THRESHOLD, L.pstride());
            }
        }
    }

    float icAngle(const Level& L, int px, int py) const
    {
        int m_01 = 0, m_10 = 0;
        const int step = (int)L.pstride();
        const uint8_t* center = L.interior() + (size_t)py * step + px;
        for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
        for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
            int v_sum = 0, d = umax[v];
            for (int u = -d; u <= d; ++u) {
                int plus = center[u + v * step], minus = center[u - v * step];
                v_sum += plus - minus;
                m_10 += u * (plus + minus);
            }
            m_01 += v * v_sum;
        }
        return fast_atan2((float)m_01, (float)m_10);
    }

    void computeKeyPoints()
    {
        using clk = std::chrono::steady_clock;
        const float W = 30;
        for (int l = 0; l < nlevels; l++) {
            auto t0 = clk::now();
            Level& L = lv[l];
            L.candidates.clear();
            L.keys.clear();
            const int minBX = EDGE_THRESHOLD - 3, minBY = minBX;
            const int maxBX = L.w - EDGE_THRESHOLD + 3, maxBY = L.h - EDGE_THRESHOLD + 3;
            const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
            const int nCols = (int)(width / W), nRows = (int)(height / W);
            if (nCols <= 0 || nRows <= 0) continue;     // reference divides by zero; shapes this small are rejected upstream
            const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
            std::vector<XYR> cell;
            for (int i = 0; i < nRows; i++) {
                const float iniY = (float)(minBY + i * hCell);
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBY - 3) continue;
                if (maxY > maxBY) maxY = (float)maxBY;
                for (int j = 0; j < nCols; j++) {
                    const float iniX = (float)(minBX + j * wCell);
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBX - 6) continue;
                    if (maxX > maxBX) maxX = (float)maxBX;
                    const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
                    const uint8_t* roi = L.interior() + (size_t)y0 * L.pstride() + x0;
                    fast_detect(roi, x1 - x0, y1 - y0, L.pstride(), iniThFAST, true, cell);
                    if (cell.empty()) fast_detect(roi, x1 - x0, y1 - y0, L.pstride(), minThFAST, true, cell);
                    for (const XYR& k : cell) L.candidates.push_back(XYR{k.x + j * wCell, k.y + i * hCell, k.r});
                }
            }
            auto t1 = clk::now();
            std::vector<int> sel = distribute_octtree(L.candidates, minBX, maxBX, minBY, maxBY, perLevel[l], tie_mode);
            const int scaledPatchSize = (int)(PATCH_SIZE * sf[l]);
            for (int idx : sel) {
                const XYR& c = L.candidates[idx];
                orbo_kp k;
                k.x = (float)c.x + minBX; k.y = (float)c.y + minBY;
                k.size = (float)scaledPatchSize; k.angle = -1.f; k.response = (float)c.r;
                k.octave = l; k.class_id = -1;
                L.keys.push_back(k);
            }
            auto t2 = clk::now();
            secs[1] += std::chrono::duration<double>(t1 - t0).count();
            secs[2] += std::chrono::duration<double>(t2 - t1).count();
        }
        auto t3 = clk::now();
        for (int l = 0; l < nlevels; l++)
            for (orbo_kp& k : lv[l].keys) k.angle = icAngle(lv[l], cv_round(k.x), cv_round(k.y));
        secs[3] += std::chrono::duration<double>(clk::now() - t3).count();
    }

    void describe(const Level& L, const orbo_kp& kp, uint8_t* desc) const
    {
        const float factorPI = (float)(3.14159265358979323846 / 180.f);
        float angle = (float)kp.angle * factorPI;
        float a = std::cos(angle), b = std::sin(angle);      // float overloads -> cosf / sinf, as in the reference
        const uint8_t* img = L.blurred.data();
        const int step = L.w;
        const uint8_t* center = img + (size_t)cv_round(kp.y) * step + cv_round(kp.x);
        auto sample = [&](const int* p) -> int {
            const float px = (float)p[0], py = (float)p[1];
            int r, c;
            if (fma_mode) {
                r = cv_round(std::fmaf(px, b, py * a));
                c = cv_round(std::fmaf(px, a, -(py * b)));
            } else {
                volatile float t1 = px * b, t2 = py * a, t3 = px * a, t4 = py * b;
                r = cv_round((float)(t1 + t2));
                c = cv_round((float)(t3 - t4));
            }
            return center[r * step + c];
        };
        const int* pat = PATTERN;
        for (int i = 0; i < 32; ++i, pat += 32) {
            int val = 0;
            for (int bit = 0; bit < 8; bit++) {
                int t0 = sample(pat + 4 * bit), t1 = sample(pat + 4 * bit + 2);
                val |= (t0 < t1) << bit;
            }
            desc[i] = (uint8_t)val;
        }
    }

    int extract(const uint8_t* img, int w, int h, size_t stride, orbo_kp* kps, uint8_t* desc, int cap)
    {
        using clk = std::chrono::steady_clock;
        for (double& s : secs) s = 0;
        if (!img || w <= 0 || h <= 0) return 0;
        auto t0 = clk::now();
        computePyramid(img, w, h, stride);
        secs[0] = std::chrono::duration<double>(clk::now() - t0).count();
        computeKeyPoints();
        int n = 0;
        for (int l = 0; l < nlevels; l++) {
            Level& L = lv[l];
            if (L.keys.empty()) { L.blurred.clear(); continue; }
            auto t1 = clk::now();
            L.blurred.resize((size_t)L.w * L.h);
            gauss7(L.interior(), L.w, L.h, L.pstride(), L.blurred.data(), L.w);
            auto t2 = clk::now();
            for (const orbo_kp& k : L.keys) {
                if (n < cap) {
                    if (desc) describe(L, k, desc + (size_t)n * 32);
                    if (kps) {
                        kps[n] = k;
                        if (l != 0) { kps[n].x = k.x * sf[l]; kps[n].y = k.y * sf[l]; }
                    }
                }
                n++;
            }
            secs[4] += std::chrono::duration<double>(t2 - t1).count();
            secs[5] += std::chrono::duration<double>(clk::now() - t2).count();
        }
        return n;
    }
};

} // namespace orbo

struct orbo_extractor { orbo::Extractor impl; orbo_extractor(int a, float b, int c, int d, int e) : impl(a, b, c, d, e) {} };

extern "C" {
orbo_extractor* orbo_create(int nf, float s, int nl, int ini, int mn) { return new orbo_extractor(nf, s, nl, ini, mn); }
void orbo_destroy(orbo_extractor* e) { delete e; }
void orbo_set_modes(orbo_extractor* e, int tie, int fma) { e->impl.tie_mode = tie; e->impl.fma_mode = fma; }
int orbo_extract(orbo_extractor* e, const uint8_t* img, int w, int h, size_t stride, orbo_kp* kps, uint8_t* desc, int cap)
{ return e->impl.extract(img, w, h, stride, kps, desc, cap); }
void orbo_scale_tables(orbo_extractor* e, float* a, float* b, float* c, float* d)
{
    for (int i = 0; i < e->impl.nlevels; i++) {
        if (a) a[i] = e->impl.sf[i];
        if (b) b[i] = e->impl.isf[i];
        if (c) c[i] = e->impl.sig2[i];
        if (d) d[i] = e->impl.isig2[i];
    }
}
void orbo_features_per_level(orbo_extractor* e, int* out) { for (int i = 0; i < e->impl.nlevels; i++) out[i] = e->impl.perLevel[i]; }
void orbo_umax(orbo_extractor* e, int* out) { for (int i = 0; i < 16; i++) out[i] = e->impl.umax[i]; }
void orbo_level_size(orbo_extractor* e, int l, int* w, int* h) { *w = e->impl.lv[l].w; *h = e->impl.lv[l].h; }
void orbo_level_padded(orbo_extractor* e, int l, uint8_t* dst)
{ const auto& L = e->impl.lv[l]; std::memcpy(dst, L.padded.data(), L.padded.size()); }
void orbo_level_blurred(orbo_extractor* e, int l, uint8_t* dst)
{
    const auto& L = e->impl.lv[l];
    if (!L.blurred.empty()) { std::memcpy(dst, L.blurred.data(), L.blurred.size()); return; }
    orbo::gauss7(L.interior(), L.w, L.h, L.pstride(), dst, L.w);
}
int orbo_level_candidates(orbo_extractor* e, int l, int* xyr, int cap)
{
    const auto& c = e->impl.lv[l].candidates;
    for (size_t i = 0; i < c.size() && (int)i < cap; i++) { xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].r; }
    return (int)c.size();
}
int orbo_level_nkeys(orbo_extractor* e, int l) { return (int)e->impl.lv[l].keys.size(); }
void orbo_stage_seconds(orbo_extractor* e, double* out) { for (int i = 0; i < 6; i++) out[i] = e->impl.secs[i]; }
}
