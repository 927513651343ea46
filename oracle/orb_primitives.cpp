// orb_primitives.cpp -- CPU ORACLE (test infrastructure, not product): scalar restatements of the
// OpenCV primitives the reference's ORBextractor calls.  OpenCV is NOT vendored by the reference
// (CMakeLists.txt:52-58, un-pinned); semantics here are pinned to cv2 4.13.0 by tests/test_oracle_vs_cv2.py.
//   cv::resize INTER_LINEAR u8   <- called at ORBextractor.cc:1120
//   cv::copyMakeBorder REFLECT_101 <- ORBextractor.cc:1122,1127
//   cv::GaussianBlur 7x7 s=2     <- ORBextractor.cc:1086
//   cv::FAST(9_16, nms)          <- ORBextractor.cc:809,814
//   cv::fastAtan2                <- ORBextractor.cc:103
//   cvRound                      <- ORBextractor.cc:81,115,119-120,442,460,1112
#include "orb_oracle.h"
#include "orb_internal.h"
#include <cmath>
#include <cfloat>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orbo {

// ---------------------------------------------------------------- resize (INTER_LINEAR, 8-bit, 1 channel)
// OpenCV fixed-point path: 11-bit coefficients, horizontal pass in int32, vertical pass
// (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2.
void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                      uint8_t* dst, int dw, int dh, size_t dstride)
{
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> xa(2 * dw), ya(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0.f; sx = 0; }
        if (sx + 1 >= sw) { fx = 0.f; sx = sw - 1; }
        xofs[dx] = sx;
        xa[2 * dx] = (short)cv_round((1.f - fx) * 2048.f);
        xa[2 * dx + 1] = (short)cv_round(fx * 2048.f);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        yofs[dy] = sy;
        ya[2 * dy] = (short)cv_round((1.f - fy) * 2048.f);
        ya[2 * dy + 1] = (short)cv_round(fy * 2048.f);
    }
    std::vector<int> row0(dw), row1(dw);
    for (int dy = 0; dy < dh; dy++) {
        int sy0 = std::min(std::max(yofs[dy], 0), sh - 1);
        int sy1 = std::min(std::max(yofs[dy] + 1, 0), sh - 1);
        const uint8_t* S0 = src + (size_t)sy0 * sstride;
        const uint8_t* S1 = src + (size_t)sy1 * sstride;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx];
            int sx1 = std::min(sx + 1, sw - 1);
            int a0 = xa[2 * dx], a1 = xa[2 * dx + 1];
            row0[dx] = S0[sx] * a0 + S0[sx1] * a1;
            row1[dx] = S1[sx] * a0 + S1[sx1] * a1;
        }
        int b0 = ya[2 * dy], b1 = ya[2 * dy + 1];
        uint8_t* D = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; dx++) {
            int v = (((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------- copyMakeBorder(BORDER_REFLECT_101)
static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

void border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, int b, size_t dstride)
{
    for (int y = -b; y < h + b; y++) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + b) * dstride;
        for (int x = -b; x < w + b; x++) D[x + b] = S[reflect101(x, w)];
    }
}

// ---------------------------------------------------------------- GaussianBlur(7x7, sigma=2), 8-bit fixed-point path
// OpenCV 4.x bit-exact u8 path: taps 18,34,48,56,48,34,18 (sum 256), horizontal 8.8, vertical 8.16, round to nearest.
void gauss7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> tmp((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (size_t)y * sstride;
        uint16_t* T = &tmp[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            int acc = 0;
            if (x >= 3 && x + 3 < w) for (int k = 0; k < 7; k++) acc += K[k] * S[x + k - 3];
            else for (int k = 0; k < 7; k++) acc += K[k] * S[reflect101(x + k - 3, w)];
            T[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t* D = dst + (size_t)y * dstride;
        const uint16_t* R[7];
        for (int k = 0; k < 7; k++) R[k] = &tmp[(size_t)reflect101(y + k - 3, h) * w];
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)K[k] * R[k][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// ---------------------------------------------------------------- FAST-9/16 with 3x3 non-max suppression
// Ring in OpenCV's order (circular); raw = max over 16 arcs of 9 of min(ring - centre) resp. (centre - ring).
// corner iff raw > t;  score (= KeyPoint::response) = raw - 1.
static const int RING[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

int fast_raw(const uint8_t* p, size_t stride)
{
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; k++) d[k] = (int)p[(ptrdiff_t)RING[k][1] * (ptrdiff_t)stride + RING[k][0]] - v;
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = -255;
    for (int s = 0; s < 16; s++) {
        int mb = 255, md = 255;
        for (int k = 0; k < 9; k++) { mb = std::min(mb, d[s + k]); md = std::min(md, -d[s + k]); }
        best = std::max(best, std::max(mb, md));
    }
    return best;
}

void fast_detect(const uint8_t* roi, int w, int h, size_t stride, int threshold, bool nms, std::vector<XYR>& out)
{
    out.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return;
    std::vector<int> score((size_t)w * h, 0);   // 0 for non-corners and the 3-px frame, like OpenCV's zeroed row buffers
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            // OpenCV-style early exit (does not change the result): an arc of 9 contiguous ring pixels always contains two
            // adjacent compass pixels (ring positions 0, 4, 8, 12), so both must be brighter (or darker) than centre +- t.
            const uint8_t* p = roi + (size_t)y * stride + x;
            const int v = p[0], hi = v + threshold, lo = v - threshold;
            const int p0 = p[3 * (ptrdiff_t)stride], p4 = p[3], p8 = p[-3 * (ptrdiff_t)stride], p12 = p[-3];
            const bool br = ((p0 > hi) || (p8 > hi)) && ((p4 > hi) || (p12 > hi));
            const bool dk = ((p0 < lo) || (p8 < lo)) && ((p4 < lo) || (p12 < lo));
            if (!br && !dk) { score[(size_t)y * w + x] = -1; continue; }
            int raw = fast_raw(p, stride);
            if (raw > threshold) score[(size_t)y * w + x] = raw - 1;
            else score[(size_t)y * w + x] = -1;  // marker: not a corner
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int s = score[(size_t)y * w + x];
            if (s < 0) continue;
            bool keep = true;
            if (nms) {
                for (int dy = -1; dy <= 1 && keep; dy++)
                    for (int dx = -1; dx <= 1; dx++) {
                        if (!dx && !dy) continue;
                        int n = score[(size_t)(y + dy) * w + (x + dx)];
                        if (n < 0) n = 0;
                        if (!(s > n)) { keep = false; break; }
                    }
            }
            if (keep) out.push_back(XYR{x, y, s});
        }
}

// ---------------------------------------------------------------- cv::fastAtan2 (degrees, float32 polynomial)
float fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    volatile float ax = std::fabs(x), ay = std::fabs(y);   // volatile: keep every op a separately rounded float op
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        volatile float t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1; t = t * c;
        a = t;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        volatile float t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1; t = t * c;
        a = 90.f - t;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

} // namespace orbo

extern "C" {
void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh, size_t dstride)
{ orbo::resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride); }
void orbo_border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, int b, size_t dstride)
{ orbo::border_reflect101(src, w, h, sstride, dst, b, dstride); }
void orbo_gauss7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride)
{ orbo::gauss7(src, w, h, sstride, dst, dstride); }
int orbo_fast(const uint8_t* roi, int w, int h, size_t stride, int threshold, int nms, int* xyr, int cap)
{
    std::vector<orbo::XYR> v;
    orbo::fast_detect(roi, w, h, stride, threshold, nms != 0, v);
    int n = (int)v.size();
    for (int i = 0; i < n && i < cap; i++) { xyr[3 * i] = v[i].x; xyr[3 * i + 1] = v[i].y; xyr[3 * i + 2] = v[i].r; }
    return n;
}
float orbo_fast_atan2(float y, float x) { return orbo::fast_atan2(y, x); }

// cvtColor(src, dst, CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) as called by Tracking::GrabImage* (Tracking.cc:280-305,
// :320-334, :350-364).  OpenCV (un-vendored dependency; arithmetic checked against cv2 4.13.0, tests/golden/cvtcolor_*.npz):
// 8-bit path RGB2Gray<uchar>, coefficients RY15 = 9798, GY15 = 19235, BY15 = 3735, CV_DESCALE(x, 15) = (x + (1 << 14)) >> 15.
void orbo_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int rgb, uint8_t* dst, size_t dstride)
{
    const int cr = 9798, cg = 19235, cb = 3735;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const uint8_t* p = src + (size_t)y * sstride + (size_t)x * channels;
            const int r = rgb ? p[0] : p[2], g = p[1], b = rgb ? p[2] : p[0];
            dst[(size_t)y * dstride + x] = (uint8_t)((r * cr + g * cg + b * cb + (1 << 14)) >> 15);
        }
}
}
