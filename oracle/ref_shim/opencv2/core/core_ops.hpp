// core_ops.hpp -- CV SHIM FOR THE REFERENCE BUILD (test infrastructure, not product): the few CV_32F
// matrix operations the reference's ORBmatcher.cc writes (products, sums, transposes, scalar scaling,
// dot, L2 norm).  Evaluated eagerly; products, dot and norm accumulate in double and round once.  Real OpenCV evaluates lazily: `A*b + c`
// with a 3 x 3 A is ONE gemm whose small-matrix path sums float products and then forms (float)(t*1.0 + c*1.0) in double, so a projection
// can differ from this shim's in the last bit.  That does not enter any comparison: wherever the reference's sources and the drop-in headers
// are held against each other, both are compiled over this same algebra (and in a real build both sit on the same OpenCV) -- the drop-in
// templates write the reference's own cv::Mat expressions (include/ORBmatcher_reference_api.inl).
#pragma once

namespace cv {

namespace shim {
static inline void need32f(const Mat& m) { if (m.depth() != CV_32F) throw std::runtime_error("cv shim: CV_32F expected"); }
template <class F> static inline Mat zip(const Mat& a, const Mat& b, F f)
{
    need32f(a); need32f(b);
    if (a.rows != b.rows || a.cols != b.cols) throw std::runtime_error("cv shim: size mismatch");
    Mat r(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) {
        const float *pa = a.ptr<float>(y), *pb = b.ptr<float>(y);
        float* pr = r.ptr<float>(y);
        for (int x = 0; x < a.cols; x++) pr[x] = f(pa[x], pb[x]);
    }
    return r;
}
template <class F> static inline Mat map(const Mat& a, F f)
{
    need32f(a);
    Mat r(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) {
        const float* pa = a.ptr<float>(y);
        float* pr = r.ptr<float>(y);
        for (int x = 0; x < a.cols; x++) pr[x] = f(pa[x]);
    }
    return r;
}
} // namespace shim

static inline Mat operator+(const Mat& a, const Mat& b) { return shim::zip(a, b, [](float p, float q) { return p + q; }); }
static inline Mat operator-(const Mat& a, const Mat& b) { return shim::zip(a, b, [](float p, float q) { return p - q; }); }
static inline Mat operator-(const Mat& a) { return shim::map(a, [](float p) { return -p; }); }
static inline Mat operator*(const Mat& a, double s) { return shim::map(a, [s](float p) { return (float)(p * s); }); }
static inline Mat operator*(double s, const Mat& a) { return a * s; }
static inline Mat operator/(const Mat& a, double s) { return shim::map(a, [s](float p) { return (float)(p / s); }); }
static inline MatExpr operator*(double s, const MatExpr& e) { MatExpr r = e; r.fill = (float)(e.fill * s); return r; }     // s * Mat::ones(...)
static inline MatExpr operator*(const MatExpr& e, double s) { return s * e; }
// a -/+ (s * Mat::ones): per element a -/+ fill in float, what the evaluated expression gives (the constant matrix is not materialised)
static inline Mat operator-(const Mat& a, const MatExpr& e)
{
    if (e.eye || a.rows != e.rows || a.cols != e.cols) return a - Mat(e);
    const float c = e.fill;
    return shim::map(a, [c](float p) { return p - c; });
}
static inline Mat operator+(const Mat& a, const MatExpr& e)
{
    if (e.eye || a.rows != e.rows || a.cols != e.cols) return a + Mat(e);
    const float c = e.fill;
    return shim::map(a, [c](float p) { return p + c; });
}
static inline Mat operator*(const Mat& a, const Mat& b)
{
    shim::need32f(a); shim::need32f(b);
    if (a.cols != b.rows) throw std::runtime_error("cv shim: product size mismatch");
    Mat r(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            double s = 0;
            for (int k = 0; k < a.cols; k++) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
            r.at<float>(y, x) = (float)s;
        }
    return r;
}
inline Mat Mat::t() const
{
    shim::need32f(*this);
    Mat r(cols, rows, CV_32F);
    for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) r.at<float>(x, y) = at<float>(y, x);
    return r;
}
inline double Mat::dot(const Mat& o) const
{
    shim::need32f(*this); shim::need32f(o);
    if (total() != o.total()) throw std::runtime_error("cv shim: dot size mismatch");
    double s = 0;
    const int n = (int)total();
    for (int i = 0; i < n; i++) s += (double)at<float>(i / cols, i % cols) * (double)o.at<float>(i / o.cols, i % o.cols);
    return s;
}
inline Mat Mat::mul(const Mat& o) const { return shim::zip(*this, o, [](float p, float q) { return p * q; }); }
inline Mat Mat::inv() const { throw std::runtime_error("cv shim: Mat::inv not provided"); }
inline void Mat::convertTo(Mat& dst, int rtype) const
{
    Mat r(rows, cols, rtype & 7);
    if (depth() == CV_8U && (rtype & 7) == CV_32F) {        // the one conversion on a hot path (Frame.cc:579, :596): no per-element dispatch
        for (int y = 0; y < rows; y++) {
            const uchar* s = ptr<uchar>(y);
            float* d = r.ptr<float>(y);
            for (int x = 0; x < cols; x++) d[x] = (float)s[x];
        }
        dst = r;
        return;
    }
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            double v = depth() == CV_8U ? (double)at<uchar>(y, x) : depth() == CV_32F ? (double)at<float>(y, x)
                     : depth() == CV_64F ? at<double>(y, x) : (double)at<int>(y, x);
            switch (rtype & 7) {
            case CV_8U: r.at<uchar>(y, x) = (uchar)std::min(std::max(cvRound(v), 0), 255); break;
            case CV_32F: r.at<float>(y, x) = (float)v; break;
            case CV_64F: r.at<double>(y, x) = v; break;
            default: r.at<int>(y, x) = cvRound(v); break;
            }
        }
    dst = r;
}
static inline double norm(const Mat& a, int normType = NORM_L2)
{
    shim::need32f(a);
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            const double v = a.at<float>(y, x);
            if (normType == NORM_L1) s += std::fabs(v); else if (normType == NORM_INF) s = std::max(s, std::fabs(v)); else s += v * v;
        }
    return normType == NORM_L2 ? std::sqrt(s) : s;
}
// norm(a, b, type) = norm(a - b, type), CV_32F, accumulated in double (normDiffL1_32f and friends)
static inline double norm(const Mat& a, const Mat& b, int normType = NORM_L2)
{
    shim::need32f(a); shim::need32f(b);
    if (a.rows != b.rows || a.cols != b.cols) throw std::runtime_error("cv shim: norm size mismatch");
    double s = 0;
    for (int y = 0; y < a.rows; y++) {
        const float *pa = a.ptr<float>(y), *pb = b.ptr<float>(y);
        for (int x = 0; x < a.cols; x++) {
            const double v = (double)pa[x] - (double)pb[x];
            if (normType == NORM_L1) s += std::fabs(v); else if (normType == NORM_INF) s = std::max(s, std::fabs(v)); else s += v * v;
        }
    }
    return normType == NORM_L2 ? std::sqrt(s) : s;
}

} // namespace cv
