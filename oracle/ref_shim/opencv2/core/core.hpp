// opencv2/core/core.hpp -- CV SHIM FOR THE REFERENCE BUILD (test infrastructure, not product).
//
// The smallest `cv::` surface over which the UNMODIFIED reference sources
// (/root/reference/src/ORBextractor.cc, src/ORBmatcher.cc) compile with g++ alone: this image has no
// OpenCV C++ headers.  Written from OpenCV's public API; nothing here is taken from the reference.  The
// pixel primitives (FAST, resize, copyMakeBorder, GaussianBlur, fastAtan2) forward to the scalar
// restatements in oracle/orb_primitives.cpp, which tests/test_oracle_vs_cv2.py pins bit for bit to cv2
// 4.13.0.  Built only by `make -C oracle _ref` into oracle/_ref/.
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <assert.h>
#include <cmath>
#include <vector>
#include <algorithm>
#include <stdexcept>
#include <initializer_list>
#include <limits.h>
#include <float.h>
#include <string>
#include <sstream>
#include <fstream>
#include <iostream>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) (((depth) & 7) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8U 0
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32F_ CV_32F
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

// cvRound / cvFloor / cvCeil (core/fast_math.hpp): round half to even under the default rounding mode.
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
    T dot(const Point_& o) const { return x * o.x + y * o.y; }
};
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
typedef Point2i Point;
template <typename T> static inline Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> static inline Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x - b.x, a.y - b.y); }
// core/types.hpp: the scalar is applied per coordinate in its own type, the result cast back (saturate_cast is a plain cast for float)
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, int b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, float b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, double b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
template <typename T> static inline bool operator==(const Point_<T>& a, const Point_<T>& b) { return a.x == b.x && a.y == b.y; }

template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

// features2d KeyPoint: the field order every caller relies on (28 bytes).
class KeyPoint {
public:
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f p, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(p), size(s), angle(a), response(r), octave(o), class_id(c) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
};

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };

struct MatStep {
    size_t p[2];
    MatStep() { p[0] = p[1] = 0; }
    operator size_t() const { return p[0]; }
    size_t operator[](int i) const { return p[i]; }
};

// Mat::zeros / Mat::eye return an expression in OpenCV; ASSIGNING one to an existing Mat of the same size and type
// fills that Mat's buffer in place instead of rebinding it.  The reference relies on this at ORBextractor.cc:1037
// (`descriptors = Mat::zeros(...)` on a row range of the output matrix).
struct MatExpr { int rows, cols, type; bool eye; float fill; };     // fill: value of every element (Mat::ones, scaled), 0 for zeros / eye

// Dense 2-D single-channel matrix with OpenCV's sharing semantics: copies and ROIs alias one
// reference-counted buffer; clone() is deep.
class Mat {
public:
    int flags, rows, cols;
    uchar* data;
    const uchar* datastart;
    const uchar* dataend;
    MatStep step;

    Mat() : flags(0), rows(0), cols(0), data(0), datastart(0), dataend(0), buf(0) {}
    Mat(const Mat& m) : flags(m.flags), rows(m.rows), cols(m.cols), data(m.data), datastart(m.datastart), dataend(m.dataend),
                        step(m.step), buf(m.buf) { if (buf) __atomic_add_fetch(&buf->refs, 1, __ATOMIC_RELAXED); }
    Mat& operator=(const Mat& m)
    {
        if (this == &m) return *this;
        if (m.buf) __atomic_add_fetch(&m.buf->refs, 1, __ATOMIC_RELAXED);
        unref();
        flags = m.flags; rows = m.rows; cols = m.cols; data = m.data; datastart = m.datastart; dataend = m.dataend; step = m.step; buf = m.buf;
        return *this;
    }
    ~Mat() { unref(); }
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
    // user-allocated data (no ownership)
    Mat(int r, int c, int type, void* ptr, size_t stp = 0) : Mat()
    {
        flags = type; rows = r; cols = c; data = (uchar*)ptr;
        step.p[1] = elemSize(); step.p[0] = stp ? stp : (size_t)c * elemSize();
        datastart = data; dataend = data + (size_t)r * step.p[0];
    }
    Mat(const Mat& m, const Rect& roi) : Mat(m)
    {
        data = m.data + (size_t)roi.y * m.step.p[0] + (size_t)roi.x * m.elemSize();
        rows = roi.height; cols = roi.width;
    }

    void create(int r, int c, int type)
    {
        if (data && rows == r && cols == c && this->type() == type) return;
        flags = type; rows = r; cols = c;
        step.p[1] = elemSize(); step.p[0] = (size_t)c * elemSize();
        const size_t bytes = (size_t)r * step.p[0];
        // plain malloc, not operator new: pixel storage stays out of the way of the node allocator that
        // ref_api.cpp makes switchable (see there)
        unref();
        Buffer* b = (Buffer*)malloc(sizeof(Buffer) + bytes + 64);
        if (!b) throw std::bad_alloc();
        b->refs = 1;
        buf = b;
        uchar* p = (uchar*)(((uintptr_t)(b + 1) + 63) & ~(uintptr_t)63);
        data = p; datastart = p; dataend = p + bytes;
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { unref(); data = 0; datastart = dataend = 0; rows = cols = 0; }

    int type() const { return flags & 0xfff; }
    int depth() const { return flags & 7; }
    int channels() const { return 1; }
    size_t elemSize() const { static const int sz[8] = {1, 1, 2, 2, 4, 4, 8, 2}; return sz[depth()]; }
    size_t elemSize1() const { return elemSize(); }
    size_t step1(int i = 0) const { return step.p[i] / elemSize1(); }
    bool empty() const { return data == 0 || rows == 0 || cols == 0; }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step.p[0] == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }

    Mat operator()(const Rect& roi) const { return Mat(*this, roi); }
    Mat rowRange(int a, int b) const { return Mat(*this, Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return Mat(*this, Rect(a, 0, b - a, rows)); }
    Mat row(int i) const { return rowRange(i, i + 1); }
    Mat col(int j) const { return colRange(j, j + 1); }

    Mat clone() const
    {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type());
        for (int y = 0; y < rows; y++) memcpy(m.data + (size_t)y * m.step.p[0], data + (size_t)y * step.p[0], (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat& dst) const
    {
        if (empty()) { dst.release(); return; }
        dst.create(rows, cols, type());
        for (int y = 0; y < rows; y++) memmove(dst.data + (size_t)y * dst.step.p[0], data + (size_t)y * step.p[0], (size_t)cols * elemSize());
    }
    // into an existing view of the same size and type (OpenCV: copyTo(OutputArray) on a temporary header, Frame.cc:279-280)
    void copyTo(const Mat& dst) const
    {
        if (dst.rows != rows || dst.cols != cols || dst.type() != type()) throw std::runtime_error("cv shim: copyTo into a view of another size");
        for (int y = 0; y < rows; y++) memmove(dst.data + (size_t)y * dst.step.p[0], data + (size_t)y * step.p[0], (size_t)cols * elemSize());
    }
    // The shim's Mat has one channel: an N x 2 CV_32F matrix and its N x 1 two-channel view (Frame.cc:436, :438) are the same bytes,
    // and undistortPoints below reads either as N points.
    Mat reshape(int) const { return *this; }
    static MatExpr zeros(int r, int c, int type) { MatExpr e = {r, c, type, false, 0.f}; return e; }
    static MatExpr eye(int r, int c, int type) { MatExpr e = {r, c, type, true, 0.f}; return e; }
    static MatExpr ones(int r, int c, int type) { MatExpr e = {r, c, type, false, 1.f}; return e; }
    Mat(const MatExpr& e) : Mat() { *this = e; }
    Mat& operator=(const MatExpr& e)
    {
        create(e.rows, e.cols, e.type);         // keeps the buffer when size and type already match
        for (int y = 0; y < rows; y++) memset(data + (size_t)y * step.p[0], 0, (size_t)cols * elemSize());
        if (e.fill != 0.f) {
            if (depth() != CV_32F) throw std::runtime_error("cv shim: Mat::ones is CV_32F only");
            for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) at<float>(y, x) = e.fill;
        }
        if (e.eye)
            for (int i = 0; i < std::min(rows, cols); i++) {
                if (depth() == CV_32F) at<float>(i, i) = 1.f; else if (depth() == CV_64F) at<double>(i, i) = 1.0; else at<uchar>(i, i) = 1;
            }
        return *this;
    }

    uchar* ptr(int y = 0) { return data + (size_t)y * step.p[0]; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step.p[0]; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.p[0]); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.p[0]); }
    template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.p[0]))[x]; }
    template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.p[0]))[x]; }
    // single index: element i of a row or column vector (OpenCV's 1-D accessor)
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : (cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols)); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : (cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols)); }

    // position of this view inside the buffer it was cut from (copyMakeBorder without BORDER_ISOLATED looks there)
    void locateROI(Size& whole, Point& ofs) const
    {
        const size_t es = elemSize();
        const ptrdiff_t d0 = data - datastart, d1 = dataend - datastart;
        if (d0 == 0) { ofs.x = ofs.y = 0; }
        else { ofs.y = (int)(d0 / (ptrdiff_t)step.p[0]); ofs.x = (int)((d0 - (ptrdiff_t)step.p[0] * ofs.y) / (ptrdiff_t)es); }
        whole.height = std::max((int)((d1 - (ptrdiff_t)es * cols) / (ptrdiff_t)step.p[0] + 1), ofs.y + rows);
        whole.width = std::max((int)((d1 - (ptrdiff_t)step.p[0] * (whole.height - 1)) / (ptrdiff_t)es), ofs.x + cols);
    }

    // float matrix algebra used by the reference's ORBmatcher.cc (CV_32F only); defined in core_ops.hpp
    Mat t() const;
    Mat inv() const;
    double dot(const Mat& o) const;
    Mat mul(const Mat& o) const;
    void convertTo(Mat& dst, int rtype) const;

private:
    struct Buffer { int refs; };
    Buffer* buf;
    void unref() { if (buf && __atomic_sub_fetch(&buf->refs, 1, __ATOMIC_ACQ_REL) == 0) free(buf); buf = 0; }
};

template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, depthOf((T*)0)) {}
    Mat_(const Mat& m) : Mat(m) {}
    T& operator()(int y, int x) { return this->template at<T>(y, x); }
    const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
    // Mat_<float>(3,1) << a, b, c;
    struct Filler {
        Mat_* m; int i;
        Filler& operator,(T v) { m->template at<T>(i / m->cols, i % m->cols) = v; i++; return *this; }
        operator Mat_() const { return *m; }
        operator Mat() const { return *m; }
    };
    Filler operator<<(T v) { Filler f{this, 0}; f, v; return f; }
private:
    static int depthOf(float*) { return CV_32F; }
    static int depthOf(double*) { return CV_64F; }
    static int depthOf(uchar*) { return CV_8U; }
    static int depthOf(int*) { return CV_32S; }
};

class _InputArray {
public:
    _InputArray() : m(0) {}
    _InputArray(const Mat& mat) : m(&mat) {}
    Mat getMat(int = -1) const { return m ? *m : Mat(); }
    bool empty() const { return !m || m->empty(); }
protected:
    const Mat* m;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() : w(0) {}
    _OutputArray(Mat& mat) : _InputArray(mat), w(&mat) {}
    void create(int r, int c, int type) const { w->create(r, c, type); }
    void create(Size sz, int type) const { w->create(sz.height, sz.width, type); }
    void release() const { if (w) w->release(); }
    Mat getMat(int = -1) const { return w ? *w : Mat(); }
    Mat& getMatRef() const { return *w; }
private:
    Mat* w;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
static inline InputArray noArray() { static _InputArray none; return none; }

float fastAtan2(float y, float x);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);

// cv::undistortPoints(src, dst, K, dist, R = noArray(), P = K): oracle/orb_primitives.cpp's restatement (pinned to cv2 by
// tests/test_oracle_vs_cv2.py); src / dst: N x 2 CV_32F (see Mat::reshape), K 3 x 3 CV_32F, dist 4 or 5 x 1 (or 1 x 4 / 5) CV_32F
void undistortPoints(InputArray src, OutputArray dst, InputArray cameraMatrix, InputArray distCoeffs, InputArray R = noArray(), InputArray P = noArray());

// YAML / XML storage: not provided.  DBoW2's TemplatedVocabulary declares virtual save / load over it; this build loads vocabularies
// with loadFromTextFile only, so every operation here throws.
class FileNode {
public:
    FileNode operator[](const std::string&) const { return fail(); }
    FileNode operator[](const char*) const { return fail(); }
    FileNode operator[](int) const { return fail(); }
    size_t size() const { fail(); return 0; }
    operator int() const { fail(); return 0; }
    operator float() const { fail(); return 0; }
    operator double() const { fail(); return 0; }
    operator std::string() const { fail(); return std::string(); }
private:
    static FileNode fail() { throw std::runtime_error("cv shim: FileStorage not provided"); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const std::string&) const { throw std::runtime_error("cv shim: FileStorage not provided"); }
    FileNode operator[](const char*) const { throw std::runtime_error("cv shim: FileStorage not provided"); }
};
template <typename T> static inline FileStorage& operator<<(FileStorage&, const T&) { throw std::runtime_error("cv shim: FileStorage not provided"); }

class KeyPointsFilter {
public:
    static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};

} // namespace cv

#include "core_ops.hpp"
