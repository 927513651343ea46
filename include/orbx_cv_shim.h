// orbx_cv_shim.h -- the few OpenCV types the drop-in headers need, for builds WITHOUT OpenCV (this image has no OpenCV
// C++ headers).  Define ORBX_USE_OPENCV to compile include/ORBextractor.h / ORBmatcher.h against the real library
// instead (what a user of kimwin2/ORB_SLAM2v2-1 does); field order and sizes of cv::KeyPoint match either way
// (pt.x, pt.y, size, angle, response, octave, class_id = 28 bytes).
#ifndef ORBX_CV_SHIM_H_
#define ORBX_CV_SHIM_H_
#ifdef ORBX_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <memory>
#include <vector>

// macros, as in OpenCV's cvdef.h (user code writes CV_8U, never cv::CV_8U)
#ifndef CV_8U
#define CV_8U 0
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_8U
#define CV_32FC1 CV_32F
#endif

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point2i;
typedef Point2i Point;

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

// single-channel matrix (CV_8U descriptors and images, CV_32F poses): shared, reference-counted storage or a non-owning
// view with a row step in bytes
class Mat {
public:
    int rows, cols;
    size_t step;
    uint8_t* data;
    Mat() : rows(0), cols(0), step(0), data(nullptr), type_(CV_8U) {}
    Mat(int r, int c, int type) : rows(0), cols(0), step(0), data(nullptr), type_(CV_8U) { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t st = 0)
        : rows(r), cols(c), step(st ? st : (size_t)c * esz(type)), data((uint8_t*)ext), type_(type) {}
    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && own_ && type == type_ && step == (size_t)c * esz(type)) return;
        own_.reset(new std::vector<uint8_t>((size_t)r * c * esz(type)));
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        data = own_->empty() ? nullptr : own_->data();
    }
    void release() { own_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return esz(type_); }
    bool isContinuous() const { return step == (size_t)cols * esz(type_); }
    template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    template <typename T> T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T> const T& at(int r, int c) const { return ptr<T>(r)[c]; }
    Mat row(int r) const { Mat m(1, cols, type_, data + (size_t)r * step, step); m.own_ = own_; return m; }
    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; r++) memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * esz(type_));
        return m;
    }
    Mat getMat() const { return *this; }
private:
    static size_t esz(int type) { return (type & 7) == CV_64F ? 8 : ((type & 7) >= CV_32S ? 4 : 1); }
    std::shared_ptr<std::vector<uint8_t>> own_;
    int type_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif  // ORBX_USE_OPENCV
#endif  // ORBX_CV_SHIM_H_
