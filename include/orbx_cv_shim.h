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

namespace cv {

enum { CV_8U = 0, CV_8UC1 = 0 };

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

// 8-bit single-channel matrix: shared, reference-counted storage or a non-owning view with a row step
class Mat {
public:
    int rows, cols;
    size_t step;
    uint8_t* data;
    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int /*type*/) : rows(0), cols(0), step(0), data(nullptr) { create(r, c, CV_8U); }
    Mat(int r, int c, int /*type*/, void* ext, size_t st = 0) : rows(r), cols(c), step(st ? st : (size_t)c), data((uint8_t*)ext) {}
    void create(int r, int c, int /*type*/)
    {
        if (r == rows && c == cols && own_ && step == (size_t)c) return;
        own_.reset(new std::vector<uint8_t>((size_t)r * c));
        rows = r; cols = c; step = (size_t)c;
        data = own_->empty() ? nullptr : own_->data();
    }
    void release() { own_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    bool isContinuous() const { return step == (size_t)cols; }
    template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    Mat row(int r) const { Mat m(1, cols, CV_8U, data + (size_t)r * step, step); m.own_ = own_; return m; }
    Mat getMat() const { return *this; }
private:
    std::shared_ptr<std::vector<uint8_t>> own_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif  // ORBX_USE_OPENCV
#endif  // ORBX_CV_SHIM_H_
