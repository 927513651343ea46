// cv_shim.cpp -- CV SHIM FOR THE REFERENCE BUILD (test infrastructure, not product).
// The OpenCV functions the unmodified reference sources call, forwarded to the cv2-pinned scalar
// restatements of oracle/orb_primitives.cpp (tests/test_oracle_vs_cv2.py checks those bit for bit
// against cv2 4.13.0).  Call sites in the reference: FAST ORBextractor.cc:809,814; resize :1120;
// copyMakeBorder :1122,1127; GaussianBlur :1086; fastAtan2 :103; KeyPointsFilter::retainBest :1006,1024
// (dead code, must link).
#include <opencv2/core/core.hpp>
#include "../orb_internal.h"

namespace cv {

float fastAtan2(float y, float x) { return orbo::fast_atan2(y, x); }

static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType)
{
    Mat src = src_.getMat();
    if (src.depth() != CV_8U) throw std::runtime_error("cv shim: copyMakeBorder is 8-bit only");
    if ((borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101) throw std::runtime_error("cv shim: BORDER_REFLECT_101 only");
    const uchar* sdata = src.data;
    int rows = src.rows, cols = src.cols;
    if (!(borderType & BORDER_ISOLATED)) {
        // a view into a larger buffer: OpenCV takes the true neighbours where they exist
        Size whole; Point ofs;
        src.locateROI(whole, ofs);
        const int dtop = std::min(ofs.y, top), dbottom = std::min(whole.height - rows - ofs.y, bottom);
        const int dleft = std::min(ofs.x, left), dright = std::min(whole.width - cols - ofs.x, right);
        sdata -= (size_t)dtop * src.step.p[0] + dleft;
        rows += dtop + dbottom; cols += dleft + dright;
        top -= dtop; bottom -= dbottom; left -= dleft; right -= dright;
    }
    dst_.create(rows + top + bottom, cols + left + right, CV_8U);
    Mat dst = dst_.getMat();
    if (top == bottom && top == left && top == right) {
        orbo::border_reflect101(sdata, cols, rows, src.step.p[0], dst.data, top, dst.step.p[0]);
        return;
    }
    for (int y = -top; y < rows + bottom; y++) {
        const uchar* S = sdata + (size_t)reflect101(y, rows) * src.step.p[0];
        uchar* D = dst.data + (size_t)(y + top) * dst.step.p[0];
        for (int x = -left; x < cols + right; x++) D[x + left] = S[reflect101(x, cols)];
    }
}

void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation)
{
    Mat src = src_.getMat();
    if (src.depth() != CV_8U || interpolation != INTER_LINEAR) throw std::runtime_error("cv shim: resize is 8-bit INTER_LINEAR only");
    if (dsize.width == 0 || dsize.height == 0) dsize = Size(cvRound(src.cols * fx), cvRound(src.rows * fy));
    dst_.create(dsize.height, dsize.width, CV_8U);
    Mat dst = dst_.getMat();
    orbo::resize_linear_u8(src.data, src.cols, src.rows, src.step.p[0], dst.data, dst.cols, dst.rows, dst.step.p[0]);
}

void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType)
{
    Mat src = src_.getMat();
    if (src.depth() != CV_8U || ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || (sigmaY != 2 && sigmaY != 0) ||
        (borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101)
        throw std::runtime_error("cv shim: GaussianBlur is 8-bit 7x7 sigma 2 REFLECT_101 only");
    dst_.create(src.rows, src.cols, CV_8U);
    Mat dst = dst_.getMat();
    orbo::gauss7(src.data, src.cols, src.rows, src.step.p[0], dst.data, dst.step.p[0]);   // reads src fully before writing dst
}

void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression)
{
    Mat img = image.getMat();
    if (img.depth() != CV_8U) throw std::runtime_error("cv shim: FAST is 8-bit only");
    std::vector<orbo::XYR> found;
    orbo::fast_detect(img.data, img.cols, img.rows, img.step.p[0], threshold, nonmaxSuppression, found);
    keypoints.clear();
    keypoints.reserve(found.size());
    for (const orbo::XYR& k : found) keypoints.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1.f, (float)k.r));
}

// features2d KeyPointsFilter::retainBest: the npoints strongest by response, plus every keypoint that ties with the last kept.
void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int npoints)
{
    if (npoints < 0 || (int)keypoints.size() <= npoints) return;
    if (npoints == 0) { keypoints.clear(); return; }
    std::nth_element(keypoints.begin(), keypoints.begin() + npoints - 1, keypoints.end(),
                     [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
    const float ambiguous = keypoints[npoints - 1].response;
    std::vector<KeyPoint>::iterator end = std::partition(keypoints.begin() + npoints, keypoints.end(),
                                                         [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
    keypoints.resize(end - keypoints.begin());
}

} // namespace cv
