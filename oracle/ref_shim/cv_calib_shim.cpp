// cv_calib_shim.cpp -- CV SHIM FOR THE REFERENCE BUILD (test infrastructure, not product): cv::undistortPoints as the reference's Frame.cc
// calls it (:437, :463: undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on an N x 2 CV_32F matrix), forwarded to the restatement in
// oracle/orb_matcher.cpp that tests/test_oracle_vs_cv2.py pins bit for bit to cv2 4.13.0.  That function lives in liborb_oracle.so, which
// carries its own static libstdc++; it is therefore reached with dlopen(RTLD_LOCAL) + dlsym -- linking against the library would let
// its copies of the iostream symbols pre-empt the system's inside this library.
#include <opencv2/core/core.hpp>
#include <dlfcn.h>
#include <string>

namespace {
typedef void (*undistort_fn)(const float* pts, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* out);
undistort_fn oracle_undistort_points()
{
    static undistort_fn fn = nullptr;
    if (!fn) {
        Dl_info info;
        if (!dladdr((void*)&oracle_undistort_points, &info) || !info.dli_fname) throw std::runtime_error("cv shim: dladdr failed");
        std::string dir(info.dli_fname);
        dir = dir.substr(0, dir.find_last_of('/'));                    // .../oracle/_ref
        const std::string path = dir + "/../liborb_oracle.so";
        void* h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!h) throw std::runtime_error("cv shim: cannot load " + path);
        fn = (undistort_fn)dlsym(h, "orbo_undistort_points");
        if (!fn) throw std::runtime_error("cv shim: orbo_undistort_points not found");
    }
    return fn;
}
}  // namespace

namespace cv {

void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray dist_, InputArray R, InputArray P)
{
    const Mat src = src_.getMat(), K = K_.getMat(), dist = dist_.getMat();
    if (src.depth() != CV_32F || src.cols != 2 || !src.isContinuous()) throw std::runtime_error("cv shim: undistortPoints wants N x 2 CV_32F");
    if (!R.empty()) throw std::runtime_error("cv shim: undistortPoints with a rectification is not provided");
    const Mat Pm = P.empty() ? K : P.getMat();
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++)
            if (Pm.at<float>(r, c) != K.at<float>(r, c)) throw std::runtime_error("cv shim: undistortPoints is provided for P == K only");
    const int nd = (int)dist.total();
    float d[12] = {0};
    for (int i = 0; i < nd && i < 12; i++) d[i] = dist.at<float>(i);
    std::vector<float> out(2 * (size_t)src.rows);
    oracle_undistort_points()(src.ptr<float>(0), src.rows, K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2), d, nd, out.data());
    Mat dst(src.rows, 2, CV_32F);                       // src and dst are the same Mat at both call sites: write a fresh buffer, then rebind
    memcpy(dst.data, out.data(), out.size() * sizeof(float));
    dst_.getMatRef() = dst;
}

} // namespace cv
