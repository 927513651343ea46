// ORBextractor.h -- drop-in replacement of the reference's include/ORBextractor.h (kimwin2/ORB_SLAM2v2-1
// include/ORBextractor.h:45-111, src/ORBextractor.cc): same namespace, class name, constructor, operator(), accessors
// and public mvImagePyramid, so Frame.cc / Tracking.cc compile unchanged.  Every method forwards to the C ABI of
// liborbx.so (include/orbx.h), whose kernels are hand-written sm_100a CUDA.  There is no CPU fallback: a CUDA failure
// throws std::runtime_error (the reference has no error path of its own; OpenCV would throw cv::Exception).
//
// Link with -lorbx; leave the reference's src/ORBextractor.cc out of the build.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "orbx.h"
#include "orbx_cv_shim.h"

namespace ORB_SLAM2
{

class ORBextractor
{
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // reference: ORBextractor.cc:410-470.  device: CUDA ordinal (extra, defaulted argument)
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
        : mvImagePyramid(nlevels), h_(nullptr), nlevels_(nlevels), scaleFactor_(scaleFactor)
    {
        if (orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, 1, &h_) != ORBX_OK) fail("orbx_create");
        if (orbx_set_pyramid_mirror(h_, 1) != ORBX_OK) fail("orbx_set_pyramid_mirror");
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        orbx_scale_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data());
    }
    ~ORBextractor() { orbx_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image.  Mask is ignored, as in the reference (ORBextractor.cc:1043).
    void operator()(cv::InputArray image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors)
    {
        if (image.empty()) return;                                    // :1046-1047
        cv::Mat im = image.getMat();
        run(im.ptr<uint8_t>(0), im.cols, im.rows, im.step, 1, false, keypoints, descriptors);
    }

    // The same on a colour image: cvtColor(im, im, CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of Tracking::GrabImage*
    // (Tracking.cc:280-305, :320-334, :350-364) runs on the device in front of the pyramid; channels = 3 or 4, bRGB = mbRGB.
    // mvImagePyramid[0] is then the grey image the reference would have produced.
    void ExtractColor(const uint8_t* data, int cols, int rows, size_t step, int channels, bool bRGB, std::vector<cv::KeyPoint>& keypoints,
                      cv::OutputArray descriptors)
    {
        if (!data || cols <= 0 || rows <= 0) return;
        run(data, cols, rows, step, channels, bRGB, keypoints, descriptors);
    }

private:
    void run(const uint8_t* data, int cols, int rows, size_t step, int channels, bool bRGB, std::vector<cv::KeyPoint>& keypoints,
             cv::OutputArray descriptors)
    {
        int cap = orbx_max_keypoints(h_), n = 0;
        for (int attempt = 0; attempt < 2; attempt++) {
            kbuf_.resize(cap);
            dbuf_.resize((size_t)cap * 32);
            const uint8_t* one[1] = {data};
            const int rc = channels == 1 ? orbx_extract(h_, data, cols, rows, step, kbuf_.data(), dbuf_.data(), cap, &n)
                                         : orbx_extract_color_batch(h_, one, 1, cols, rows, step, channels, bRGB, kbuf_.data(), dbuf_.data(), cap, &n);
            if (rc == ORBX_OK) break;
            if (rc != ORBX_ERR_CAPACITY || attempt == 1) fail("orbx_extract");
            cap = n;
        }
        keypoints.clear();
        keypoints.reserve(n);
        if (n == 0) descriptors.release();                             // :1064-1065
        else descriptors.create(n, 32, CV_8U);                        // :1068
        for (int i = 0; i < n; i++) {
            cv::KeyPoint kp;
            kp.pt.x = kbuf_[i].x; kp.pt.y = kbuf_[i].y; kp.size = kbuf_[i].size; kp.angle = kbuf_[i].angle;
            kp.response = kbuf_[i].response; kp.octave = kbuf_[i].octave; kp.class_id = kbuf_[i].class_id;
            keypoints.push_back(kp);
        }
        if (n) {
            cv::Mat d = descriptors.getMat();
            for (int i = 0; i < n; i++) memcpy(d.ptr<uint8_t>(i), &dbuf_[(size_t)i * 32], 32);
        }
        // mvImagePyramid (include/ORBextractor.h:85 of the reference; read by Frame::ComputeStereoMatches, Frame.cc:488,578,590,595):
        // views of the handle's page-locked mirror, which the call above filled beside its kernels -- no copy, no launch and no
        // synchronisation here.  Like the reference's buffers (re-allocated per call, :1114-1115) they hold the LAST frame only.
        if (downloadPyramid_)
            for (int l = 0; l < nlevels_; l++) {
                int w, hh;
                const uint8_t* p = nullptr;
                size_t pitch = 0;
                if (orbx_level_size(h_, l, &w, &hh) != ORBX_OK || orbx_pyramid_mirror(h_, l, &p, &pitch) != ORBX_OK) fail("orbx_pyramid_mirror");
                mvImagePyramid[l] = cv::Mat(hh, w, CV_8U, const_cast<uint8_t*>(p), pitch);
            }
    }

public:

    int inline GetLevels() { return nlevels_; }
    float inline GetScaleFactor() { return scaleFactor_; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;

    // extras of the replacement
    // off: mvImagePyramid stays empty and the call skips the 1 MB device-to-host copy (callers that associate stereo on the
    // device with ORBmatcher::ComputeStereoMatches do not read it)
    void setPyramidDownload(bool on)
    {
        downloadPyramid_ = on;
        if (orbx_set_pyramid_mirror(h_, on ? 1 : 0) != ORBX_OK) fail("orbx_set_pyramid_mirror");
        if (!on) for (size_t l = 0; l < mvImagePyramid.size(); l++) mvImagePyramid[l] = cv::Mat();
    }
    orbx_extractor* handle() { return h_; }

protected:
    void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + orbx_last_error()); }

    orbx_extractor* h_;
    int nlevels_;
    float scaleFactor_;
    bool downloadPyramid_ = true;
    std::vector<orbx_kp> kbuf_;
    std::vector<uint8_t> dbuf_;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
};

}  // namespace ORB_SLAM2

#endif
