// Stand-in for the reference's include/Converter.h in the libframe_ref.so build (test infrastructure): the real header pulls in Eigen and g2o
// for conversions Frame.cc never calls.  Frame.cc:414 uses toDescriptorVector only (Converter.cc:29-37: one row view per descriptor).
#ifndef CONVERTER_H
#define CONVERTER_H
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2
{
class Converter
{
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors)
    {
        std::vector<cv::Mat> v;
        v.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) v.push_back(Descriptors.row(j));
        return v;
    }
};
}  // namespace ORB_SLAM2
#endif
