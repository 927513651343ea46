// matcher_ref_api.cpp -- TEST INFRASTRUCTURE, not product: the C++ entry points of oracle/_ref/libmatcher_ref.so (slam/matcher_ref.h).
// Compiled with -DORB_SLAM2=ORB_SLAM2_REF next to the reference's own src/ORBmatcher.cc; this is the only file that sees the reference's
// include/ORBmatcher.h.
#include "ORBmatcher.h"        // oracle/ref_shim/slam/ORBmatcher.h -> stand-in objects, then the reference's header
#undef ORB_SLAM2
#include "matcher_ref.h"

// the static members of the stand-in Frame (static in the reference too: Frame.h:124-131, :189-192) live here
namespace ORB_SLAM2_REF {
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMaxX = 0, Frame::mnMinY = 0, Frame::mnMaxY = 0;
}

namespace refm {
typedef ORB_SLAM2_REF::ORBmatcher RefMatcher;
int SearchByProjection(float nn, bool co, Frame& F, const std::vector<MapPoint*>& v, float th) { RefMatcher m(nn, co); return m.SearchByProjection(F, v, th); }
int SearchByProjection(float nn, bool co, Frame& C, const Frame& L, float th, bool bMono) { RefMatcher m(nn, co); return m.SearchByProjection(C, L, th, bMono); }
int SearchByProjection(float nn, bool co, Frame& C, KeyFrame* pKF, const std::set<MapPoint*>& found, float th, int ORBdist)
{
    RefMatcher m(nn, co);
    return m.SearchByProjection(C, pKF, found, th, ORBdist);
}
int SearchByProjection(float nn, bool co, KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vp, std::vector<MapPoint*>& vpMatched, int th)
{
    RefMatcher m(nn, co);
    return m.SearchByProjection(pKF, Scw, vp, vpMatched, th);
}
int SearchByBoW(float nn, bool co, KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& out) { RefMatcher m(nn, co); return m.SearchByBoW(pKF, F, out); }
int SearchByBoW(float nn, bool co, KeyFrame* a, KeyFrame* b, std::vector<MapPoint*>& out) { RefMatcher m(nn, co); return m.SearchByBoW(a, b, out); }
int SearchForInitialization(float nn, bool co, Frame& F1, Frame& F2, std::vector<cv::Point2f>& prev, std::vector<int>& m12, int windowSize)
{
    RefMatcher m(nn, co);
    return m.SearchForInitialization(F1, F2, prev, m12, windowSize);
}
int SearchForTriangulation(float nn, bool co, KeyFrame* a, KeyFrame* b, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& pairs, bool bOnlyStereo)
{
    RefMatcher m(nn, co);
    return m.SearchForTriangulation(a, b, F12, pairs, bOnlyStereo);
}
int SearchBySim3(float nn, bool co, KeyFrame* a, KeyFrame* b, std::vector<MapPoint*>& m12, float s12, const cv::Mat& R12, const cv::Mat& t12, float th)
{
    RefMatcher m(nn, co);
    return m.SearchBySim3(a, b, m12, s12, R12, t12, th);
}
int Fuse(float nn, bool co, KeyFrame* pKF, const std::vector<MapPoint*>& v, float th) { RefMatcher m(nn, co); return m.Fuse(pKF, v, th); }
int Fuse(float nn, bool co, KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& v, float th, std::vector<MapPoint*>& rep)
{
    RefMatcher m(nn, co);
    return m.Fuse(pKF, Scw, v, th, rep);
}
int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return RefMatcher::DescriptorDistance(a, b); }
}  // namespace refm
