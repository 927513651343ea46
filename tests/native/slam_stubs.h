// slam_stubs.h -- stand-ins for the reference's Frame / KeyFrame / MapPoint (include/Frame.h, KeyFrame.h, MapPoint.h of
// kimwin2/ORB_SLAM2v2-1) carrying exactly the members that src/ORBmatcher.cc reads or writes, under the reference's names.
// Test infrastructure: lets the reference's call lines (Tracking.cc:742-743, :907-910, :1028, :1336, LoopClosing.cc:275) compile verbatim
// against include/ORBmatcher.h without the SLAM object graph (Map, ROS messages).
#pragma once
#include <map>
#include <vector>
#include "ORBmatcher.h"

namespace ORB_SLAM2
{

namespace DBoW2stub { typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector; }

class MapPoint
{
public:
    MapPoint() : mTrackProjX(0), mTrackProjY(0), mTrackProjXR(0), mbTrackInView(false), mnTrackScaleLevel(0), mTrackViewCos(1), bad(false), nObs(1) {}
    // Tracking variables (MapPoint.h:116-123)
    float mTrackProjX, mTrackProjY, mTrackProjXR;
    bool mbTrackInView;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    bool isBad() { return bad; }
    int Observations() { return nObs; }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }

    bool bad;
    int nObs;
    cv::Mat mDescriptor;      // 1 x 32 CV_8U
    cv::Mat mWorldPos;        // 3 x 1 CV_32F
};

class Frame
{
public:
    Frame() : N(0), fx(0), fy(0), cx(0), cy(0), mbf(0), mb(0) {}
    int N;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<float> mvScaleFactors;
    DBoW2stub::FeatureVector mFeatVec;
    cv::Mat mTcw;             // 4 x 4 CV_32F
    float fx, fy, cx, cy, mbf, mb;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
};
float Frame::mnMinX = 0, Frame::mnMaxX = 0, Frame::mnMinY = 0, Frame::mnMaxY = 0;

class KeyFrame
{
public:
    KeyFrame() : N(0), mnMinX(0), mnMinY(0), mnMaxX(0), mnMaxY(0) {}
    int N;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    DBoW2stub::FeatureVector mFeatVec;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::vector<MapPoint*> mvpMapPoints;
};

}  // namespace ORB_SLAM2
