// prelude.h -- force-included (-include) in the build of oracle/_ref/libslam_ref.so (test infrastructure, not product).
//
// That library is the reference's object graph from its own, unmodified sources: src/Frame.cc, src/KeyFrame.cc, src/MapPoint.cc,
// src/ORBmatcher.cc, src/ORBextractor.cc, their headers, and the vendored DBoW2.  What those files point to beyond the path -- Map,
// KeyFrameDatabase, the ServerMap classes (ROS messages), boost::serialization, Converter (Eigen, g2o) -- is not in this image; this file
// defines their include guards, so that the real headers become empty, and stands in the few members the five source files name.
#pragma once
#define MAP_H
#define KEYFRAMEDATABASE_H
#define SERVERMAP_H
#define BOOST_ARCHIVER_H
#define CONVERTER_H
#include <limits.h>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <vector>
#include <opencv2/core/core.hpp>
#include <Thirdparty/DBoW2/DBoW2/BowVector.h>
#include <Thirdparty/DBoW2/DBoW2/FeatureVector.h>

namespace boost
{
namespace serialization { class access; }
namespace archive
{
// KeyFrame::serialize and MapPoint::serialize are instantiated for both (KeyFrame.cc:775-776, MapPoint.cc:477-478); nothing is stored
class binary_iarchive { public: template <class T> binary_iarchive& operator&(T&) { return *this; } template <class T> binary_iarchive& operator&(const T&) { return *this; } };
class binary_oarchive { public: template <class T> binary_oarchive& operator&(T&) { return *this; } template <class T> binary_oarchive& operator&(const T&) { return *this; } };
}
}  // namespace boost

namespace ORB_SLAM2
{
class KeyFrame;
class MapPoint;

class Map
{
public:
    std::mutex mMutexPointCreation, mMutexMapUpdate;
    std::vector<MapPoint*> GetAllMapPoints() { return std::vector<MapPoint*>(points.begin(), points.end()); }
    void AddMapPoint(MapPoint* p) { points.insert(p); }
    void EraseMapPoint(MapPoint* p) { points.erase(p); }
    void UpdateMapPoint(MapPoint*) {}
    void AddKeyFrame(KeyFrame* k) { keyframes.insert(k); }
    void EraseKeyFrame(KeyFrame* k) { keyframes.erase(k); }
    void UpdateKeyFrame(KeyFrame*) {}
    std::set<MapPoint*> points;
    std::set<KeyFrame*> keyframes;
};

class KeyFrameDatabase
{
public:
    void erase(KeyFrame*) {}
};

class ServerMapPoint
{
public:
    ServerMapPoint() : mnId(0), nObs(0), mnFirstKFid(0), mnFirstFrame(0), uid(0) {}
    long unsigned int mnId;
    int nObs;
    long int mnFirstKFid, mnFirstFrame;
    cv::Mat mWorldPos, mDescriptor;
    unsigned int GetUID() { return uid; }
    unsigned int uid;
};

class ServerKeyFrame
{
public:
    ServerKeyFrame() : mnId(0), mnFrameId(0), mTimeStamp(0), mfGridElementWidthInv(0), mfGridElementHeightInv(0), mnTrackReferenceForFrame(0),
                       mnFuseTargetForKF(0), mnBALocalForKF(0), mnBAFixedForKF(0), mnLoopQuery(0), mnLoopWords(0), mnRelocQuery(0), mnRelocWords(0) {}
    long unsigned int mnId, mnFrameId;
    double mTimeStamp;
    float mfGridElementWidthInv, mfGridElementHeightInv;
    long unsigned int mnTrackReferenceForFrame, mnFuseTargetForKF, mnBALocalForKF, mnBAFixedForKF, mnLoopQuery;
    int mnLoopWords;
    long unsigned int mnRelocQuery;
    int mnRelocWords;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
};

// Converter.cc:29-37, the one conversion Frame.cc:414 and KeyFrame.cc:63 call: one row view per descriptor
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

using namespace std;      // the reference's headers rely on a using-directive that one of the headers blanked above leaks
