// Wrapper found first on the include path when the reference's sources say #include "MapPoint.h" in the build of
// oracle/_ref/libmappoint_ref.so (test infrastructure, not product): MapPoint is the reference's own class (include/MapPoint.h,
// src/MapPoint.cc, unmodified); what it points to is stood in.  KeyFrame and Frame come from ref_shim/slam/slam_objects.h; Map, ServerMapPoint
// and the boost archives are the few members and operators MapPoint.cc names (Map.h / ServerMap.h drag in ROS message headers,
// BoostArchiver.h boost::serialization, none of which this image has).
#ifndef ORBX_MAPPOINT_WRAPPER_H
#define ORBX_MAPPOINT_WRAPPER_H
#define ORBX_REAL_MAPPOINT
#define KEYFRAME_H
#define FRAME_H
#define MAP_H
#define SERVERMAP_H
#define BOOST_ARCHIVER_H
#include <limits.h>
#include <map>
#include <mutex>
#include <set>
#include <vector>
#include <opencv2/core/core.hpp>

namespace boost
{
namespace serialization { class access; }
namespace archive
{
// MapPoint::serialize is instantiated for both (MapPoint.cc:477-478); nothing here stores anything
class binary_iarchive { public: template <class T> binary_iarchive& operator&(T&) { return *this; } };
class binary_oarchive { public: template <class T> binary_oarchive& operator&(T&) { return *this; } };
}
}  // namespace boost

namespace ORB_SLAM2
{
class MapPoint;
class Map
{
public:
    std::mutex mMutexPointCreation;
    std::vector<MapPoint*> GetAllMapPoints() { return std::vector<MapPoint*>(points.begin(), points.end()); }
    void EraseMapPoint(MapPoint* p) { points.erase(p); erased.push_back(p); }
    void UpdateMapPoint(MapPoint* p) { updated.push_back(p); }
    std::set<MapPoint*> points;
    std::vector<MapPoint*> erased, updated;
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
}  // namespace ORB_SLAM2

using namespace std;      // the reference's headers rely on a using-directive that one of the real object headers leaks
#include_next "MapPoint.h"
#include "../slam/slam_objects.h"
#endif
