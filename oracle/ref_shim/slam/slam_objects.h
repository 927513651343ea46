// slam_objects.h -- TEST INFRASTRUCTURE, not product.  Stand-ins for the reference's MapPoint / KeyFrame / Frame
// (kimwin2/ORB_SLAM2v2-1 include/MapPoint.h, KeyFrame.h, Frame.h) carrying exactly the members that the UNMODIFIED
// src/ORBmatcher.cc reads, writes or calls, under the reference's names, so that file compiles and runs without the SLAM object
// graph (Map, ServerMap, ROS messages, g2o).  They define the reference's include guards, so the real headers become empty when
// ORBmatcher.h pulls them in.  The handful of member functions restate the reference's (file:line given); the matching logic
// itself -- everything in ORBmatcher.cc -- stays the reference's own code.
//
// Everything lives in `namespace ORB_SLAM2`; the reference build and the tests include this file under
// `#define ORB_SLAM2 ORB_SLAM2_REF` so that the reference's ORBmatcher cannot collide with the product's class of the same name.
#ifndef ORBX_SLAM_OBJECTS_H
#define ORBX_SLAM_OBJECTS_H
// With ORBX_REAL_FRAME (the build of the reference's src/Frame.cc, ref_shim/frame/) only MapPoint and KeyFrame are stood in: Frame is the
// reference's own class and DBoW2's vendored headers are the real ones.  With ORBX_REAL_MAPPOINT (the build of src/MapPoint.cc,
// ref_shim/mappoint/) MapPoint is the reference's own class, already defined when this file is read, and KeyFrame / Frame are stood in.
#define MAPPOINT_H
#define KEYFRAME_H
#ifndef ORBX_REAL_FRAME
#define FRAME_H
#endif
#include <cmath>
#include <map>
#include <mutex>
#include <set>
#include <vector>
#include <opencv2/core/core.hpp>
#ifdef ORBX_REAL_FRAME
#include <Thirdparty/DBoW2/DBoW2/FeatureVector.h>      // found on the include path only: the reference's vendored DBoW2
#else
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#endif

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

namespace ORB_SLAM2
{

class KeyFrame;
class Frame;

#ifndef ORBX_REAL_MAPPOINT
class MapPoint
{
public:
    MapPoint() : mTrackProjX(0), mTrackProjY(0), mTrackProjXR(0), mbTrackInView(false), mnTrackScaleLevel(0), mTrackViewCos(1),
                 mnId(0), bad(false), nObs(1), mfMinDistance(0), mfMaxDistance(0), mpReplaced(0) {}
    // tracking variables (MapPoint.h:116-123)
    float mTrackProjX, mTrackProjY, mTrackProjXR;
    bool mbTrackInView;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    long unsigned int mnId;

    bool isBad() { return bad; }
    int Observations() { return nObs; }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }          // MapPoint.cc:385-389
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }          // MapPoint.cc:391-395
    inline int PredictScale(const float& currentDist, KeyFrame* pKF);          // MapPoint.cc:397-412
    inline int PredictScale(const float& currentDist, Frame* pF);              // MapPoint.cc:414-429
    bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
    int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
    void AddObservation(KeyFrame* pKF, size_t idx) { if (!mObservations.count(pKF)) { mObservations[pKF] = idx; nObs++; } }
    void Replace(MapPoint* pMP) { if (pMP != this) { mpReplaced = pMP; bad = true; } }

    // state the tests set directly
    bool bad;
    int nObs;
    float mfMinDistance, mfMaxDistance;
    cv::Mat mDescriptor;       // 1 x 32 CV_8U
    cv::Mat mWorldPos;         // 3 x 1 CV_32F
    cv::Mat mNormalVector;     // 3 x 1 CV_32F
    std::map<KeyFrame*, size_t> mObservations;
    MapPoint* mpReplaced;
};

#endif  // ORBX_REAL_MAPPOINT

// the 64 x 48 bucket grid over the undistorted keypoints: Frame::AssignFeaturesToGrid / PosInGrid (Frame.cc:230-245, :397-407),
// KeyFrame copies it (KeyFrame.cc:31-57)
struct FeatureGrid {
    float minX, minY, maxX, maxY, invW, invH;
    std::vector<size_t> cell[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    void build(const std::vector<cv::KeyPoint>& keysUn, float mnMinX, float mnMaxX, float mnMinY, float mnMaxY)
    {
        minX = mnMinX; maxX = mnMaxX; minY = mnMinY; maxY = mnMaxY;
        invW = (float)FRAME_GRID_COLS / (maxX - minX);                          // Frame.cc:101-102
        invH = (float)FRAME_GRID_ROWS / (maxY - minY);
        for (int i = 0; i < FRAME_GRID_COLS; i++) for (int j = 0; j < FRAME_GRID_ROWS; j++) cell[i][j].clear();
        for (size_t i = 0; i < keysUn.size(); i++) {
            const int px = (int)round((keysUn[i].pt.x - minX) * invW), py = (int)round((keysUn[i].pt.y - minY) * invH);
            if (px < 0 || px >= FRAME_GRID_COLS || py < 0 || py >= FRAME_GRID_ROWS) continue;
            cell[px][py].push_back(i);
        }
    }
    // Frame::GetFeaturesInArea (Frame.cc:342-395; with levels) and KeyFrame::GetFeaturesInArea (KeyFrame.cc:572-611; checkLevels false)
    std::vector<size_t> query(const std::vector<cv::KeyPoint>& keysUn, float x, float y, float r, bool useLevels, int minLevel, int maxLevel) const
    {
        std::vector<size_t> out;
        const int x0 = std::max(0, (int)floor((x - minX - r) * invW)), x1 = std::min(FRAME_GRID_COLS - 1, (int)ceil((x - minX + r) * invW));
        const int y0 = std::max(0, (int)floor((y - minY - r) * invH)), y1 = std::min(FRAME_GRID_ROWS - 1, (int)ceil((y - minY + r) * invH));
        if (x0 >= FRAME_GRID_COLS || x1 < 0 || y0 >= FRAME_GRID_ROWS || y1 < 0) return out;
        const bool check = useLevels && (minLevel > 0 || maxLevel >= 0);
        for (int ix = x0; ix <= x1; ix++)
            for (int iy = y0; iy <= y1; iy++)
                for (size_t j = 0; j < cell[ix][iy].size(); j++) {
                    const cv::KeyPoint& kp = keysUn[cell[ix][iy][j]];
                    if (check && (kp.octave < minLevel || (maxLevel >= 0 && kp.octave > maxLevel))) continue;
                    if (fabs(kp.pt.x - x) < r && fabs(kp.pt.y - y) < r) out.push_back(cell[ix][iy][j]);
                }
        return out;
    }
};

class KeyFrame
{
public:
    KeyFrame() : mnId(0), mnFrameId(0), bad(false), N(0), fx(0), fy(0), cx(0), cy(0), invfx(0), invfy(0), mbf(0), mb(0), mThDepth(0), mnScaleLevels(8), mfScaleFactor(1.2f),
                 mfLogScaleFactor(std::log(1.2f)), mnMinX(0), mnMinY(0), mnMaxX(0), mnMaxY(0) {}
    long unsigned int mnId, mnFrameId;
    bool bad;
    bool isBad() { return bad; }                                                // KeyFrame.cc:548-552
    void EraseMapPointMatch(const size_t& idx) { mvpMapPoints[idx] = static_cast<MapPoint*>(NULL); }          // KeyFrame.cc:217-221
    void ReplaceMapPointMatch(const size_t& idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }                  // KeyFrame.cc:231-234
    int N;
    float fx, fy, cx, cy, invfx, invfy, mbf, mb, mThDepth;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    int mnScaleLevels;
    float mfScaleFactor, mfLogScaleFactor;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;                                        // KeyFrame.h:146-149 (ints there)

    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::set<MapPoint*> GetMapPoints()                                          // KeyFrame.cc:236-249: the non-null, non-bad ones
    {
        std::set<MapPoint*> s;
        for (size_t i = 0; i < mvpMapPoints.size(); i++)
            if (mvpMapPoints[i] && !mvpMapPoints[i]->isBad()) s.insert(mvpMapPoints[i]);
        return s;
    }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
    cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }       // KeyFrame.cc:111-121
    cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }      // KeyFrame.cc:613-616
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const { return grid.query(mvKeysUn, x, y, r, false, -1, -1); }

    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Tcw, Ow;           // 4 x 4 and 3 x 1 CV_32F; SetPose (KeyFrame.cc:70-84) computes Ow = -Rcw.t() * tcw
    FeatureGrid grid;
};

#ifndef ORBX_REAL_FRAME
class Frame
{
public:
    Frame() : mnId(0), N(0), mbf(0), mb(0), mThDepth(0), mnScaleLevels(8), mfScaleFactor(1.2f), mfLogScaleFactor(std::log(1.2f)) {}
    long unsigned int mnId;
    cv::Mat GetCameraCenter() { return mOw.clone(); }                            // Frame.h:68-70
    cv::Mat mOw;
    int N;
    static float fx, fy, cx, cy, invfx, invfy;
    float mbf, mb, mThDepth;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    cv::Mat mTcw;              // 4 x 4 CV_32F
    int mnScaleLevels;
    float mfScaleFactor, mfLogScaleFactor;
    std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;

    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1) const
    {
        return grid.query(mvKeysUn, x, y, r, true, minLevel, maxLevel);
    }
    FeatureGrid grid;
};

#endif

#ifndef ORBX_REAL_MAPPOINT
inline int MapPoint::PredictScale(const float& currentDist, KeyFrame* pKF)
{
    const float ratio = mfMaxDistance / currentDist;
    const int n = (int)ceil(log(ratio) / pKF->mfLogScaleFactor);
    return n < 0 ? 0 : (n >= pKF->mnScaleLevels ? pKF->mnScaleLevels - 1 : n);
}
#endif
// MapPoint.cc:414-429; with ORBX_REAL_FRAME this text is compiled after the reference's Frame.h (ref_shim/frame/Frame.h)
#define ORBX_PREDICT_SCALE_FRAME \
    inline int MapPoint::PredictScale(const float& currentDist, Frame* pF) \
    { \
        const float ratio = mfMaxDistance / currentDist; \
        const int n = (int)ceil(log(ratio) / pF->mfLogScaleFactor); \
        return n < 0 ? 0 : (n >= pF->mnScaleLevels ? pF->mnScaleLevels - 1 : n); \
    }
#if !defined(ORBX_REAL_FRAME) && !defined(ORBX_REAL_MAPPOINT)
ORBX_PREDICT_SCALE_FRAME
#endif

}  // namespace ORB_SLAM2
#endif
