// mappoint_ref_api.cpp -- C ABI over the UNMODIFIED reference MapPoint (test infrastructure, not product).
//
// oracle/_ref/libmappoint_ref.so = /root/reference/src/MapPoint.cc and src/ORBmatcher.cc (for ORBmatcher::DescriptorDistance), compiled where
// they lie over the cv shim; KeyFrame / Frame / Map / ServerMapPoint / the boost archives are stand-ins (ref_shim/mappoint/MapPoint.h,
// ref_shim/slam/slam_objects.h), namespace renamed to ORB_SLAM2_REFM.  This file builds a MapPoint through the reference's own
// constructor, adds observations with the reference's AddObservation and calls the reference's ComputeDistinctiveDescriptors
// (MapPoint.cc:252-312), UpdateNormalAndDepth (:340-383), Get{Min,Max}DistanceInvariance (:385-395) and PredictScale (:397-429).
#include <opencv2/core/core.hpp>
#include "MapPoint.h"
#include "ORBmatcher.h"
#include <memory>

using namespace ORB_SLAM2;

namespace ORB_SLAM2
{
float Frame::fx, Frame::fy, Frame::cx, Frame::cy, Frame::invfx, Frame::invfy, Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
}

extern "C" {

// One MapPoint observed by n key frames.  Key frame i holds ONE feature: descriptor desc[32 i .. 32 i + 32), octave[i], camera centre
// centers[3 i ..], bad flag kf_bad[i]; key frame 0 is the reference key frame.  The key frames live in one array, so the address order the
// reference's std::map<KeyFrame*, size_t> iterates in (:271) is the index order.
// out_desc: GetDescriptor() (32 bytes; untouched and return value 0 if the reference left it empty); out7 = normal (3), GetMinDistanceInvariance,
// GetMaxDistanceInvariance, then for each of the m query distances PredictScale(dist, KeyFrame*) and PredictScale(dist, Frame*) in levels[2 m].
int refp_map_point(const float* pos3, int n, const uint8_t* desc, const int* octave, const float* centers, const uint8_t* kf_bad, float scaleFactor,
                   int nlevels, const float* query_dist, int m, uint8_t* out_desc, float* out5, int* levels)
{
    std::vector<KeyFrame> kfs(n);
    std::vector<float> sf(nlevels, 1.f);
    for (int l = 1; l < nlevels; l++) sf[l] = sf[l - 1] * scaleFactor;                         // ORBextractor.cc:419-423
    for (int i = 0; i < n; i++) {
        KeyFrame& K = kfs[i];
        K.mnId = i; K.mnFrameId = i; K.N = 1; K.bad = kf_bad[i] != 0;
        K.mvKeysUn.resize(1); K.mvKeysUn[0].octave = octave[i];
        K.mvuRight.assign(1, -1.f);
        K.mDescriptors = cv::Mat(1, 32, CV_8U);
        memcpy(K.mDescriptors.ptr<uint8_t>(0), desc + 32 * (size_t)i, 32);
        K.Ow = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) K.Ow.at<float>(k) = centers[3 * i + k];
        K.mvScaleFactors = sf; K.mnScaleLevels = nlevels; K.mfScaleFactor = scaleFactor; K.mfLogScaleFactor = log(scaleFactor);     // Frame.cc:186-187
        K.mvpMapPoints.assign(1, static_cast<MapPoint*>(NULL));
    }
    Map map;
    cv::Mat pos(3, 1, CV_32F);
    for (int k = 0; k < 3; k++) pos.at<float>(k) = pos3[k];
    MapPoint mp(pos, &kfs[0], &map);
    for (int i = 0; i < n; i++) mp.AddObservation(&kfs[i], 0);
    mp.ComputeDistinctiveDescriptors();
    mp.UpdateNormalAndDepth();
    const cv::Mat d = mp.GetDescriptor(), nv = mp.GetNormal();
    if (!d.empty()) memcpy(out_desc, d.ptr<uint8_t>(0), 32);
    for (int k = 0; k < 3; k++) out5[k] = nv.empty() ? 0.f : nv.at<float>(k);
    out5[3] = mp.GetMinDistanceInvariance();
    out5[4] = mp.GetMaxDistanceInvariance();
    Frame F;
    F.mnScaleLevels = nlevels; F.mfScaleFactor = scaleFactor; F.mfLogScaleFactor = log(scaleFactor);
    for (int q = 0; q < m; q++) {
        levels[2 * q] = mp.PredictScale(query_dist[q], &kfs[0]);
        levels[2 * q + 1] = mp.PredictScale(query_dist[q], &F);
    }
    return d.empty() ? 0 : 1;
}

}  // extern "C"
