// matcher_ref.h -- TEST INFRASTRUCTURE, not product.  Entry points of oracle/_ref/libmatcher_ref.so: the UNMODIFIED reference
// src/ORBmatcher.cc (kimwin2/ORB_SLAM2v2-1), compiled where it lies over the cv shim (oracle/ref_shim) and the stand-in SLAM objects
// (slam_objects.h), with its namespace renamed to ORB_SLAM2_REF so that it can sit in one process with the product's ORBmatcher.
// Each function builds the reference's ORBmatcher(nnratio, checkOri) and forwards to the member of the same name (file:line given);
// the reference's header never has to be visible to the caller (it does not exist on the GPU box).
#pragma once
#define ORB_SLAM2 ORB_SLAM2_REF
#include "slam_objects.h"
#undef ORB_SLAM2
#include <set>
#include <utility>
#include <vector>

namespace refm {
typedef ORB_SLAM2_REF::Frame Frame;
typedef ORB_SLAM2_REF::KeyFrame KeyFrame;
typedef ORB_SLAM2_REF::MapPoint MapPoint;

int SearchByProjection(float nnratio, bool checkOri, Frame& F, const std::vector<MapPoint*>& vpMapPoints, float th);                 // ORBmatcher.cc:45
int SearchByProjection(float nnratio, bool checkOri, Frame& CurrentFrame, const Frame& LastFrame, float th, bool bMono);             // :1330
int SearchByProjection(float nnratio, bool checkOri, Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
                       float th, int ORBdist);                                                                                          // :1474
int SearchByProjection(float nnratio, bool checkOri, KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                       std::vector<MapPoint*>& vpMatched, int th);                                                                      // :290
int SearchByBoW(float nnratio, bool checkOri, KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);                     // :159
int SearchByBoW(float nnratio, bool checkOri, KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);                    // :522
int SearchForInitialization(float nnratio, bool checkOri, Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                            std::vector<int>& vnMatches12, int windowSize);                                                            // :405
int SearchForTriangulation(float nnratio, bool checkOri, KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                           std::vector<std::pair<size_t, size_t> >& vMatchedPairs, bool bOnlyStereo);                                  // :657
int SearchBySim3(float nnratio, bool checkOri, KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, float s12,
                 const cv::Mat& R12, const cv::Mat& t12, float th);                                                                     // :1104
int Fuse(float nnratio, bool checkOri, KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, float th);                            // :827
int Fuse(float nnratio, bool checkOri, KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
         std::vector<MapPoint*>& vpReplacePoint);                                                                                       // :979
int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);                                                                            // :1649
}  // namespace refm
