// orb_oracle.h -- CPU ORACLE for the ORB front end (TEST INFRASTRUCTURE, NOT PRODUCT).
//
// A scalar C++ restatement of the reference's ORBextractor / ORBmatcher /
// Frame::ComputeStereoMatches (kimwin2/ORB_SLAM2v2-1), with the OpenCV primitives it
// calls restated from their published semantics and pinned against cv2 4.13.0 by
// tests/test_oracle_vs_cv2.py.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library.  The product
// (orb_slam2v2-1_b200/) never links, imports or calls it.
//
// Parity status: the reference ships no tests or golden vectors ("parity unpinned" by
// the reference itself); the oracle is pinned to OpenCV 4.13.0 primitive semantics and
// to a second, independent Python+cv2 orchestration (oracle/cv2_cross_oracle.py).
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

// Same field order and size (28 B) as cv::KeyPoint {pt.x, pt.y, size, angle, response, octave, class_id}.
typedef struct orbo_kp {
    float x, y, size, angle, response;
    int octave, class_id;
} orbo_kp;

typedef struct orbo_extractor orbo_extractor;

// ORBextractor::ORBextractor (ORBextractor.cc:410-470)
orbo_extractor* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
void orbo_destroy(orbo_extractor*);
// tie_mode: 0 = creation-order tie-break in the largest-first phase (portable definition),
//           1 = verbatim std::list + heap-address tie-break (what the reference binary does on this libc).
// fma_mode: 0 = uncontracted rotation in computeOrbDescriptor (portable), 1 = FMA-contracted as g++ -O3 -march=native emits.
void orbo_set_modes(orbo_extractor*, int tie_mode, int fma_mode);
// ORBextractor::operator() (ORBextractor.cc:1043-1105). Returns the keypoint count (may exceed cap: then only cap are written).
int orbo_extract(orbo_extractor*, const uint8_t* img, int w, int h, size_t stride, orbo_kp* kps, uint8_t* desc, int cap);
// Tables: out[0]=mvScaleFactor, out[1]=mvInvScaleFactor, out[2]=mvLevelSigma2, out[3]=mvInvLevelSigma2 (nlevels floats each).
void orbo_scale_tables(orbo_extractor*, float* sf, float* isf, float* sig2, float* isig2);
void orbo_features_per_level(orbo_extractor*, int* out);
void orbo_umax(orbo_extractor*, int* out16);
// Stage dumps of the LAST orbo_extract call (known-answer tests per kernel).
void orbo_level_size(orbo_extractor*, int level, int* w, int* h);
// Padded (w+38)x(h+38) pyramid buffer of a level, dst stride = w+38.
void orbo_level_padded(orbo_extractor*, int level, uint8_t* dst);
// Blurred level (w x h), dst stride = w. Levels with zero keypoints are still blurred here for testing.
void orbo_level_blurred(orbo_extractor*, int level, uint8_t* dst);
// FAST candidates handed to DistributeOctTree, in order: x,y relative to minBorder (16), response. Returns count.
int orbo_level_candidates(orbo_extractor*, int level, int* xyr, int cap);
int orbo_level_nkeys(orbo_extractor*, int level);
// per-stage CPU seconds of the last call: pyramid, fast, octree, orient, blur, desc
void orbo_stage_seconds(orbo_extractor*, double* out6);

// ---- primitives (unit-tested against cv2) ----
void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh, size_t dstride);
void orbo_border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, int b, size_t dstride);
void orbo_gauss7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride);
// cv::FAST(roi, kps, threshold, true) on a w x h view; writes x,y,response triples; returns count.
int orbo_fast(const uint8_t* roi, int w, int h, size_t stride, int threshold, int nms, int* xyr, int cap);
float orbo_fast_atan2(float y, float x);
/* cvtColor RGB/BGR/RGBA/BGRA -> GRAY, OpenCV 8-bit fixed point (Tracking.cc:280-305) */
void orbo_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int rgb, uint8_t* dst, size_t dstride);
// DistributeOctTree alone (ORBextractor.cc:539-763): in xyr triples -> selected indices (into the input) in output order.
int orbo_distribute(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int tie_mode, int* out_idx, int cap);

// ---- matcher ----
// ORBmatcher::DescriptorDistance (ORBmatcher.cc:1649-1665)
int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b);
// Inner loop of SearchByBoW (ORBmatcher.cc:205-230) over whole sets: per query best/second/argbest in train index order
// (strict < for best, else-if < for second, both start at 256).
void orbo_match_bruteforce(const uint8_t* q, int nq, const uint8_t* t, int nt,
                           int* best_idx, int* best_dist, int* second_dist);
// Generic candidate-list search: query i scans train rows cand_idx[cand_off[i]..cand_off[i+1]) in that order.
// Also returns second-best index (-1 if none) for the wrappers that need levels.
void orbo_match_candidates(const uint8_t* q, int nq, const uint8_t* t, int nt,
                           const int* cand_off, const int* cand_idx,
                           int* best_idx, int* best_dist, int* second_idx, int* second_dist);

// Frame grid (Frame.cc:230-245, 342-407): 64 x 48 cells over [minX,maxX) x [minY,maxY).
typedef struct orbo_grid orbo_grid;
orbo_grid* orbo_grid_build(const orbo_kp* kps, int n, float minX, float maxX, float minY, float maxY);
void orbo_grid_destroy(orbo_grid*);
// Frame::GetFeaturesInArea (Frame.cc:342-395). Returns count (indices in reference order).
int orbo_grid_query(const orbo_grid*, const orbo_kp* kps, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap);

// ORBmatcher::SearchForInitialization (ORBmatcher.cc:405-520). prev_matched: in/out N1 x 2 floats. matches12: N1 ints.
int orbo_search_for_initialization(const orbo_kp* k1, const uint8_t* d1, int n1,
                                   const orbo_kp* k2, const uint8_t* d2, int n2,
                                   float minX, float maxX, float minY, float maxY,
                                   float* prev_matched, int* matches12, int windowSize, float nnratio, int checkOri);

// Frame::ComputeStereoMatches (Frame.cc:481-655) with maxD explicit (the reference reads an uninitialised mb).
// Pyramids are given as per-level interior pointers (unblurred), sizes and strides.
/* plain-array restatements of ORBmatcher::SearchByProjection(Frame, MapPoints) (ORBmatcher.cc:45-129),
 * SearchByProjection(CurrentFrame, LastFrame) (:1330-1472) and both SearchByBoW (:159-288, :522-655); see orb_matcher.cpp */
int orbo_search_by_projection(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied, int n,
                              float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                              const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos,
                              const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, float nnratio, int* feature_match);
int orbo_search_by_projection_frame(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied, int n,
                                    float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                    const float* proj_xy, const float* proj_ur, const int* last_octave, const float* last_angle,
                                    const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                    int* feature_match);
int orbo_search_by_bow(int variant, const orbo_kp* k1, const uint8_t* d1, int n1, const uint8_t* valid1,
                       const int* node1, const int* off1, const int* idx1, int nn1,
                       const orbo_kp* k2, const uint8_t* d2, int n2, const uint8_t* valid2,
                       const int* node2, const int* off2, const int* idx2, int nn2,
                       float nnratio, int checkOri, int* matches);

/* the LocalMapping / LoopClosing searches (Fuse x2 ORBmatcher.cc:827-1102, SearchBySim3 :1104-1328, SearchByProjection(KF, Scw)
 * :290-403, SearchByProjection(Frame, KF) :1474-1601, SearchForTriangulation :657-825); see orb_matcher.cpp */
void orbo_search_in_area(const orbo_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip, int n,
                         float minX, float maxX, float minY, float maxY, const float* qxy, const float* qur, const float* qr,
                         const int* qlevel, const uint8_t* qdesc, int nq, const float* inv_level_sigma2, int* best_idx, int* best_dist);
int orbo_fuse(const orbo_kp* kps, const uint8_t* desc, const float* u_right, int n, float minX, float maxX, float minY, float maxY,
              const float* sf, const float* inv_level_sigma2, int nlevels, const float* proj_xy, const float* proj_ur,
              const int* level, const uint8_t* mp_desc, int nmp, float th, int stereo_chi2, int* best_idx);
int orbo_search_by_sim3(const orbo_kp* k1, const uint8_t* d1, int n1, const orbo_kp* k2, const uint8_t* d2, int n2,
                        float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                        const float* proj12_xy, const int* level12, const uint8_t* mp1_desc,
                        const float* proj21_xy, const int* level21, const uint8_t* mp2_desc, float th, int* match12);
int orbo_search_by_projection_sim3(const orbo_kp* kps, const uint8_t* desc, const uint8_t* matched_in, int n,
                                   float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                   const float* proj_xy, const int* level, const uint8_t* mp_desc, int nmp, float th, int* feature_match);
int orbo_search_by_projection_kf(const orbo_kp* kps, const uint8_t* desc, const uint8_t* occupied_in, int n,
                                 float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                 const float* proj_xy, const int* level, const float* kf_angle, const uint8_t* mp_desc, int nmp,
                                 float th, int ORBdist, int checkOri, int* feature_match);
int orbo_search_for_triangulation(const orbo_kp* k1, const uint8_t* d1, int n1, const uint8_t* has_mp1, const float* u_right1,
                                  const int* node1, const int* off1, const int* idx1, int nn1,
                                  const orbo_kp* k2, const uint8_t* d2, int n2, const uint8_t* has_mp2, const float* u_right2,
                                  const int* node2, const int* off2, const int* idx2, int nn2,
                                  const float* F12, float ex, float ey, const float* sf2, const float* sigma2_2, int nlevels,
                                  int onlyStereo, int checkOri, int* matches12);

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:281-311): index of the descriptor with the least median distance */
int orbo_distinctive_descriptor(const uint8_t* desc, int n);
void orbo_distinctive_descriptors(const uint8_t* desc, const int* off, int npoints, int* best);

/* Frame glue: cv::undistortPoints as called at Frame.cc:438/:464, Frame::UndistortKeyPoints (:419-449), ComputeImageBounds
 * (:451-479), ComputeStereoFromRGBD (:658-679) */
void orbo_undistort_points(const float* pts, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* out);
void orbo_undistort_keypoints(const orbo_kp* kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, orbo_kp* out);
void orbo_image_bounds(int cols, int rows, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* bounds4);
void orbo_stereo_from_rgbd(const orbo_kp* kps, const orbo_kp* kps_un, int n, const float* im_depth, size_t stride_bytes, float bf,
                           float* u_right, float* depth);

/* ORBVocabulary::transform (DBoW2 TemplatedVocabulary.h:1140-1270) on a tree given as arrays; see orb_vocabulary.cpp */
typedef struct orbo_vocabulary orbo_vocabulary;
orbo_vocabulary* orbo_voc_create(int k, int L, int scoring, int weighting, const int* parent, const uint8_t* desc, const double* weight, int nnodes);
void orbo_voc_destroy(orbo_vocabulary* v);
int orbo_voc_size(const orbo_vocabulary* v);
void orbo_voc_transform_raw(const orbo_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word_id, double* weight, int* node_id);
int orbo_voc_transform(const orbo_vocabulary* voc, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val,
                       int* fv_node, int* fv_off, int* fv_idx, int* nfv);

int orbo_stereo_matches(const orbo_kp* kl, const uint8_t* dl, int nl,
                        const orbo_kp* kr, const uint8_t* dr, int nr,
                        const uint8_t* const* pyrL, const uint8_t* const* pyrR,
                        const int* lw, const int* lh, const size_t* lstrideL, const size_t* lstrideR,
                        const float* scaleFactors, const float* invScaleFactors, int nlevels,
                        float bf, float minD, float maxD, float* uRight, float* depth);

#ifdef __cplusplus
}
#endif
