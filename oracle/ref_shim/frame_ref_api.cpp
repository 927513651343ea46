// frame_ref_api.cpp -- C ABI over the UNMODIFIED reference Frame (test infrastructure, not product).
//
// oracle/_ref/libframe_ref.so = /root/reference/src/Frame.cc, src/ORBmatcher.cc, src/ORBextractor.cc and the vendored
// Thirdparty/DBoW2 (BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp, FORB.cpp, TemplatedVocabulary.h, DUtils/Random.cpp), each compiled
// where it lies (never copied), over the cv shim (ref_shim/opencv2) and the stand-in MapPoint / KeyFrame (ref_shim/slam/slam_objects.h),
// namespace renamed to ORB_SLAM2_REFF.  (The same file also builds oracle/_ref/libframe_dropin.so, where ORBextractor and ORBmatcher are THIS
// REPOSITORY'S drop-in headers instead of the reference's: see ref_shim/dropin/prelude_dropin.h.)  This file only builds Frames through the reference's own constructors, calls the reference's own
// member functions and copies members out; what Frame computes -- UndistortKeyPoints, ComputeImageBounds, AssignFeaturesToGrid,
// GetFeaturesInArea, ComputeStereoMatches, ComputeStereoFromRGBD, isInFrustum, ComputeBoW, UnprojectStereo -- is the reference's code.
//
// The stereo constructor (Frame.cc:61-117) calls ComputeStereoMatches() BEFORE it sets mb (:114), so the stock build reads an
// uninitialised float as minZ (:511-513).  kind 1 below therefore builds the left side through the monocular constructor (which sets mb),
// extracts the right image with the second extractor exactly as ExtractORB(1, imRight) does (:247-253), and then calls the public
// ComputeStereoMatches(): the same statements in the same order, with mb defined.
#include <opencv2/core/core.hpp>
#include "Frame.h"
#include "ORBmatcher.h"
#include "../orb_oracle.h"
#include <fstream>
#include <memory>

using namespace ORB_SLAM2;

struct reff_frame {
    std::unique_ptr<ORBextractor> left, right;
    std::unique_ptr<Frame> F;
    cv::Mat K, dist;
};

struct reff_voc {
    ORBVocabulary voc;
};

static void put_kps(const std::vector<cv::KeyPoint>& v, orbo_kp* out)
{
    for (size_t i = 0; i < v.size(); i++) {
        out[i].x = v[i].pt.x; out[i].y = v[i].pt.y; out[i].size = v[i].size; out[i].angle = v[i].angle; out[i].response = v[i].response;
        out[i].octave = v[i].octave; out[i].class_id = v[i].class_id;
    }
}

extern "C" {

// kind: 0 = monocular constructor (Frame.cc:174-228), 1 = stereo (see the note above), 2 = RGB-D constructor (:119-171; depth: h x w floats).
// K4 = (fx, fy, cx, cy); dist: 4 or 5 coefficients (k1 k2 p1 p2 [k3]); voc may be null.  The static "first frame" state is reset, so every
// frame computes its own image bounds (Frame::mbInitialComputations, :97-112).
reff_frame* reff_frame_create(int kind, const uint8_t* im, const uint8_t* im_right, const float* depth, int w, int h, int nfeatures, float scaleFactor,
                              int nlevels, int iniTh, int minTh, const float* K4, const float* dist, int ndist, float bf, float thDepth, reff_voc* voc)
{
    try {
        std::unique_ptr<reff_frame> r(new reff_frame);
        r->K = cv::Mat::eye(3, 3, CV_32F);
        r->K.at<float>(0, 0) = K4[0]; r->K.at<float>(1, 1) = K4[1]; r->K.at<float>(0, 2) = K4[2]; r->K.at<float>(1, 2) = K4[3];
        r->dist = cv::Mat(ndist, 1, CV_32F);                                  // Tracking.cc:63-73: a 4 x 1 (or 5 x 1) CV_32F column
        for (int i = 0; i < ndist; i++) r->dist.at<float>(i) = dist[i];
        r->left.reset(new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh));
        cv::Mat imL(h, w, CV_8UC1, const_cast<uint8_t*>(im));
        ORBVocabulary* pv = voc ? &voc->voc : static_cast<ORBVocabulary*>(NULL);
        Frame::mbInitialComputations = true;
        if (kind == 2) {
            cv::Mat imD(h, w, CV_32F, const_cast<float*>(depth));
            r->F.reset(new Frame(imL, imD, 0.0, r->left.get(), pv, r->K, r->dist, bf, thDepth));
        } else {
            r->F.reset(new Frame(imL, 0.0, r->left.get(), pv, r->K, r->dist, bf, thDepth));
            if (kind == 1) {
                r->right.reset(new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh));
                cv::Mat imR(h, w, CV_8UC1, const_cast<uint8_t*>(im_right));
                r->F->mpORBextractorRight = r->right.get();
                r->F->ExtractORB(1, imR);
                r->F->ComputeStereoMatches();
            }
        }
        return r.release();
    } catch (const std::exception& e) {
        fprintf(stderr, "reff_frame_create: %s\n", e.what());
        return NULL;
    }
}
void reff_frame_destroy(reff_frame* f) { delete f; }

int reff_frame_n(const reff_frame* f) { return f->F->N; }
int reff_frame_n_right(const reff_frame* f) { return (int)f->F->mvKeysRight.size(); }
float reff_frame_mb(const reff_frame* f) { return f->F->mb; }
// mvKeys, mvKeysUn, mDescriptors, mvuRight, mvDepth; bounds4 = (mnMinX, mnMaxX, mnMinY, mnMaxY); grid_inv2 = (mfGridElementWidthInv, ..HeightInv)
void reff_frame_get(const reff_frame* f, orbo_kp* keys, orbo_kp* keys_un, uint8_t* desc, float* u_right, float* depth, float* bounds4, float* grid_inv2)
{
    const Frame& F = *f->F;
    put_kps(F.mvKeys, keys);
    put_kps(F.mvKeysUn, keys_un);
    for (int i = 0; i < F.N; i++) memcpy(desc + 32 * (size_t)i, F.mDescriptors.ptr<uint8_t>(i), 32);
    for (int i = 0; i < F.N; i++) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    bounds4[0] = Frame::mnMinX; bounds4[1] = Frame::mnMaxX; bounds4[2] = Frame::mnMinY; bounds4[3] = Frame::mnMaxY;
    grid_inv2[0] = Frame::mfGridElementWidthInv; grid_inv2[1] = Frame::mfGridElementHeightInv;
}
void reff_frame_get_right(const reff_frame* f, orbo_kp* keys, uint8_t* desc)
{
    const Frame& F = *f->F;
    put_kps(F.mvKeysRight, keys);
    for (size_t i = 0; i < F.mvKeysRight.size(); i++) memcpy(desc + 32 * i, F.mDescriptorsRight.ptr<uint8_t>((int)i), 32);
}
// mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2 as the Frame copied them from the extractor (Frame.cc:184-190)
int reff_frame_scales(const reff_frame* f, float* sf, float* isf, float* sigma2, float* isigma2)
{
    const Frame& F = *f->F;
    for (int l = 0; l < F.mnScaleLevels; l++) { sf[l] = F.mvScaleFactors[l]; isf[l] = F.mvInvScaleFactors[l]; sigma2[l] = F.mvLevelSigma2[l]; isigma2[l] = F.mvInvLevelSigma2[l]; }
    return F.mnScaleLevels;
}
// ORBextractor::mvImagePyramid[level] of the left (side 0) or right (side 1) extractor: the view the reference's ComputeStereoMatches reads
int reff_frame_level(const reff_frame* f, int side, int level, int* w, int* h, const uint8_t** ptr, size_t* step)
{
    const ORBextractor* ex = side ? f->right.get() : f->left.get();
    if (!ex || level < 0 || level >= (int)ex->mvImagePyramid.size()) return -1;
    const cv::Mat& m = ex->mvImagePyramid[level];
    *w = m.cols; *h = m.rows; *ptr = m.data; *step = m.step;
    return 0;
}
// Frame::GetFeaturesInArea (Frame.cc:342-395), in the order the reference returns them
int reff_frame_features_in_area(const reff_frame* f, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap)
{
    const std::vector<size_t> v = f->F->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int)v[i];
    return (int)v.size();
}
// mGrid[col][row] as AssignFeaturesToGrid (Frame.cc:230-245) filled it
int reff_frame_grid_cell(const reff_frame* f, int col, int row, int* out, int cap)
{
    const std::vector<size_t>& v = f->F->mGrid[col][row];
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int)v[i];
    return (int)v.size();
}
// Frame::PosInGrid (Frame.cc:397-407) of an arbitrary undistorted position
int reff_frame_pos_in_grid(const reff_frame* f, float x, float y, int* posX, int* posY)
{
    cv::KeyPoint kp;
    kp.pt.x = x; kp.pt.y = y;
    return f->F->PosInGrid(kp, *posX, *posY) ? 1 : 0;
}
#ifndef ORBX_FRAME_API_REAL_MAPPOINT     // fills the stand-in MapPoint's fields directly; left out where MapPoint is the reference's own class (libframe_dropin.so)
// Frame::SetPose + isInFrustum (Frame.cc:255-340) for n MapPoints (world position, normal, distance range each).
// out: n x 6 floats = (returned bool, mTrackProjX, mTrackProjY, mTrackProjXR, mnTrackScaleLevel, mTrackViewCos)
void reff_frame_is_in_frustum(reff_frame* f, const float* Tcw16, const float* pos, const float* normal, const float* min_dist, const float* max_dist, int n,
                              float viewingCosLimit, float* out)
{
    cv::Mat T(4, 4, CV_32F);
    for (int i = 0; i < 16; i++) T.at<float>(i / 4, i % 4) = Tcw16[i];
    f->F->SetPose(T);
    for (int i = 0; i < n; i++) {
        MapPoint mp;
        mp.mWorldPos = cv::Mat(3, 1, CV_32F);
        mp.mNormalVector = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) { mp.mWorldPos.at<float>(k) = pos[3 * i + k]; mp.mNormalVector.at<float>(k) = normal[3 * i + k]; }
        mp.mfMinDistance = min_dist[i];
        mp.mfMaxDistance = max_dist[i];
        const bool in = f->F->isInFrustum(&mp, viewingCosLimit);
        float* o = out + 6 * (size_t)i;
        o[0] = in ? 1.f : 0.f; o[1] = mp.mTrackProjX; o[2] = mp.mTrackProjY; o[3] = mp.mTrackProjXR; o[4] = (float)mp.mnTrackScaleLevel; o[5] = mp.mTrackViewCos;
    }
}
#endif
// Frame::UnprojectStereo (Frame.cc:681-696) under the pose set by the last reff_frame_is_in_frustum: 3 floats, returns 0 where the reference returns an empty Mat
int reff_frame_unproject_stereo(reff_frame* f, int i, float* xyz)
{
    const cv::Mat p = f->F->UnprojectStereo(i);
    if (p.empty()) return 0;
    for (int k = 0; k < 3; k++) xyz[k] = p.at<float>(k);
    return 1;
}

// ---- one stereo unit after another with the extractors kept, as Tracking keeps them (Tracking.cc:88-95): the timed CPU baseline of the
// stereo configurations.  Per pair: Frame(imLeft, ...) [ExtractORB(0), UndistortKeyPoints, bounds, AssignFeaturesToGrid], ExtractORB(1, imRight),
// ComputeStereoMatches(); the Frame is destroyed before returning.  counts2 = (N, features with a stereo match).
struct reff_stereo_pipeline {
    std::unique_ptr<ORBextractor> left, right;
    cv::Mat K, dist;
    float bf, thDepth;
};
reff_stereo_pipeline* reff_stereo_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const float* K4, float bf, float thDepth)
{
    std::unique_ptr<reff_stereo_pipeline> p(new reff_stereo_pipeline);
    p->left.reset(new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh));
    p->right.reset(new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh));
    p->K = cv::Mat::eye(3, 3, CV_32F);
    p->K.at<float>(0, 0) = K4[0]; p->K.at<float>(1, 1) = K4[1]; p->K.at<float>(0, 2) = K4[2]; p->K.at<float>(1, 2) = K4[3];
    p->dist = cv::Mat::zeros(4, 1, CV_32F);                                  // stereo input is rectified (Frame.h:150-152)
    p->bf = bf; p->thDepth = thDepth;
    return p.release();
}
void reff_stereo_destroy(reff_stereo_pipeline* p) { delete p; }
int reff_stereo_run(reff_stereo_pipeline* p, const uint8_t* im_left, const uint8_t* im_right, int w, int h, int* counts2)
{
    try {
        cv::Mat imL(h, w, CV_8UC1, const_cast<uint8_t*>(im_left)), imR(h, w, CV_8UC1, const_cast<uint8_t*>(im_right));
        Frame F(imL, 0.0, p->left.get(), static_cast<ORBVocabulary*>(NULL), p->K, p->dist, p->bf, p->thDepth);
        F.mpORBextractorRight = p->right.get();
        F.ExtractORB(1, imR);
        F.ComputeStereoMatches();
        int matched = 0;
        for (int i = 0; i < F.N; i++) matched += F.mvuRight[i] >= 0;
        counts2[0] = F.N; counts2[1] = matched;
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "reff_stereo_run: %s\n", e.what());
        return -1;
    }
}

// ---- the vendored DBoW2 vocabulary (ORBVocabulary = TemplatedVocabulary<FORB::TDescriptor, FORB>), loaded with the reference's own
// loadFromTextFile (TemplatedVocabulary.h:1351-1436)
reff_voc* reff_voc_load_text(const char* path)
{
    std::unique_ptr<reff_voc> v(new reff_voc);
    try {
        if (!v->voc.loadFromTextFile(path)) return NULL;
    } catch (...) {
        return NULL;
    }
    return v.release();
}
void reff_voc_destroy(reff_voc* v) { delete v; }
int reff_voc_size(const reff_voc* v) { return (int)v->voc.size(); }
// Frame::ComputeBoW (Frame.cc:410-417): mBowVec as (word id, value) pairs and mFeatVec as CSR, both in map order.
// bow_id / bow_val: room for N entries; fv_node / fv_idx: room for N, fv_off for N + 1.  Returns the number of BowVector entries, -1 without a vocabulary.
int reff_frame_compute_bow(reff_frame* f, int* bow_id, double* bow_val, int* fv_node, int* fv_off, int* fv_idx, int* n_fv)
{
    Frame& F = *f->F;
    if (!F.mpORBvocabulary) return -1;
    F.mBowVec.clear();
    F.mFeatVec.clear();
    F.ComputeBoW();
    int nb = 0;
    for (DBoW2::BowVector::const_iterator it = F.mBowVec.begin(); it != F.mBowVec.end(); ++it, ++nb) { bow_id[nb] = (int)it->first; bow_val[nb] = it->second; }
    int nn = 0, ni = 0;
    fv_off[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = F.mFeatVec.begin(); it != F.mFeatVec.end(); ++it, ++nn) {
        fv_node[nn] = (int)it->first;
        for (size_t k = 0; k < it->second.size(); k++) fv_idx[ni++] = (int)it->second[k];
        fv_off[nn + 1] = ni;
    }
    *n_fv = nn;
    return nb;
}
// ORBVocabulary::transform on bare descriptors (n x 32 bytes), the call of Frame.cc:415 without a Frame
int reff_voc_transform(reff_voc* v, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val, int* fv_node, int* fv_off, int* fv_idx, int* n_fv)
{
    std::vector<cv::Mat> vd;
    vd.reserve(n);
    for (int i = 0; i < n; i++) vd.push_back(cv::Mat(1, 32, CV_8U, const_cast<uint8_t*>(desc + 32 * (size_t)i)));
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    v->voc.transform(vd, bv, fv, levelsup);
    int nb = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++nb) { bow_id[nb] = (int)it->first; bow_val[nb] = it->second; }
    int nn = 0, ni = 0;
    fv_off[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        fv_node[nn] = (int)it->first;
        for (size_t k = 0; k < it->second.size(); k++) fv_idx[ni++] = (int)it->second[k];
        fv_off[nn + 1] = ni;
    }
    *n_fv = nn;
    return nb;
}
// DBoW2's score between two BowVectors with the vocabulary's scoring object (TemplatedVocabulary::score, used by KeyFrameDatabase)
double reff_voc_score(reff_voc* v, const int* id1, const double* val1, int n1, const int* id2, const double* val2, int n2)
{
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; i++) a[(DBoW2::WordId)id1[i]] = val1[i];
    for (int i = 0; i < n2; i++) b[(DBoW2::WordId)id2[i]] = val2[i];
    return v->voc.score(a, b);
}
// loadFromBinaryFile (TemplatedVocabulary.h:1467-1511; System.cc:84)
reff_voc* reff_voc_load_binary(const char* path)
{
    std::unique_ptr<reff_voc> v(new reff_voc);
    try {
        if (!v->voc.loadFromBinaryFile(path)) return NULL;
    } catch (...) {
        return NULL;
    }
    return v.release();
}
#ifndef ORBX_FRAME_API_REAL_MAPPOINT     // DBoW2 only: the drop-in vocabulary reads this format, it does not write it
void reff_voc_save_binary(reff_voc* v, const char* path) { v->voc.saveToBinaryFile(path); }
#endif

}  // extern "C"
