// ORBmatcher.h -- drop-in for the descriptor-level part of the reference's include/ORBmatcher.h
// (kimwin2/ORB_SLAM2v2-1 include/ORBmatcher.h:37-102, src/ORBmatcher.cc) and of Frame::ComputeStereoMatches
// (src/Frame.cc:481-655) over the C ABI of liborbx.so (include/orbx.h).  Same class name, constructor, constants and
// return conventions.  Two layers:
//   * the tracking searches under the reference's own signatures -- SearchByProjection(Frame&, vpMapPoints, th),
//     SearchByProjection(CurrentFrame, LastFrame, th, bMono), SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches),
//     SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12), SearchForInitialization(Frame&, Frame&, ...) -- as member templates
//     over the user's Frame / KeyFrame / MapPoint types (ORBmatcher_reference_api.inl), so the call sites in Tracking.cc and
//     LoopClosing.cc compile unchanged while this repository stays free of the SLAM object graph (Map, ROS messages);
//   * the same searches, and the LocalMapping / LoopClosing ones, over a FrameView + plain arrays (the members the routines
//     read), which is what the templates call.  Predicates that read map state (isBad, Observations, epipolar checks) stay on
//     the host side, as in SURVEY.md section 8(a) a15.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <set>
#include <vector>
#include "orbx.h"
#include "orbx_cv_shim.h"
#include "ORBextractor.h"

namespace ORB_SLAM2
{

// What the matcher reads from a Frame: mvKeysUn, mDescriptors (N x 32, CV_8U, continuous) and the grid bounds
// mnMinX, mnMaxX, mnMinY, mnMaxY (Frame.h:188-191).
struct FrameView {
    const std::vector<cv::KeyPoint>* keys;
    const cv::Mat* descriptors;
    float minX, maxX, minY, maxY;
};

class ORBmatcher;

// A Frame that lives on the device (orbm_frame, include/orbx.h): mvKeysUn, mDescriptors and the 64 x 48 grid of
// Frame::AssignFeaturesToGrid (src/Frame.cc:230-245) built once per frame in HBM, so that the windowed searches of one Tracking step
// (SearchByProjection x2, SearchForInitialization) move only the MapPoint side.  Build it once per frame, next to Frame's own grid.
class DeviceFrame
{
public:
    // from the host members of a Frame: keysUn = mvKeysUn, descriptors = mDescriptors, bounds = mnMinX, mnMaxX, mnMinY, mnMaxY
    DeviceFrame(ORBmatcher& matcher, const std::vector<cv::KeyPoint>& keysUn, const cv::Mat& descriptors, float minX, float maxX, float minY, float maxY);
    // straight from the device results of the extractor's last call (no descriptor round trip); Frame::UndistortKeyPoints
    // (src/Frame.cc:419-449) runs on the device when distCoef (mDistCoef as floats) is given and distCoef[0] != 0
    DeviceFrame(ORBmatcher& matcher, ORBextractor& extractor, const float K4[4], const std::vector<float>& distCoef, float minX, float maxX, float minY,
                float maxY, int frame = 0);
    ~DeviceFrame() { orbm_frame_destroy(f_); }
    DeviceFrame(const DeviceFrame&) = delete;
    DeviceFrame& operator=(const DeviceFrame&) = delete;
    int size() const { return orbm_frame_size(f_); }
    std::vector<cv::KeyPoint> keypoints() const            // mvKeysUn as the device holds it
    {
        std::vector<orbx_kp> k(size());
        if (orbm_frame_keypoints(f_, k.data()) != ORBX_OK) throw std::runtime_error(std::string("orbm_frame_keypoints: ") + orbx_last_error());
        std::vector<cv::KeyPoint> out(k.size());
        for (size_t i = 0; i < k.size(); i++) {
            out[i].pt.x = k[i].x; out[i].pt.y = k[i].y; out[i].size = k[i].size; out[i].angle = k[i].angle;
            out[i].response = k[i].response; out[i].octave = k[i].octave; out[i].class_id = k[i].class_id;
        }
        return out;
    }
    orbm_frame* handle() const { return f_; }
private:
    orbm_frame* f_ = nullptr;
};

class ORBmatcher
{
public:
    // device: CUDA ordinal, -1 = the calling thread's current device.  Construction is cheap enough for the reference's habit
    // of building a matcher on the stack per call (Tracking.cc:742,907,1012,...): the library recycles the context (stream,
    // device and page-locked arenas) of the last matcher this thread destroyed.
    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = -1) : mfNNratio(nnratio), mbCheckOrientation(checkOri), m_(nullptr)
    {
        if (orbm_create(device, &m_) != ORBX_OK) fail("orbm_create");
    }
    ~ORBmatcher() { orbm_destroy(m_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // Computes the Hamming distance between two ORB descriptors (ORBmatcher.cc:1649-1665).  A single pair is 8 XOR + popcount
    // on the host -- the reference calls this inline N^2 times (MapPoint.cc:286-296), where a kernel launch per pair would be
    // ~10^3 x slower; the device forms are DescriptorDistances() (many pairs), SearchBruteForce() and
    // ComputeDistinctiveDescriptors() (the whole MapPoint.cc:252-312 loop for a batch of points).
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
    {
        uint32_t pa[8], pb[8];
        memcpy(pa, a.ptr<uint8_t>(0), 32);
        memcpy(pb, b.ptr<uint8_t>(0), 32);
        int dist = 0;
        for (int i = 0; i < 8; i++) dist += __builtin_popcount(pa[i] ^ pb[i]);
        return dist;
    }
    std::vector<int> DescriptorDistances(const cv::Mat& a, const cv::Mat& b)      // row i of a against row i of b
    {
        std::vector<int> d(a.rows);
        if (orbm_descriptor_distance(m_, a.ptr<uint8_t>(0), b.ptr<uint8_t>(0), a.rows, d.data()) != ORBX_OK) fail("orbm_descriptor_distance");
        return d;
    }

    // Matching for the Map Initialization (ORBmatcher.cc:405-520).  Returns the number of matches.
    int SearchForInitialization(const FrameView& F1, const FrameView& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10)
    {
        std::vector<orbx_kp> k1 = pack(*F1.keys), k2 = pack(*F2.keys);
        std::vector<float> prev(2 * k1.size());
        for (size_t i = 0; i < k1.size(); i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        vnMatches12.assign(k1.size(), -1);
        int n = 0;
        if (orbm_search_for_initialization(m_, k1.data(), F1.descriptors->ptr<uint8_t>(0), (int)k1.size(), k2.data(),
                                           F2.descriptors->ptr<uint8_t>(0), (int)k2.size(), F2.minX, F2.maxX, F2.minY, F2.maxY,
                                           prev.data(), vnMatches12.data(), windowSize, mfNNratio, mbCheckOrientation, &n) != ORBX_OK)
            fail("orbm_search_for_initialization");
        for (size_t i = 0; i < k1.size(); i++) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
        return n;
    }

    // ---- the reference's own signatures (include/ORBmatcher.h:47-82 of the reference); definitions in ORBmatcher_reference_api.inl
    template <class FrameT, class MapPointT> int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3);
    template <class FrameT> int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono);
    template <class KeyFrameT, class FrameT, class MapPointT> int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches);
    template <class KeyFrameT, class MapPointT> int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12);
    template <class FrameT, class = typename std::enable_if<!std::is_same<typename std::remove_cv<FrameT>::type, FrameView>::value &&
                                                            !std::is_same<typename std::remove_cv<FrameT>::type, DeviceFrame>::value>::type>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10);
#ifdef ORBX_USE_OPENCV
    // The LocalMapping / LoopClosing / relocalisation searches under the reference's signatures (include/ORBmatcher.h:56-82 there; call
    // sites Tracking.cc:1597,1611, LocalMapping.cc:276,497,522, LoopClosing.cc:333,385,608).  They project with the same cv::Mat
    // expressions as the reference (so the floats are OpenCV's own) and therefore exist only in builds against OpenCV.
    template <class FrameT, class KeyFrameT, class MapPointT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th, const int ORBdist);
    template <class KeyFrameT, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th);
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo);
    template <class KeyFrameT, class MapPointT>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12,
                     const float th);
    template <class KeyFrameT, class MapPointT> int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0);
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint);
#endif

    // ---- the same searches against a DeviceFrame (definitions after the class)
    int SearchForInitialization(const DeviceFrame& F1, const DeviceFrame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10)
    {
        std::vector<float> prev(2 * vbPrevMatched.size());
        for (size_t i = 0; i < vbPrevMatched.size(); i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        vnMatches12.assign(F1.size(), -1);
        int n = 0;
        if (orbm_search_for_initialization_dev(m_, F1.handle(), F2.handle(), prev.data(), vnMatches12.data(), windowSize, mfNNratio, mbCheckOrientation,
                                               &n) != ORBX_OK)
            fail("orbm_search_for_initialization_dev");
        for (size_t i = 0; i < vbPrevMatched.size(); i++) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
        return n;
    }
    orbx_matcher* handle() const { return m_; }

    // Best / second-best scan of SearchByBoW's inner loop (ORBmatcher.cc:205-230) over whole descriptor sets, with the
    // acceptance test bestDist1 <= TH_LOW && bestDist1 < mfNNratio * bestDist2.  Returns the number accepted;
    // vnMatches[i] = train row or -1.
    int SearchBruteForce(const cv::Mat& query, const cv::Mat& train, std::vector<int>& vnMatches)
    {
        const int nq = query.rows;
        std::vector<int> bi(nq), bd(nq), sd(nq);
        std::vector<uint8_t> acc(nq);
        if (orbm_match_ratio(m_, query.ptr<uint8_t>(0), nq, train.ptr<uint8_t>(0), train.rows, TH_LOW, mfNNratio, bi.data(), bd.data(),
                             sd.data(), acc.data()) != ORBX_OK)
            fail("orbm_match_ratio");
        vnMatches.assign(nq, -1);
        int n = 0;
        for (int i = 0; i < nq; i++)
            if (acc[i]) { vnMatches[i] = bi[i]; n++; }
        return n;
    }

    // Windowed / bucketed searches (SearchByProjection, SearchByBoW): the caller supplies, per query, the candidate rows
    // in the reference's visiting order (GetFeaturesInArea order or FeatureVector bucket order) as CSR lists.
    void SearchCandidates(const cv::Mat& query, const cv::Mat& train, const std::vector<int>& candOff, const std::vector<int>& candIdx,
                          std::vector<int>& bestIdx, std::vector<int>& bestDist, std::vector<int>& secondIdx, std::vector<int>& secondDist)
    {
        const int nq = query.rows;
        bestIdx.resize(nq); bestDist.resize(nq); secondIdx.resize(nq); secondDist.resize(nq);
        if (orbm_match_candidates(m_, query.ptr<uint8_t>(0), nq, train.ptr<uint8_t>(0), train.rows, candOff.data(), candIdx.data(),
                                  bestIdx.data(), bestDist.data(), secondIdx.data(), secondDist.data()) != ORBX_OK)
            fail("orbm_match_candidates");
    }

    // What the projection searches read from a Frame besides FrameView: mvuRight, which features already hold a MapPoint with
    // Observations() > 0, and mvScaleFactors.
    struct FrameState {
        const std::vector<float>* uRight;          // Frame::mvuRight
        const std::vector<uint8_t>* occupied;      // F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0
        const std::vector<float>* scaleFactors;    // Frame::mvScaleFactors
    };
    // The MapPoint fields the routine reads, one entry per point that passed the reference's own skip tests, in visit order.
    struct ProjectedPoints {
        std::vector<float> xy;                      // (mTrackProjX, mTrackProjY) resp. the projection (u, v)
        std::vector<float> xr;                      // mTrackProjXR resp. u - mbf * invzc
        std::vector<int> level;                     // mnTrackScaleLevel resp. LastFrame.mvKeys[i].octave
        std::vector<float> viewCosOrAngle;          // mTrackViewCos resp. LastFrame.mvKeysUn[i].angle
        std::vector<uint8_t> descriptors;           // GetDescriptor(), 32 bytes per point
        std::vector<uint8_t> hasObservations;       // Observations() > 0
    };

    // Search matches between Frame keypoints and projected MapPoints, used to track the local map (ORBmatcher.cc:45-129).
    // vnMatch[j] = index (into P) of the point the reference would store in F.mvpMapPoints[j], -1 = untouched.
    int SearchByProjection(const FrameView& F, const FrameState& S, const ProjectedPoints& P, std::vector<int>& vnMatch, const float th = 3)
    {
        std::vector<orbx_kp> k = pack(*F.keys);
        vnMatch.assign(k.size(), -1);
        int n = 0;
        if (orbm_search_by_projection(m_, k.data(), F.descriptors->ptr<uint8_t>(0), S.uRight->data(), S.occupied->data(), (int)k.size(), F.minX,
                                      F.maxX, F.minY, F.maxY, S.scaleFactors->data(), (int)S.scaleFactors->size(), P.xy.data(), P.xr.data(),
                                      P.level.data(), P.viewCosOrAngle.data(), P.descriptors.data(), P.hasObservations.data(),
                                      (int)P.level.size(), th, mfNNratio, vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection");
        return n;
    }

    // Project MapPoints tracked in the last frame into the current frame and search matches, used to track from the previous
    // frame (ORBmatcher.cc:1330-1472).  mode: 0 = neither, 1 = bForward, 2 = bBackward (:1351-1352).
    int SearchByProjection(const FrameView& CurrentFrame, const FrameState& S, const ProjectedPoints& P, std::vector<int>& vnMatch, const float th,
                           const int mode)
    {
        std::vector<orbx_kp> k = pack(*CurrentFrame.keys);
        vnMatch.assign(k.size(), -1);
        int n = 0;
        if (orbm_search_by_projection_frame(m_, k.data(), CurrentFrame.descriptors->ptr<uint8_t>(0), S.uRight->data(), S.occupied->data(),
                                            (int)k.size(), CurrentFrame.minX, CurrentFrame.maxX, CurrentFrame.minY, CurrentFrame.maxY,
                                            S.scaleFactors->data(), (int)S.scaleFactors->size(), P.xy.data(), P.xr.data(), P.level.data(),
                                            P.viewCosOrAngle.data(), P.descriptors.data(), P.hasObservations.data(), (int)P.level.size(), th, mode,
                                            mbCheckOrientation, vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection_frame");
        return n;
    }

    int SearchByProjection(const DeviceFrame& F, const FrameState& S, const ProjectedPoints& P, std::vector<int>& vnMatch, const float th = 3)
    {
        vnMatch.assign(F.size(), -1);
        int n = 0;
        if (orbm_search_by_projection_dev(m_, F.handle(), S.uRight->data(), S.occupied->data(), S.scaleFactors->data(), (int)S.scaleFactors->size(),
                                          P.xy.data(), P.xr.data(), P.level.data(), P.viewCosOrAngle.data(), P.descriptors.data(), P.hasObservations.data(),
                                          (int)P.level.size(), th, mfNNratio, vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection_dev");
        return n;
    }
    int SearchByProjection(const DeviceFrame& CurrentFrame, const FrameState& S, const ProjectedPoints& P, std::vector<int>& vnMatch, const float th,
                           const int mode)
    {
        vnMatch.assign(CurrentFrame.size(), -1);
        int n = 0;
        if (orbm_search_by_projection_frame_dev(m_, CurrentFrame.handle(), S.uRight->data(), S.occupied->data(), S.scaleFactors->data(),
                                                (int)S.scaleFactors->size(), P.xy.data(), P.xr.data(), P.level.data(), P.viewCosOrAngle.data(),
                                                P.descriptors.data(), P.hasObservations.data(), (int)P.level.size(), th, mode, mbCheckOrientation,
                                                vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection_frame_dev");
        return n;
    }

    // A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned int>>) flattened to CSR, in map order.
    struct FeatureVectorCSR {
        std::vector<int> node, off, idx;
        template <class Map> static FeatureVectorCSR from(const Map& fv)
        {
            FeatureVectorCSR c;
            c.off.push_back(0);
            for (typename Map::const_iterator it = fv.begin(); it != fv.end(); ++it) {
                c.node.push_back((int)it->first);
                for (size_t i = 0; i < it->second.size(); i++) c.idx.push_back((int)it->second[i]);
                c.off.push_back((int)c.idx.size());
            }
            return c;
        }
    };

    // Search matches between MapPoints in a KeyFrame and ORB in a Frame, brute force constrained to ORB of the same vocabulary
    // node (ORBmatcher.cc:159-288).  vbValidKF[i]: feature i of the key frame has a MapPoint that is not bad.
    // vnMatchF[j] = key-frame feature whose MapPoint goes to frame feature j, -1 = none.
    int SearchByBoW(const FrameView& KF, const std::vector<uint8_t>& vbValidKF, const FeatureVectorCSR& fvKF, const FrameView& F,
                    const FeatureVectorCSR& fvF, std::vector<int>& vnMatchF)
    {
        std::vector<orbx_kp> k1 = pack(*KF.keys), k2 = pack(*F.keys);
        vnMatchF.assign(k2.size(), -1);
        int n = 0;
        if (orbm_search_by_bow(m_, 0, k1.data(), KF.descriptors->ptr<uint8_t>(0), (int)k1.size(), vbValidKF.data(), fvKF.node.data(),
                               fvKF.off.data(), fvKF.idx.data(), (int)fvKF.node.size(), k2.data(), F.descriptors->ptr<uint8_t>(0), (int)k2.size(),
                               nullptr, fvF.node.data(), fvF.off.data(), fvF.idx.data(), (int)fvF.node.size(), mfNNratio, mbCheckOrientation,
                               vnMatchF.data(), &n) != ORBX_OK)
            fail("orbm_search_by_bow");
        return n;
    }
    // Key frame against key frame (loop closing, ORBmatcher.cc:522-655).  vnMatches12[i] = matched feature of the second key frame.
    int SearchByBoW(const FrameView& KF1, const std::vector<uint8_t>& vbValid1, const FeatureVectorCSR& fv1, const FrameView& KF2,
                    const std::vector<uint8_t>& vbValid2, const FeatureVectorCSR& fv2, std::vector<int>& vnMatches12)
    {
        std::vector<orbx_kp> k1 = pack(*KF1.keys), k2 = pack(*KF2.keys);
        vnMatches12.assign(k1.size(), -1);
        int n = 0;
        if (orbm_search_by_bow(m_, 1, k1.data(), KF1.descriptors->ptr<uint8_t>(0), (int)k1.size(), vbValid1.data(), fv1.node.data(),
                               fv1.off.data(), fv1.idx.data(), (int)fv1.node.size(), k2.data(), KF2.descriptors->ptr<uint8_t>(0), (int)k2.size(),
                               vbValid2.data(), fv2.node.data(), fv2.off.data(), fv2.idx.data(), (int)fv2.node.size(), mfNNratio,
                               mbCheckOrientation, vnMatches12.data(), &n) != ORBX_OK)
            fail("orbm_search_by_bow");
        return n;
    }

    // ---- LocalMapping / LoopClosing / relocalisation searches.  ProjectedPoints carries, per MapPoint that passed the reference's
    // skip tests, xy = projection (u, v), xr = u - bf * invz (Fuse only), level = nPredictedLevel, descriptors = GetDescriptor().

    // Project MapPoints into a KeyFrame and search for duplicated MapPoints (ORBmatcher.cc:827-975).  vnBestIdx[i] = key-frame
    // feature point i fuses into, -1 = none; the caller then applies :940-957 (Replace / AddObservation) to those features.
    int Fuse(const FrameView& KF, const std::vector<float>& vuRight, const std::vector<float>& vScaleFactors,
             const std::vector<float>& vInvLevelSigma2, const ProjectedPoints& P, std::vector<int>& vnBestIdx, const float th = 3.0)
    {
        return fuse(KF, &vuRight, vScaleFactors, &vInvLevelSigma2, P, vnBestIdx, th, 1);
    }
    // Project MapPoints into a KeyFrame using a given Sim3 and search for duplicated MapPoints (ORBmatcher.cc:977-1102).
    int Fuse(const FrameView& KF, const std::vector<float>& vScaleFactors, const ProjectedPoints& P, std::vector<int>& vnBestIdx, const float th)
    {
        return fuse(KF, nullptr, vScaleFactors, nullptr, P, vnBestIdx, th, 0);
    }

    // Search matches between MapPoints seen in KF1 and KF2 transforming by a Sim3 [s12*R12|t12] (ORBmatcher.cc:1104-1328).
    // P12: one entry PER FEATURE of key frame 1 (level < 0 where :1145-1173 skip the feature), projected into key frame 2; P21 likewise.
    // vnMatches12[i1] = i2 where the two directions agree, -1 otherwise.
    int SearchBySim3(const FrameView& KF1, const FrameView& KF2, const std::vector<float>& vScaleFactors, const ProjectedPoints& P12,
                     const ProjectedPoints& P21, std::vector<int>& vnMatches12, const float th)
    {
        std::vector<orbx_kp> k1 = pack(*KF1.keys), k2 = pack(*KF2.keys);
        vnMatches12.assign(k1.size(), -1);
        int n = 0;
        if (orbm_search_by_sim3(m_, k1.data(), KF1.descriptors->ptr<uint8_t>(0), (int)k1.size(), k2.data(), KF2.descriptors->ptr<uint8_t>(0),
                                (int)k2.size(), KF2.minX, KF2.maxX, KF2.minY, KF2.maxY, vScaleFactors.data(), (int)vScaleFactors.size(),
                                P12.xy.data(), P12.level.data(), P12.descriptors.data(), P21.xy.data(), P21.level.data(), P21.descriptors.data(),
                                th, vnMatches12.data(), &n) != ORBX_OK)
            fail("orbm_search_by_sim3");
        return n;
    }

    // Project MapPoints using a Similarity Transformation and search matches, used in loop detection (ORBmatcher.cc:290-403).
    // vbMatched[j] = vpMatched[j] != NULL on entry; vnMatch[j] = point this call stores into vpMatched[j], -1 = untouched.
    int SearchByProjection(const FrameView& KF, const std::vector<uint8_t>& vbMatched, const std::vector<float>& vScaleFactors,
                           const ProjectedPoints& P, std::vector<int>& vnMatch, int th)
    {
        std::vector<orbx_kp> k = pack(*KF.keys);
        vnMatch.assign(k.size(), -1);
        int n = 0;
        if (orbm_search_by_projection_sim3(m_, k.data(), KF.descriptors->ptr<uint8_t>(0), vbMatched.data(), (int)k.size(), KF.minX, KF.maxX, KF.minY,
                                           KF.maxY, vScaleFactors.data(), (int)vScaleFactors.size(), P.xy.data(), P.level.data(),
                                           P.descriptors.data(), (int)P.level.size(), (float)th, vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection_sim3");
        return n;
    }

    // Project MapPoints seen in a KeyFrame into the Frame and search matches, used in relocalisation (ORBmatcher.cc:1474-1601).
    // vbOccupied[j] = CurrentFrame.mvpMapPoints[j] != NULL; P.viewCosOrAngle = pKF->mvKeysUn[i].angle.
    int SearchByProjection(const FrameView& CurrentFrame, const std::vector<uint8_t>& vbOccupied, const std::vector<float>& vScaleFactors,
                           const ProjectedPoints& P, std::vector<int>& vnMatch, const float th, const int ORBdist)
    {
        std::vector<orbx_kp> k = pack(*CurrentFrame.keys);
        vnMatch.assign(k.size(), -1);
        int n = 0;
        if (orbm_search_by_projection_kf(m_, k.data(), CurrentFrame.descriptors->ptr<uint8_t>(0), vbOccupied.data(), (int)k.size(), CurrentFrame.minX,
                                         CurrentFrame.maxX, CurrentFrame.minY, CurrentFrame.maxY, vScaleFactors.data(), (int)vScaleFactors.size(),
                                         P.xy.data(), P.level.data(), P.viewCosOrAngle.data(), P.descriptors.data(), (int)P.level.size(), th, ORBdist,
                                         mbCheckOrientation, vnMatch.data(), &n) != ORBX_OK)
            fail("orbm_search_by_projection_kf");
        return n;
    }

    // What SearchForTriangulation reads from a KeyFrame besides FrameView.
    struct KeyFrameState {
        const std::vector<uint8_t>* hasMapPoint;    // pKF->GetMapPoint(idx) != NULL
        const std::vector<float>* uRight;           // KeyFrame::mvuRight
        const FeatureVectorCSR* featVec;            // KeyFrame::mFeatVec
    };
    // Matching to triangulate new MapPoints, checking the epipolar constraint (ORBmatcher.cc:657-825).  F12: 3x3 CV_32F;
    // (ex, ey): the epipole in the second image (:667-672); vScaleFactors2 / vLevelSigma2_2 of pKF2.
    int SearchForTriangulation(const FrameView& KF1, const KeyFrameState& S1, const FrameView& KF2, const KeyFrameState& S2, const cv::Mat& F12,
                               float ex, float ey, const std::vector<float>& vScaleFactors2, const std::vector<float>& vLevelSigma2_2,
                               std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo)
    {
        std::vector<orbx_kp> k1 = pack(*KF1.keys), k2 = pack(*KF2.keys);
        std::vector<int> m12(k1.size(), -1);
        float F[9];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) F[3 * r + c] = F12.ptr<float>(r)[c];
        int n = 0;
        if (orbm_search_for_triangulation(m_, k1.data(), KF1.descriptors->ptr<uint8_t>(0), (int)k1.size(), S1.hasMapPoint->data(), S1.uRight->data(),
                                          S1.featVec->node.data(), S1.featVec->off.data(), S1.featVec->idx.data(), (int)S1.featVec->node.size(),
                                          k2.data(), KF2.descriptors->ptr<uint8_t>(0), (int)k2.size(), S2.hasMapPoint->data(), S2.uRight->data(),
                                          S2.featVec->node.data(), S2.featVec->off.data(), S2.featVec->idx.data(), (int)S2.featVec->node.size(),
                                          F, ex, ey, vScaleFactors2.data(), vLevelSigma2_2.data(), (int)vScaleFactors2.size(), bOnlyStereo,
                                          mbCheckOrientation, m12.data(), &n) != ORBX_OK)
            fail("orbm_search_for_triangulation");
        vMatchedPairs.clear();
        vMatchedPairs.reserve(n);
        for (size_t i = 0; i < m12.size(); i++)                                  // :818-823
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)m12[i]));
        return n;
    }

    // MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:252-312) for a batch of MapPoints: vvDescriptors[p] = the vDescriptors the
    // reference collects for point p (:270-276).  Returns BestIdx per point (-1 when it has no descriptor); the caller then does
    // mDescriptor = vDescriptors[BestIdx].clone() (:308-311).
    std::vector<int> ComputeDistinctiveDescriptors(const std::vector<std::vector<cv::Mat> >& vvDescriptors)
    {
        std::vector<int> off(1, 0), best(vvDescriptors.size(), -1);
        std::vector<uint8_t> rows;
        for (size_t p = 0; p < vvDescriptors.size(); p++) {
            for (size_t i = 0; i < vvDescriptors[p].size(); i++)
                rows.insert(rows.end(), vvDescriptors[p][i].ptr<uint8_t>(0), vvDescriptors[p][i].ptr<uint8_t>(0) + 32);
            off.push_back((int)(rows.size() / 32));
        }
        if (orbm_distinctive_descriptors(m_, rows.data(), off.data(), (int)vvDescriptors.size(), best.data()) != ORBX_OK)
            fail("orbm_distinctive_descriptors");
        return best;
    }

    // ---- Frame glue (the per-keypoint maps of the Frame constructors).  K4 = (fx, fy, cx, cy); distCoef = mDistCoef as floats
    // (k1, k2, p1, p2[, k3]).
    // Frame::UndistortKeyPoints (Frame.cc:419-449): mvKeysUn from mvKeys.
    void UndistortKeyPoints(const std::vector<cv::KeyPoint>& vKeys, const float K4[4], const std::vector<float>& distCoef,
                            std::vector<cv::KeyPoint>& vKeysUn)
    {
        std::vector<orbx_kp> k = pack(vKeys), u(k.size());
        if (orbm_undistort_keypoints(m_, k.data(), (int)k.size(), K4[0], K4[1], K4[2], K4[3], distCoef.data(), (int)distCoef.size(), u.data()) != ORBX_OK)
            fail("orbm_undistort_keypoints");
        vKeysUn = vKeys;
        for (size_t i = 0; i < u.size(); i++) { vKeysUn[i].pt.x = u[i].x; vKeysUn[i].pt.y = u[i].y; }
    }
    // Frame::ComputeImageBounds (Frame.cc:451-479): bounds4 = (mnMinX, mnMaxX, mnMinY, mnMaxY).
    void ComputeImageBounds(int cols, int rows, const float K4[4], const std::vector<float>& distCoef, float bounds4[4])
    {
        if (orbm_image_bounds(m_, cols, rows, K4[0], K4[1], K4[2], K4[3], distCoef.data(), (int)distCoef.size(), bounds4) != ORBX_OK)
            fail("orbm_image_bounds");
    }
    // Frame::ComputeStereoFromRGBD (Frame.cc:658-679): imDepth = CV_32F image (data pointer, size, row step in bytes).
    void ComputeStereoFromRGBD(const std::vector<cv::KeyPoint>& vKeys, const std::vector<cv::KeyPoint>& vKeysUn, const float* imDepth, int cols,
                               int rows, size_t step, float bf, std::vector<float>& vuRight, std::vector<float>& vDepth)
    {
        std::vector<orbx_kp> k = pack(vKeys), u = pack(vKeysUn);
        vuRight.assign(k.size(), -1.0f);
        vDepth.assign(k.size(), -1.0f);
        if (orbm_stereo_from_rgbd(m_, k.data(), u.data(), (int)k.size(), imDepth, cols, rows, step, bf, vuRight.data(), vDepth.data()) != ORBX_OK)
            fail("orbm_stereo_from_rgbd");
    }

    // Frame::ComputeStereoMatches (Frame.cc:481-655): fills mvuRight / mvDepth, returns the surviving matches.  maxD is
    // explicit: the reference reads an uninitialised mb there (Frame.cc:511 vs :114); the intended value is fx.
    int ComputeStereoMatches(ORBextractor& left, ORBextractor& right, const std::vector<cv::KeyPoint>& keysLeft, const cv::Mat& descLeft,
                             const std::vector<cv::KeyPoint>& keysRight, const cv::Mat& descRight, float bf, float maxD,
                             std::vector<float>& mvuRight, std::vector<float>& mvDepth)
    {
        std::vector<orbx_kp> kl = pack(keysLeft), kr = pack(keysRight);
        mvuRight.assign(kl.size(), -1.0f);
        mvDepth.assign(kl.size(), -1.0f);
        int n = 0;
        if (orbm_stereo_matches(m_, left.handle(), right.handle(), 0, kl.data(), descLeft.ptr<uint8_t>(0), (int)kl.size(), kr.data(),
                                descRight.ptr<uint8_t>(0), (int)kr.size(), bf, 0.0f, maxD, mvuRight.data(), mvDepth.data(), &n) != ORBX_OK)
            fail("orbm_stereo_matches");
        return n;
    }

public:
    static const int TH_LOW = 50;          // ORBmatcher.cc:37-39
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;

protected:
    int fuse(const FrameView& KF, const std::vector<float>* vuRight, const std::vector<float>& vScaleFactors,
             const std::vector<float>* vInvLevelSigma2, const ProjectedPoints& P, std::vector<int>& vnBestIdx, float th, int stereoChi2)
    {
        std::vector<orbx_kp> k = pack(*KF.keys);
        vnBestIdx.assign(P.level.size(), -1);
        int n = 0;
        if (orbm_fuse(m_, k.data(), KF.descriptors->ptr<uint8_t>(0), vuRight ? vuRight->data() : nullptr, (int)k.size(), KF.minX, KF.maxX, KF.minY,
                      KF.maxY, vScaleFactors.data(), vInvLevelSigma2 ? vInvLevelSigma2->data() : nullptr, (int)vScaleFactors.size(), P.xy.data(),
                      P.xr.data(), P.level.data(), P.descriptors.data(), (int)P.level.size(), th, stereoChi2, vnBestIdx.data(), &n) != ORBX_OK)
            fail("orbm_fuse");
        return n;
    }
    static std::vector<orbx_kp> pack(const std::vector<cv::KeyPoint>& v)
    {
        std::vector<orbx_kp> o(v.size());
        for (size_t i = 0; i < v.size(); i++) {
            o[i].x = v[i].pt.x; o[i].y = v[i].pt.y; o[i].size = v[i].size; o[i].angle = v[i].angle;
            o[i].response = v[i].response; o[i].octave = v[i].octave; o[i].class_id = v[i].class_id;
        }
        return o;
    }
    void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + orbx_last_error()); }

    float mfNNratio;
    bool mbCheckOrientation;
    orbx_matcher* m_;
};

inline DeviceFrame::DeviceFrame(ORBmatcher& matcher, const std::vector<cv::KeyPoint>& keysUn, const cv::Mat& descriptors, float minX, float maxX,
                                float minY, float maxY)
{
    std::vector<orbx_kp> k(keysUn.size());
    for (size_t i = 0; i < k.size(); i++) {
        k[i].x = keysUn[i].pt.x; k[i].y = keysUn[i].pt.y; k[i].size = keysUn[i].size; k[i].angle = keysUn[i].angle;
        k[i].response = keysUn[i].response; k[i].octave = keysUn[i].octave; k[i].class_id = keysUn[i].class_id;
    }
    if (orbm_frame_create(matcher.handle(), k.data(), k.empty() ? nullptr : descriptors.ptr<uint8_t>(0), (int)k.size(), minX, maxX, minY, maxY, &f_) != ORBX_OK)
        throw std::runtime_error(std::string("orbm_frame_create: ") + orbx_last_error());
}

inline DeviceFrame::DeviceFrame(ORBmatcher& matcher, ORBextractor& extractor, const float K4[4], const std::vector<float>& distCoef, float minX,
                                float maxX, float minY, float maxY, int frame)
{
    if (orbm_frame_from_extractor(matcher.handle(), extractor.handle(), frame, K4[0], K4[1], K4[2], K4[3], distCoef.empty() ? nullptr : distCoef.data(),
                                  (int)distCoef.size(), minX, maxX, minY, maxY, &f_) != ORBX_OK)
        throw std::runtime_error(std::string("orbm_frame_from_extractor: ") + orbx_last_error());
}

}  // namespace ORB_SLAM2

#include "ORBmatcher_reference_api.inl"

#endif  // ORBMATCHER_H
