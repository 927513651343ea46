// ORBmatcher.h -- drop-in for the descriptor-level part of the reference's include/ORBmatcher.h
// (kimwin2/ORB_SLAM2v2-1 include/ORBmatcher.h:37-102, src/ORBmatcher.cc) and of Frame::ComputeStereoMatches
// (src/Frame.cc:481-655) over the C ABI of liborbx.so (include/orbx.h).  Same class name, constructor, constants and
// return conventions.  The searches take a FrameView (the members of Frame/KeyFrame that the routines read) instead
// of Frame&, because Frame drags the whole SLAM object graph (Map, MapPoint, ROS messages) into the translation
// unit; INTEGRATION.md shows the three-line adapter from the reference's Frame.  Predicates that read map state
// (isBad, Observations, epipolar checks) stay on the host side of the caller, as in SURVEY.md section 8(a) a15.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <stdexcept>
#include <string>
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

class ORBmatcher
{
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0) : mfNNratio(nnratio), mbCheckOrientation(checkOri), m_(nullptr)
    {
        if (orbm_create(device, &m_) != ORBX_OK) fail("orbm_create");
    }
    ~ORBmatcher() { orbm_destroy(m_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // Computes the Hamming distance between two ORB descriptors (ORBmatcher.cc:1649-1665).  One pair per call costs a
    // kernel launch; DescriptorDistances() is the batched form for callers such as MapPoint::ComputeDistinctiveDescriptors.
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
    {
        int d = 0;
        if (orbm_descriptor_distance(shared(), a.ptr<uint8_t>(0), b.ptr<uint8_t>(0), 1, &d) != ORBX_OK)
            throw std::runtime_error(std::string("orbm_descriptor_distance: ") + orbx_last_error());
        return d;
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
    static std::vector<orbx_kp> pack(const std::vector<cv::KeyPoint>& v)
    {
        std::vector<orbx_kp> o(v.size());
        for (size_t i = 0; i < v.size(); i++) {
            o[i].x = v[i].pt.x; o[i].y = v[i].pt.y; o[i].size = v[i].size; o[i].angle = v[i].angle;
            o[i].response = v[i].response; o[i].octave = v[i].octave; o[i].class_id = v[i].class_id;
        }
        return o;
    }
    static orbx_matcher* shared()
    {
        static thread_local struct Holder {
            orbx_matcher* m = nullptr;
            ~Holder() { orbm_destroy(m); }
        } h;
        if (!h.m && orbm_create(0, &h.m) != ORBX_OK) throw std::runtime_error(std::string("orbm_create: ") + orbx_last_error());
        return h.m;
    }
    void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + orbx_last_error()); }

    float mfNNratio;
    bool mbCheckOrientation;
    orbx_matcher* m_;
};

}  // namespace ORB_SLAM2

#endif  // ORBMATCHER_H
