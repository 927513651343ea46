// ORBmatcher_reference_api.inl -- the searches of include/ORBmatcher.h under the REFERENCE'S OWN SIGNATURES
// (kimwin2/ORB_SLAM2v2-1 include/ORBmatcher.h:47-82), so that call sites such as Tracking.cc:742-743, :907-910, :1012-1034,
// :1329-1336, :1502-1522, LoopClosing.cc:249-275 compile unchanged.  Frame / KeyFrame / MapPoint are template parameters: this repository does not
// pull the SLAM object graph (Map, ROS messages) into the translation unit; the user's own headers supply the types.  Each
// overload runs the reference's host-side skip tests while flattening the objects into the arrays of the FrameView overloads,
// calls those (device windows + Hamming, host replay of the order-dependent bookkeeping), and writes the results back the way
// the reference does.  Included at the end of ORBmatcher.h; not a stand-alone header.
#ifndef ORBMATCHER_REFERENCE_API_INL
#define ORBMATCHER_REFERENCE_API_INL

namespace ORB_SLAM2
{

namespace detail
{
template <class FrameT> inline FrameView view_un(const FrameT& F)
{
    FrameView v = {&F.mvKeysUn, &F.mDescriptors, (float)F.mnMinX, (float)F.mnMaxX, (float)F.mnMinY, (float)F.mnMaxY};
    return v;
}
template <class MapPointT> inline void push_descriptor(std::vector<uint8_t>& out, MapPointT* pMP)
{
    const cv::Mat d = pMP->GetDescriptor();
    const uint8_t* p = d.ptr<uint8_t>(0);
    out.insert(out.end(), p, p + 32);
}
// F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0 (ORBmatcher.cc:88-90, :1404-1406), as it stands on entry
template <class FrameT> inline std::vector<uint8_t> occupied_with_observations(const FrameT& F)
{
    std::vector<uint8_t> o(F.mvpMapPoints.size(), 0);
    for (size_t j = 0; j < o.size(); j++) o[j] = (F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0) ? 1 : 0;
    return o;
}
// For builds WITHOUT OpenCV (with it the templates below write the reference's cv::Mat expressions themselves).
// Rows 0..2 of a 4x4 CV_32F pose, evaluated the way cv::Mat expressions evaluate them for 3x3 * 3x1 products:
// A*b + c runs through gemm's small-matrix path (float products summed left to right, then (float)(t*1.0 + c*1.0) in double);
// -A.t()*b runs through the general path (double accumulation, scaled by -1, rounded once).
struct Pose34 {
    float R[3][3], t[3];
    template <class MatT> explicit Pose34(const MatT& T)
    {
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) R[r][c] = T.template at<float>(r, c);
            t[r] = T.template at<float>(r, 3);
        }
    }
    void apply(const float x[3], float out[3]) const                  // R*x + t
    {
        for (int r = 0; r < 3; r++) {
            const float t0 = R[r][0] * x[0] + R[r][1] * x[1] + R[r][2] * x[2];
            out[r] = (float)((double)t0 * 1.0 + (double)t[r] * 1.0);
        }
    }
    void centre(float out[3]) const                                   // -R.t()*t
    {
        for (int c = 0; c < 3; c++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += (double)R[k][c] * (double)t[k];
            out[c] = (float)(s * -1.0);
        }
    }
};
}  // namespace detail

// ORBmatcher.cc:45-129 -- TrackLocalMap (call site Tracking.cc:1329-1336: matcher.SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th))
template <class FrameT, class MapPointT>
int ORBmatcher::SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th)
{
    ProjectedPoints P;
    std::vector<MapPointT*> who;
    for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
        MapPointT* pMP = vpMapPoints[iMP];
        if (!pMP->mbTrackInView) continue;                             // :52-53
        if (pMP->isBad()) continue;                                    // :55-56
        P.xy.push_back(pMP->mTrackProjX);
        P.xy.push_back(pMP->mTrackProjY);
        P.xr.push_back(pMP->mTrackProjXR);
        P.level.push_back(pMP->mnTrackScaleLevel);
        P.viewCosOrAngle.push_back(pMP->mTrackViewCos);
        detail::push_descriptor(P.descriptors, pMP);
        P.hasObservations.push_back(pMP->Observations() > 0 ? 1 : 0);
        who.push_back(pMP);
    }
    const std::vector<uint8_t> occupied = detail::occupied_with_observations(F);
    const FrameState S = {&F.mvuRight, &occupied, &F.mvScaleFactors};
    std::vector<int> vnMatch;
    const int nmatches = SearchByProjection(detail::view_un(F), S, P, vnMatch, th);
    for (size_t j = 0; j < vnMatch.size(); j++)
        if (vnMatch[j] >= 0) F.mvpMapPoints[j] = who[vnMatch[j]];     // :123
    return nmatches;
}

// ORBmatcher.cc:1330-1472 -- TrackWithMotionModel (call sites Tracking.cc:1028, :1034: matcher.SearchByProjection(mCurrentFrame, mLastFrame, th, mSensor == System::MONOCULAR))
template <class FrameT>
int ORBmatcher::SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono)
{
#ifdef ORBX_USE_OPENCV
    // with OpenCV the reference's own expressions (:1340-1348, :1362-1363), so that every float is what the installed OpenCV makes of them
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat tlcM = Rlw * twc + tlw;
    const float tlc[3] = {tlcM.template at<float>(0), tlcM.template at<float>(1), tlcM.template at<float>(2)};
#else
    const detail::Pose34 Tcw(CurrentFrame.mTcw), Tlw(LastFrame.mTcw);
    float twc[3], tlc[3];
    Tcw.centre(twc);                                                   // :1343
    Tlw.apply(twc, tlc);                                               // :1348
#endif
    const bool bForward = tlc[2] > CurrentFrame.mb && !bMono;          // :1350-1351
    const bool bBackward = -tlc[2] > CurrentFrame.mb && !bMono;
    ProjectedPoints P;
    typedef typename std::remove_reference<decltype(*LastFrame.mvpMapPoints[0])>::type MapPointT;
    std::vector<MapPointT*> who;
    for (int i = 0; i < LastFrame.N; i++) {
        MapPointT* pMP = LastFrame.mvpMapPoints[i];
        if (!pMP) continue;
        if (LastFrame.mvbOutlier[i]) continue;
        const cv::Mat x3Dw = pMP->GetWorldPos();                       // 3x1 CV_32F
#ifdef ORBX_USE_OPENCV
        const cv::Mat x3DcM = Rcw * x3Dw + tcw;                        // :1363
        const float x3Dc[3] = {x3DcM.template at<float>(0), x3DcM.template at<float>(1), x3DcM.template at<float>(2)};
#else
        const float xw[3] = {x3Dw.at<float>(0, 0), x3Dw.at<float>(1, 0), x3Dw.at<float>(2, 0)};
        float x3Dc[3];
        Tcw.apply(xw, x3Dc);                                           // :1363
#endif
        const float xc = x3Dc[0];
        const float yc = x3Dc[1];
        const float invzc = 1.0 / x3Dc[2];
        if (invzc < 0) continue;
        float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;      // :1372-1373
        float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const float ur = u - CurrentFrame.mbf * invzc;                 // :1411
        P.xy.push_back(u);
        P.xy.push_back(v);
        P.xr.push_back(ur);
        P.level.push_back(LastFrame.mvKeys[i].octave);                 // :1380
        P.viewCosOrAngle.push_back(LastFrame.mvKeysUn[i].angle);       // :1436
        detail::push_descriptor(P.descriptors, pMP);
        P.hasObservations.push_back(pMP->Observations() > 0 ? 1 : 0);
        who.push_back(pMP);
    }
    const std::vector<uint8_t> occupied = detail::occupied_with_observations(CurrentFrame);
    const FrameState S = {&CurrentFrame.mvuRight, &occupied, &CurrentFrame.mvScaleFactors};
    std::vector<int> vnMatch;
    const int nmatches = SearchByProjection(detail::view_un(CurrentFrame), S, P, vnMatch, th, bForward ? 1 : (bBackward ? 2 : 0));
    for (size_t j = 0; j < vnMatch.size(); j++)
        if (vnMatch[j] >= 0) CurrentFrame.mvpMapPoints[j] = who[vnMatch[j]];      // :1430, after the rotation check :1448-1468
    return nmatches;
}

// ORBmatcher.cc:159-288 -- TrackReferenceKeyFrame / Relocalization (call sites Tracking.cc:907-910, :1502-1522)
template <class KeyFrameT, class FrameT, class MapPointT>
int ORBmatcher::SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches)
{
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(NULL));
    std::vector<uint8_t> valid(vpMapPointsKF.size(), 0);
    for (size_t i = 0; i < valid.size(); i++) valid[i] = (vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad()) ? 1 : 0;     // :190-196
    const FeatureVectorCSR fvKF = FeatureVectorCSR::from(pKF->mFeatVec), fvF = FeatureVectorCSR::from(F.mFeatVec);
    const FrameView KFv = detail::view_un(*pKF);
    // the rotation histogram reads F.mvKeys[bestIdxF].angle (:238), not mvKeysUn
    const FrameView Fv = {&F.mvKeys, &F.mDescriptors, (float)F.mnMinX, (float)F.mnMaxX, (float)F.mnMinY, (float)F.mnMaxY};
    std::vector<int> vnMatchF;
    const int nmatches = SearchByBoW(KFv, valid, fvKF, Fv, fvF, vnMatchF);
    for (size_t j = 0; j < vnMatchF.size(); j++)
        if (vnMatchF[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[vnMatchF[j]];   // :232
    return nmatches;
}

// ORBmatcher.cc:522-655 -- loop closing (call site LoopClosing.cc:249-275)
template <class KeyFrameT, class MapPointT>
int ORBmatcher::SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12)
{
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(NULL));
    std::vector<uint8_t> valid1(vpMapPoints1.size(), 0), valid2(vpMapPoints2.size(), 0);
    for (size_t i = 0; i < valid1.size(); i++) valid1[i] = (vpMapPoints1[i] && !vpMapPoints1[i]->isBad()) ? 1 : 0;     // :555-560
    for (size_t i = 0; i < valid2.size(); i++) valid2[i] = (vpMapPoints2[i] && !vpMapPoints2[i]->isBad()) ? 1 : 0;     // :571-580
    const FeatureVectorCSR fv1 = FeatureVectorCSR::from(pKF1->mFeatVec), fv2 = FeatureVectorCSR::from(pKF2->mFeatVec);
    std::vector<int> vnMatches12;
    const int nmatches = SearchByBoW(detail::view_un(*pKF1), valid1, fv1, detail::view_un(*pKF2), valid2, fv2, vnMatches12);
    for (size_t i = 0; i < vnMatches12.size(); i++)
        if (vnMatches12[i] >= 0) vpMatches12[i] = vpMapPoints2[vnMatches12[i]];    // :602
    return nmatches;
}

// ORBmatcher.cc:405-520 -- monocular initialisation (call site Tracking.cc:742-743)
template <class FrameT, class>
int ORBmatcher::SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                        int windowSize)
{
    return SearchForInitialization(detail::view_un(F1), detail::view_un(F2), vbPrevMatched, vnMatches12, windowSize);
}

#ifdef ORBX_USE_OPENCV
namespace detail
{
// KeyFrame side of FrameView
template <class KeyFrameT> inline FrameView view_kf(const KeyFrameT& K)
{
    FrameView v = {&K.mvKeysUn, &K.mDescriptors, (float)K.mnMinX, (float)K.mnMaxX, (float)K.mnMinY, (float)K.mnMaxY};
    return v;
}
inline void push_projected(ORBmatcher::ProjectedPoints& P, float u, float v, float ur, int level, float angle, const cv::Mat& d, bool hasObs)
{
    P.xy.push_back(u); P.xy.push_back(v); P.xr.push_back(ur); P.level.push_back(level); P.viewCosOrAngle.push_back(angle);
    P.descriptors.insert(P.descriptors.end(), d.ptr<uint8_t>(0), d.ptr<uint8_t>(0) + 32);
    P.hasObservations.push_back(hasObs ? 1 : 0);
}
// ORBmatcher.cc:850-889 / :1011-1056 / :321-362: projection of a MapPoint into a key frame with the reference's own skip tests
// (depth sign, IsInImage, distance range, viewing angle) and PredictScale.  false where the reference `continue`s.
template <class KeyFrameT, class MapPointT>
inline bool project_into_keyframe(KeyFrameT* pKF, const cv::Mat& Rcw, const cv::Mat& tcw, const cv::Mat& Ow, MapPointT* pMP, bool invzInDouble, float& u,
                                  float& v, float& invz, int& level)
{
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.template at<float>(2) < 0.0f) return false;
    invz = invzInDouble ? (float)(1.0 / p3Dc.template at<float>(2)) : 1 / p3Dc.template at<float>(2);      // `1.0/z` (:1022) resp. `1/z` (:861, :330)
    const float x = p3Dc.template at<float>(0) * invz, y = p3Dc.template at<float>(1) * invz;
    u = pKF->fx * x + pKF->cx;
    v = pKF->fy * y + pKF->cy;
    if (!pKF->IsInImage(u, v)) return false;
    const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) return false;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist3D) return false;
    level = pMP->PredictScale(dist3D, pKF);
    return true;
}
// one direction of SearchBySim3 (:1145-1183 resp. :1225-1263): one entry per feature of the source key frame, level -1 where skipped
template <class KeyFrameT, class MapPointT>
inline void project_sim3(KeyFrameT* pKFintr, KeyFrameT* pKFdst, const std::vector<MapPointT*>& vpSrc, const std::vector<bool>& matched, const cv::Mat& Rsw,
                         const cv::Mat& tsw, const cv::Mat& sRds, const cv::Mat& tds, ORBmatcher::ProjectedPoints& P)
{
    const cv::Mat none = cv::Mat::zeros(1, 32, CV_8U);
    for (size_t i = 0; i < vpSrc.size(); i++) {
        MapPointT* pMP = vpSrc[i];
        float u = 0, v = 0;
        int level = -1;
        if (pMP && !matched[i] && !pMP->isBad()) {
            cv::Mat p3Dw = pMP->GetWorldPos();
            cv::Mat p3Dcs = Rsw * p3Dw + tsw;
            cv::Mat p3Dcd = sRds * p3Dcs + tds;
            if (!(p3Dcd.template at<float>(2) < 0.0)) {
                const float invz = 1.0 / p3Dcd.template at<float>(2);
                const float x = p3Dcd.template at<float>(0) * invz, y = p3Dcd.template at<float>(1) * invz;
                u = pKFintr->fx * x + pKFintr->cx;                       // fx .. cy of pKF1 in both directions (:1107-1110)
                v = pKFintr->fy * y + pKFintr->cy;
                if (pKFdst->IsInImage(u, v)) {
                    const float dist3D = cv::norm(p3Dcd);
                    if (!(dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance())) level = pMP->PredictScale(dist3D, pKFdst);
                }
            }
        }
        push_projected(P, u, v, -1.f, level, 0.f, level >= 0 ? pMP->GetDescriptor() : none, true);
    }
}
}  // namespace detail

// ORBmatcher.cc:827-977 -- LocalMapping::SearchInNeighbors (call sites LocalMapping.cc:497, :522)
template <class KeyFrameT, class MapPointT> int ORBmatcher::Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th)
{
    cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation(), Ow = pKF->GetCameraCenter();
    ProjectedPoints P;
    std::vector<int> who;
    for (size_t i = 0; i < vpMapPoints.size(); i++) {
        MapPointT* pMP = vpMapPoints[i];
        if (!pMP) continue;                                               // :848-849
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;              // :851-852
        float u, v, invz;
        int level;
        if (!detail::project_into_keyframe(pKF, Rcw, tcw, Ow, pMP, false, u, v, invz, level)) continue;
        detail::push_projected(P, u, v, u - pKF->mbf * invz, level, 0.f, pMP->GetDescriptor(), true);      // ur = u - bf*invz (:872)
        who.push_back((int)i);
    }
    std::vector<int> best;
    Fuse(detail::view_kf(*pKF), pKF->mvuRight, pKF->mvScaleFactors, pKF->mvInvLevelSigma2, P, best, th);
    int nFused = 0;
    for (size_t j = 0; j < best.size(); j++) {                             // :937-958, in the reference's visiting order
        if (best[j] < 0) continue;
        MapPointT* pMP = vpMapPoints[who[j]];
        // the reference evaluates :851-852 when it reaches the point, i.e. after the Replace / AddObservation calls of the points before
        // it (a point listed twice, or one that an earlier iteration replaced): evaluated again here, at the same moment
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        MapPointT* pMPinKF = pKF->GetMapPoint(best[j]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, best[j]);
            pKF->AddMapPoint(pMP, best[j]);
        }
        nFused++;
    }
    return nFused;
}

// ORBmatcher.cc:979-1102 -- LoopClosing::SearchAndFuse (call site LoopClosing.cc:608)
template <class KeyFrameT, class MapPointT>
int ORBmatcher::Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint)
{
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);                     // :988-992
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
    ProjectedPoints P;
    std::vector<int> who;
    for (size_t i = 0; i < vpPoints.size(); i++) {
        MapPointT* pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;          // :1007-1008
        float u, v, invz;
        int level;
        if (!detail::project_into_keyframe(pKF, Rcw, tcw, Ow, pMP, true, u, v, invz, level)) continue;
        detail::push_projected(P, u, v, -1.f, level, 0.f, pMP->GetDescriptor(), true);
        who.push_back((int)i);
    }
    std::vector<int> best;
    Fuse(detail::view_kf(*pKF), pKF->mvScaleFactors, P, best, th);
    int nFused = 0;
    for (size_t j = 0; j < best.size(); j++) {                             // :1080-1096
        if (best[j] < 0) continue;
        MapPointT* pMP = vpPoints[who[j]];
        MapPointT* pMPinKF = pKF->GetMapPoint(best[j]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[who[j]] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, best[j]);
            pKF->AddMapPoint(pMP, best[j]);
        }
        nFused++;
    }
    return nFused;
}

// ORBmatcher.cc:290-403 -- LoopClosing::ComputeSim3 (call site LoopClosing.cc:385)
template <class KeyFrameT, class MapPointT>
int ORBmatcher::SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th)
{
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);                     // :299-303
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPointT*>(NULL));
    ProjectedPoints P;
    std::vector<int> who;
    for (size_t i = 0; i < vpPoints.size(); i++) {
        MapPointT* pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;          // :316-317
        float u, v, invz;
        int level;
        if (!detail::project_into_keyframe(pKF, Rcw, tcw, Ow, pMP, false, u, v, invz, level)) continue;
        detail::push_projected(P, u, v, -1.f, level, 0.f, pMP->GetDescriptor(), true);
        who.push_back((int)i);
    }
    std::vector<uint8_t> vbMatched(vpMatched.size());
    for (size_t j = 0; j < vpMatched.size(); j++) vbMatched[j] = vpMatched[j] != NULL;
    std::vector<int> vnMatch;
    const int nmatches = SearchByProjection(detail::view_kf(*pKF), vbMatched, pKF->mvScaleFactors, P, vnMatch, th);
    for (size_t j = 0; j < vnMatch.size(); j++)
        if (vnMatch[j] >= 0) vpMatched[j] = vpPoints[who[vnMatch[j]]];     // :396
    return nmatches;
}

// ORBmatcher.cc:1474-1601 -- Tracking::Relocalization (call sites Tracking.cc:1597, :1611)
template <class FrameT, class KeyFrameT, class MapPointT>
int ORBmatcher::SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th, const int ORBdist)
{
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);   // :1478-1480
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat Ow = -Rcw.t() * tcw;
    const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
    ProjectedPoints P;
    std::vector<int> who;
    for (size_t i = 0; i < vpMPs.size(); i++) {
        MapPointT* pMP = vpMPs[i];
        if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;    // :1493-1497
        cv::Mat x3Dw = pMP->GetWorldPos();                                 // :1499-1531
        cv::Mat x3Dc = Rcw * x3Dw + tcw;
        const float xc = x3Dc.template at<float>(0), yc = x3Dc.template at<float>(1);
        const float invzc = 1.0 / x3Dc.template at<float>(2);
        const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx, v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        cv::Mat PO = x3Dw - Ow;
        float dist3D = cv::norm(PO);
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
        detail::push_projected(P, u, v, -1.f, pMP->PredictScale(dist3D, &CurrentFrame), pKF->mvKeysUn[i].angle, pMP->GetDescriptor(), true);
        who.push_back((int)i);
    }
    std::vector<uint8_t> occupied(CurrentFrame.mvpMapPoints.size());
    for (size_t j = 0; j < occupied.size(); j++) occupied[j] = CurrentFrame.mvpMapPoints[j] != NULL;       // :1546-1547
    std::vector<int> vnMatch;
    const int nmatches = SearchByProjection(detail::view_un(CurrentFrame), occupied, CurrentFrame.mvScaleFactors, P, vnMatch, th, ORBdist);
    for (size_t j = 0; j < vnMatch.size(); j++)
        if (vnMatch[j] >= 0) CurrentFrame.mvpMapPoints[j] = vpMPs[who[vnMatch[j]]];                         // :1560, after the rotation check :1579-1597
    return nmatches;
}

// ORBmatcher.cc:1104-1328 -- LoopClosing::ComputeSim3 (call site LoopClosing.cc:333)
template <class KeyFrameT, class MapPointT>
int ORBmatcher::SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12,
                             const float th)
{
    cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();      // :1113-1123
    cv::Mat sR12 = s12 * R12;
    cv::Mat sR21 = (1.0 / s12) * R12.t();
    cv::Mat t21 = -sR21 * t12;
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {                                         // :1133-1144
        MapPointT* pMP = vpMatches12[i];
        if (pMP) {
            vbAlreadyMatched1[i] = true;
            const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    ProjectedPoints P12, P21;
    detail::project_sim3(pKF1, pKF2, vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, P12);
    detail::project_sim3(pKF1, pKF1, vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, P21);
    std::vector<int> m12;
    const int nFound = SearchBySim3(detail::view_kf(*pKF1), detail::view_kf(*pKF2), pKF2->mvScaleFactors, P12, P21, m12, th);
    for (int i1 = 0; i1 < N1; i1++)
        if (m12[i1] >= 0) vpMatches12[i1] = vpMapPoints2[m12[i1]];         // :1318
    return nFound;
}

// ORBmatcher.cc:657-825 -- LocalMapping::CreateNewMapPoints (call site LocalMapping.cc:276)
template <class KeyFrameT>
int ORBmatcher::SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                                       const bool bOnlyStereo)
{
    cv::Mat Cw = pKF1->GetCameraCenter();                                  // :664-672: the epipole in the second image
    cv::Mat R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    cv::Mat C2 = R2w * Cw + t2w;
    const float invz = 1.0f / C2.template at<float>(2);
    const float ex = pKF2->fx * C2.template at<float>(0) * invz + pKF2->cx, ey = pKF2->fy * C2.template at<float>(1) * invz + pKF2->cy;
    std::vector<uint8_t> has1(pKF1->N), has2(pKF2->N);
    for (int i = 0; i < pKF1->N; i++) has1[i] = pKF1->GetMapPoint(i) != NULL;      // :698, :721
    for (int i = 0; i < pKF2->N; i++) has2[i] = pKF2->GetMapPoint(i) != NULL;
    const FeatureVectorCSR fv1 = FeatureVectorCSR::from(pKF1->mFeatVec), fv2 = FeatureVectorCSR::from(pKF2->mFeatVec);
    KeyFrameState S1 = {&has1, &pKF1->mvuRight, &fv1}, S2 = {&has2, &pKF2->mvuRight, &fv2};
    return SearchForTriangulation(detail::view_kf(*pKF1), S1, detail::view_kf(*pKF2), S2, F12, ex, ey, pKF2->mvScaleFactors, pKF2->mvLevelSigma2, vMatchedPairs,
                                  bOnlyStereo);
}
#endif  // ORBX_USE_OPENCV

}  // namespace ORB_SLAM2

#endif
