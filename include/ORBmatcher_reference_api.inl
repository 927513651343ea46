// ORBmatcher_reference_api.inl -- the searches of include/ORBmatcher.h under the REFERENCE'S OWN SIGNATURES
// (kimwin2/ORB_SLAM2v2-1 include/ORBmatcher.h:47-82), so that call sites such as Tracking.cc:742-743, :907-911, :1012-1016,
// :1336, LoopClosing.cc:249 compile unchanged.  Frame / KeyFrame / MapPoint are template parameters: this repository does not
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

// ORBmatcher.cc:45-129 -- TrackLocalMap (call site Tracking.cc:1112-1116: matcher.SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th))
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

// ORBmatcher.cc:1330-1472 -- TrackWithMotionModel (call site Tracking.cc:907: matcher.SearchByProjection(mCurrentFrame, mLastFrame, th, mSensor == System::MONOCULAR))
template <class FrameT>
int ORBmatcher::SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono)
{
    const detail::Pose34 Tcw(CurrentFrame.mTcw), Tlw(LastFrame.mTcw);
    float twc[3], tlc[3];
    Tcw.centre(twc);                                                   // :1343
    Tlw.apply(twc, tlc);                                               // :1348
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
        const float xw[3] = {x3Dw.at<float>(0, 0), x3Dw.at<float>(1, 0), x3Dw.at<float>(2, 0)};
        float x3Dc[3];
        Tcw.apply(xw, x3Dc);                                           // :1363
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

// ORBmatcher.cc:159-288 -- TrackReferenceKeyFrame / Relocalization (call sites Tracking.cc:742-743, :1329-1336)
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

// ORBmatcher.cc:522-655 -- loop closing (call site LoopClosing.cc:249)
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

// ORBmatcher.cc:405-520 -- monocular initialisation (call site Tracking.cc:598)
template <class FrameT, class>
int ORBmatcher::SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                        int windowSize)
{
    return SearchForInitialization(detail::view_un(F1), detail::view_un(F2), vbPrevMatched, vnMatches12, windowSize);
}

}  // namespace ORB_SLAM2

#endif
