// The reference's own call lines against include/ORBmatcher.h, on stand-in Frame / KeyFrame / MapPoint types
// (tests/native/slam_stubs.h): each reference-signature overload must give what the FrameView overload gives on the same
// inputs flattened by hand.  usage: reference_api_test a.raw b.raw w h nfeatures   (two views of one scene)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "slam_stubs.h"

using namespace ORB_SLAM2;

static std::vector<uint8_t> read_raw(const char* path, size_t n)
{
    std::vector<uint8_t> v(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(3); }
    fclose(f);
    return v;
}

static void fill_frame(Frame& F, ORBextractor& ex, std::vector<uint8_t>& img, int w, int h)
{
    cv::Mat im(h, w, CV_8UC1, img.data());
    ex(im, cv::Mat(), F.mvKeys, F.mDescriptors);
    F.N = (int)F.mvKeys.size();
    F.mvKeysUn = F.mvKeys;
    F.mvuRight.assign(F.N, -1.0f);
    for (int i = 0; i < F.N; i += 3) F.mvuRight[i] = F.mvKeys[i].pt.x - 10.0f;
    F.mvpMapPoints.assign(F.N, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign(F.N, false);
    F.mvScaleFactors = ex.GetScaleFactors();
    F.fx = 500.f; F.fy = 500.f; F.cx = w * 0.5f; F.cy = h * 0.5f; F.mbf = 40.f; F.mb = 0.08f;
    F.mTcw = cv::Mat(4, 4, CV_32F);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) F.mTcw.at<float>(r, c) = r == c ? 1.f : 0.f;
    for (int i = 0; i < F.N; i++) F.mFeatVec[F.mDescriptors.ptr<uint8_t>(i)[0] >> 3].push_back(i);
}

static int fails = 0;
#define CHECK(cond, what) do { if (!(cond)) { fprintf(stderr, "FAIL: %s (%s:%d)\n", what, __FILE__, __LINE__); fails++; } } while (0)

static int run(int argc, char** argv);

int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    try {
        return run(argc, argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}

static int run(int, char** argv)
{
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    std::vector<uint8_t> ia = read_raw(argv[1], (size_t)w * h), ib = read_raw(argv[2], (size_t)w * h);
    ORBextractor ex(nf, 1.2f, 8, 20, 7);
    Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
    Frame mLastFrame, mCurrentFrame;
    fill_frame(mLastFrame, ex, ia, w, h);
    fill_frame(mCurrentFrame, ex, ib, w, h);
    // one MapPoint per feature of the last frame, at a depth that projects it back onto itself under the identity pose
    std::vector<MapPoint> store(mLastFrame.N);
    std::vector<MapPoint*> mvpLocalMapPoints;
    for (int i = 0; i < mLastFrame.N; i++) {
        MapPoint& p = store[i];
        const cv::KeyPoint& kp = mLastFrame.mvKeysUn[i];
        const float z = 2.0f + 0.5f * (i % 7);
        p.mWorldPos = cv::Mat(3, 1, CV_32F);
        p.mWorldPos.at<float>(0, 0) = (kp.pt.x - mLastFrame.cx) / mLastFrame.fx * z;
        p.mWorldPos.at<float>(1, 0) = (kp.pt.y - mLastFrame.cy) / mLastFrame.fy * z;
        p.mWorldPos.at<float>(2, 0) = z;
        p.mDescriptor = mLastFrame.mDescriptors.row(i).clone();
        p.mTrackProjX = kp.pt.x + (float)(i % 5) - 2.0f;
        p.mTrackProjY = kp.pt.y + (float)(i % 3) - 1.0f;
        p.mTrackProjXR = p.mTrackProjX - 10.0f;
        p.mbTrackInView = i % 6 != 0;
        p.mnTrackScaleLevel = kp.octave;
        p.mTrackViewCos = i % 2 ? 0.9995f : 0.99f;
        p.bad = i % 11 == 0;
        p.nObs = i % 4;
        mvpLocalMapPoints.push_back(&p);
        mLastFrame.mvpMapPoints[i] = i % 9 == 0 ? static_cast<MapPoint*>(NULL) : &p;
        mLastFrame.mvbOutlier[i] = i % 13 == 0;
    }
    const std::vector<MapPoint*> none(mCurrentFrame.N, static_cast<MapPoint*>(NULL));

    // ---- Tracking.cc:1329-1336 (SearchLocalPoints): ORBmatcher matcher(0.8); matcher.SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th);
    int nA = 0;
    {
        ORBmatcher matcher(0.8);
        int th = 3;
        mCurrentFrame.mvpMapPoints = none;
        nA = matcher.SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th);
        // the same by hand over the FrameView overload
        ORBmatcher::ProjectedPoints P;
        std::vector<MapPoint*> who;
        for (size_t i = 0; i < mvpLocalMapPoints.size(); i++) {
            MapPoint* pMP = mvpLocalMapPoints[i];
            if (!pMP->mbTrackInView || pMP->isBad()) continue;
            P.xy.push_back(pMP->mTrackProjX); P.xy.push_back(pMP->mTrackProjY); P.xr.push_back(pMP->mTrackProjXR);
            P.level.push_back(pMP->mnTrackScaleLevel); P.viewCosOrAngle.push_back(pMP->mTrackViewCos);
            P.descriptors.insert(P.descriptors.end(), pMP->mDescriptor.ptr<uint8_t>(0), pMP->mDescriptor.ptr<uint8_t>(0) + 32);
            P.hasObservations.push_back(pMP->Observations() > 0);
            who.push_back(pMP);
        }
        std::vector<uint8_t> occ(mCurrentFrame.N, 0);
        FrameView V = {&mCurrentFrame.mvKeysUn, &mCurrentFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        ORBmatcher::FrameState S = {&mCurrentFrame.mvuRight, &occ, &mCurrentFrame.mvScaleFactors};
        std::vector<int> vn;
        const int n2 = matcher.SearchByProjection(V, S, P, vn, (float)th);
        CHECK(n2 == nA, "SearchByProjection(Frame, MapPoints): count");
        for (int j = 0; j < mCurrentFrame.N; j++)
            CHECK(mCurrentFrame.mvpMapPoints[j] == (vn[j] >= 0 ? who[vn[j]] : static_cast<MapPoint*>(NULL)), "SearchByProjection(Frame, MapPoints): F.mvpMapPoints");
    }

    // ---- Tracking.cc:1012-1028 (TrackWithMotionModel): ORBmatcher matcher(0.9,true); fill(NULL); matcher.SearchByProjection(mCurrentFrame,mLastFrame,th,mono)
    int nB[3] = {0, 0, 0};
    for (int k = 0; k < 3; k++) {
        ORBmatcher matcher(0.9, true);
        // k = 0: camera moved forward (bForward), 1: backward, 2: monocular
        mCurrentFrame.mTcw.at<float>(2, 3) = k == 1 ? 0.2f : -0.2f;
        mCurrentFrame.mTcw.at<float>(0, 3) = 0.01f;
        const bool mono = k == 2;
        int th = 15;
        mCurrentFrame.mvpMapPoints = none;
        nB[k] = matcher.SearchByProjection(mCurrentFrame, mLastFrame, th, mono);
        std::vector<MapPoint*> got = mCurrentFrame.mvpMapPoints;
        mCurrentFrame.mvpMapPoints = none;
        const int again = matcher.SearchByProjection(mCurrentFrame, mLastFrame, th, mono);
        CHECK(again == nB[k] && got == mCurrentFrame.mvpMapPoints, "SearchByProjection(Current, Last): repeatable");
        int cnt = 0;
        for (int j = 0; j < mCurrentFrame.N; j++) cnt += got[j] != NULL;
        // a feature taken by a point without observations is not "occupied" (ORBmatcher.cc:1404-1406): a later point may take it
        // again, nmatches counts both (:1430-1431) -- so the count bounds the assigned pointers from above
        CHECK(cnt <= nB[k] && 2 * cnt > nB[k], "SearchByProjection(Current, Last): assigned pointers against the count");
    }
    CHECK(nB[0] != nB[2] || nB[1] != nB[2], "forward / backward level windows differ from the monocular one");

    // ---- Tracking.cc:907-910 (TrackReferenceKeyFrame): ORBmatcher matcher(0.7,true); vector<MapPoint*> vpMapPointMatches;
    //      int nmatches = matcher.SearchByBoW(mpReferenceKF,mCurrentFrame,vpMapPointMatches);
    KeyFrame kfA, kfB;
    KeyFrame* mpReferenceKF = &kfA;
    kfA.N = mLastFrame.N; kfA.mvKeysUn = mLastFrame.mvKeysUn; kfA.mDescriptors = mLastFrame.mDescriptors; kfA.mFeatVec = mLastFrame.mFeatVec;
    kfA.mnMaxX = w; kfA.mnMaxY = h; kfA.mvpMapPoints = mLastFrame.mvpMapPoints;
    int nC = 0;
    {
        ORBmatcher matcher(0.7, true);
        std::vector<MapPoint*> vpMapPointMatches;
        int nmatches = matcher.SearchByBoW(mpReferenceKF, mCurrentFrame, vpMapPointMatches);
        nC = nmatches;
        std::vector<uint8_t> valid(kfA.N);
        for (int i = 0; i < kfA.N; i++) valid[i] = kfA.mvpMapPoints[i] && !kfA.mvpMapPoints[i]->isBad();
        FrameView V1 = {&kfA.mvKeysUn, &kfA.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        FrameView V2 = {&mCurrentFrame.mvKeys, &mCurrentFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        std::vector<int> vn;
        const int n2 = matcher.SearchByBoW(V1, valid, ORBmatcher::FeatureVectorCSR::from(kfA.mFeatVec), V2,
                                           ORBmatcher::FeatureVectorCSR::from(mCurrentFrame.mFeatVec), vn);
        CHECK(n2 == nmatches && (int)vpMapPointMatches.size() == mCurrentFrame.N, "SearchByBoW(KF, F): count");
        for (int j = 0; j < mCurrentFrame.N; j++)
            CHECK(vpMapPointMatches[j] == (vn[j] >= 0 ? kfA.mvpMapPoints[vn[j]] : static_cast<MapPoint*>(NULL)), "SearchByBoW(KF, F): vpMapPointMatches");
    }

    // ---- LoopClosing.cc:275: int nmatches = matcher.SearchByBoW(mpCurrentKF,pKF,vvpMapPointMatches[i]);
    std::vector<MapPoint> storeB(mCurrentFrame.N);
    kfB.N = mCurrentFrame.N; kfB.mvKeysUn = mCurrentFrame.mvKeysUn; kfB.mDescriptors = mCurrentFrame.mDescriptors; kfB.mFeatVec = mCurrentFrame.mFeatVec;
    kfB.mnMaxX = w; kfB.mnMaxY = h;
    for (int i = 0; i < kfB.N; i++) { storeB[i].bad = i % 10 == 0; kfB.mvpMapPoints.push_back(i % 7 == 0 ? static_cast<MapPoint*>(NULL) : &storeB[i]); }
    int nD = 0;
    {
        ORBmatcher matcher(0.75, true);
        KeyFrame* mpCurrentKF = &kfA;
        KeyFrame* pKF = &kfB;
        std::vector<std::vector<MapPoint*> > vvpMapPointMatches(1);
        int i = 0;
        int nmatches = matcher.SearchByBoW(mpCurrentKF, pKF, vvpMapPointMatches[i]);
        nD = nmatches;
        std::vector<uint8_t> v1(kfA.N), v2(kfB.N);
        for (int k = 0; k < kfA.N; k++) v1[k] = kfA.mvpMapPoints[k] && !kfA.mvpMapPoints[k]->isBad();
        for (int k = 0; k < kfB.N; k++) v2[k] = kfB.mvpMapPoints[k] && !kfB.mvpMapPoints[k]->isBad();
        FrameView V1 = {&kfA.mvKeysUn, &kfA.mDescriptors, 0.f, (float)w, 0.f, (float)h}, V2 = {&kfB.mvKeysUn, &kfB.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        std::vector<int> vn;
        const int n2 = matcher.SearchByBoW(V1, v1, ORBmatcher::FeatureVectorCSR::from(kfA.mFeatVec), V2, v2, ORBmatcher::FeatureVectorCSR::from(kfB.mFeatVec), vn);
        CHECK(n2 == nmatches, "SearchByBoW(KF, KF): count");
        for (int k = 0; k < kfA.N; k++)
            CHECK(vvpMapPointMatches[i][k] == (vn[k] >= 0 ? kfB.mvpMapPoints[vn[k]] : static_cast<MapPoint*>(NULL)), "SearchByBoW(KF, KF): vpMatches12");
    }

    // ---- Tracking.cc:742-743 (MonocularInitialization): ORBmatcher matcher(0.9,true);
    //      int nmatches = matcher.SearchForInitialization(mInitialFrame,mCurrentFrame,mvbPrevMatched,mvIniMatches,100);
    int nE = 0;
    {
        Frame& mInitialFrame = mLastFrame;
        std::vector<cv::Point2f> mvbPrevMatched(mInitialFrame.mvKeysUn.size()), prev2;
        for (size_t i = 0; i < mvbPrevMatched.size(); i++) mvbPrevMatched[i] = mInitialFrame.mvKeysUn[i].pt;
        prev2 = mvbPrevMatched;
        std::vector<int> mvIniMatches, ini2;
        ORBmatcher matcher(0.9, true);
        int nmatches = matcher.SearchForInitialization(mInitialFrame, mCurrentFrame, mvbPrevMatched, mvIniMatches, 100);
        nE = nmatches;
        FrameView V1 = {&mInitialFrame.mvKeysUn, &mInitialFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        FrameView V2 = {&mCurrentFrame.mvKeysUn, &mCurrentFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        const int n2 = matcher.SearchForInitialization(V1, V2, prev2, ini2, 100);
        CHECK(n2 == nmatches && ini2 == mvIniMatches, "SearchForInitialization: matches");
        for (size_t i = 0; i < prev2.size(); i++) CHECK(prev2[i].x == mvbPrevMatched[i].x && prev2[i].y == mvbPrevMatched[i].y, "SearchForInitialization: vbPrevMatched");
    }
    // ---- the same searches against frames resident on the device (DeviceFrame): equal to the host-array path
    {
        ORBmatcher matcher(0.9, true);
        const float K4[4] = {mLastFrame.fx, mLastFrame.fy, mLastFrame.cx, mLastFrame.cy};
        DeviceFrame dLast(matcher, mLastFrame.mvKeysUn, mLastFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h);
        DeviceFrame dCur(matcher, ex, K4, std::vector<float>(), 0.f, (float)w, 0.f, (float)h);      // `ex` extracted the current frame last
        CHECK(dCur.size() == mCurrentFrame.N, "DeviceFrame from the extractor: size");
        std::vector<cv::KeyPoint> back = dCur.keypoints();
        for (int i = 0; i < mCurrentFrame.N; i++) CHECK(back[i].pt.x == mCurrentFrame.mvKeysUn[i].pt.x && back[i].octave == mCurrentFrame.mvKeysUn[i].octave, "DeviceFrame keypoints");
        std::vector<cv::Point2f> p1(mLastFrame.mvKeysUn.size()), p2;
        for (size_t i = 0; i < p1.size(); i++) p1[i] = mLastFrame.mvKeysUn[i].pt;
        p2 = p1;
        std::vector<int> a, b;
        FrameView V1 = {&mLastFrame.mvKeysUn, &mLastFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        FrameView V2 = {&mCurrentFrame.mvKeysUn, &mCurrentFrame.mDescriptors, 0.f, (float)w, 0.f, (float)h};
        const int n1 = matcher.SearchForInitialization(V1, V2, p1, a, 100), n2 = matcher.SearchForInitialization(dLast, dCur, p2, b, 100);
        CHECK(n1 == n2 && a == b, "SearchForInitialization(DeviceFrame) equals the FrameView search");
    }
    CHECK(nA > 50 && nB[0] > 50 && nB[2] > 50 && nC > 20 && nD > 20 && nE > 20, "every search finds matches on the synthetic pair");
    // static DescriptorDistance (MapPoint.cc:286-296 calls it N^2 times): host popcount, equal to the device pairs
    {
        ORBmatcher matcher;
        const int n = std::min(mLastFrame.N, mCurrentFrame.N);
        cv::Mat a(n, 32, CV_8U), b(n, 32, CV_8U);
        for (int i = 0; i < n; i++) { memcpy(a.ptr<uint8_t>(i), mLastFrame.mDescriptors.ptr<uint8_t>(i), 32); memcpy(b.ptr<uint8_t>(i), mCurrentFrame.mDescriptors.ptr<uint8_t>(i), 32); }
        std::vector<int> d = matcher.DescriptorDistances(a, b);
        for (int i = 0; i < n; i++) CHECK(d[i] == ORBmatcher::DescriptorDistance(a.row(i), b.row(i)), "DescriptorDistance host == device");
    }
    printf("%s projection=%d motion=%d/%d/%d bow_kf_f=%d bow_kf_kf=%d init=%d\n", fails ? "FAILED" : "ok", nA, nB[0], nB[1], nB[2], nC, nD, nE);
    return fails ? 4 : 0;
}
