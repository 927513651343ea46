// The product's ORBmatcher (include/ORBmatcher.h) compiled against the reference's OWN headers and run on the reference's OWN objects:
// Frame, KeyFrame, MapPoint and ORBmatcher here are the unmodified classes of kimwin2/ORB_SLAM2v2-1 (include/*.h, src/*.cc, built into
// oracle/_ref/libslam_ref.so as namespace ORB_SLAM2_REFS; oracle/ref_shim/slam_full/prelude.h says what is stood in beyond them).  Every
// search is called on two identical worlds of such objects with the SAME call line -- once on the reference's matcher, once on the
// product's -- and must return the same count and leave the objects in the same state.  This is what a maintainer who swaps the two
// headers compiles; the stand-in objects of reference_matcher_test.cpp are not involved.
// usage: reference_objects_test a.raw b.raw w h nfeatures dx dy      (b = a moved by (dx, dy) pixels)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <new>
#include <set>
#include <vector>
#include "slam_ref.h"
#include "../../include/ORBextractor.h"
#include "../../include/ORBmatcher.h"

namespace ref = ORB_SLAM2_REFS;
typedef ref::Frame Frame;
typedef ref::KeyFrame KeyFrame;
typedef ref::MapPoint MapPoint;

static int fails = 0;
#define CHECK(cond, what) do { if (!(cond)) { fprintf(stderr, "FAIL: %s (%s:%d)\n", what, __FILE__, __LINE__); fails++; } } while (0)

static std::vector<uint8_t> read_raw(const char* path, size_t n)
{
    std::vector<uint8_t> v(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(3); }
    fclose(f);
    return v;
}

static const float FX = 520.f, FY = 515.f, BF = 40.f, B = 0.08f, Z0 = 4.f;

static cv::Mat pose(float tx, float ty, float tz)
{
    cv::Mat T(4, 4, CV_32F);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) T.at<float>(r, c) = r == c ? 1.f : 0.f;
    T.at<float>(0, 3) = tx; T.at<float>(1, 3) = ty; T.at<float>(2, 3) = tz;
    return T;
}

// A reference Frame around the keypoints the PRODUCT's extractor found (the members the reference's constructors fill, Frame.cc:174-228,
// filled by hand so that both matchers see the same features); the grid through the reference's own PosInGrid, as AssignFeaturesToGrid does.
static void fill_frame(Frame& F, ORB_SLAM2::ORBextractor& ex, std::vector<uint8_t>& img, int w, int h)
{
    cv::Mat im(h, w, CV_8UC1, img.data());
    ex(im, cv::Mat(), F.mvKeys, F.mDescriptors);
    F.mpORBvocabulary = NULL; F.mpORBextractorLeft = NULL; F.mpORBextractorRight = NULL; F.mpReferenceKF = NULL;
    F.mTimeStamp = 0; F.mnId = Frame::nNextId++; F.mThDepth = 35.f;
    F.mK = cv::Mat::eye(3, 3, CV_32F);
    F.mK.at<float>(0, 0) = FX; F.mK.at<float>(1, 1) = FY; F.mK.at<float>(0, 2) = w * 0.5f; F.mK.at<float>(1, 2) = h * 0.5f;
    F.mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    F.N = (int)F.mvKeys.size();
    F.mvKeysUn = F.mvKeys;
    F.mvuRight.assign(F.N, -1.0f);
    F.mvDepth.assign(F.N, -1.0f);
    for (int i = 0; i < F.N; i += 3) { F.mvuRight[i] = F.mvKeys[i].pt.x - BF / Z0; F.mvDepth[i] = Z0; }
    F.mvpMapPoints.assign(F.N, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign(F.N, false);
    F.mvScaleFactors = ex.GetScaleFactors();
    F.mvInvScaleFactors = ex.GetInverseScaleFactors();
    F.mvLevelSigma2 = ex.GetScaleSigmaSquares();
    F.mvInvLevelSigma2 = ex.GetInverseScaleSigmaSquares();
    F.mnScaleLevels = ex.GetLevels();
    F.mfScaleFactor = ex.GetScaleFactor();
    F.mfLogScaleFactor = std::log(F.mfScaleFactor);
    F.mbf = BF; F.mb = B;
    F.SetPose(pose(0, 0, 0));
    F.mFeatVec.clear();
    for (int i = 0; i < F.N; i++) F.mFeatVec.addFeature(F.mDescriptors.ptr<uint8_t>(i)[0] >> 3, (unsigned)i);       // a stand-in vocabulary: 32 nodes
    for (int i = 0; i < FRAME_GRID_COLS; i++) for (int j = 0; j < FRAME_GRID_ROWS; j++) F.mGrid[i][j].clear();
    for (int i = 0; i < F.N; i++) {                                            // Frame.cc:237-244
        int gx, gy;
        if (F.PosInGrid(F.mvKeysUn[i], gx, gy)) F.mGrid[gx][gy].push_back(i);
    }
}

static int assigned(const std::vector<MapPoint*>& v) { int n = 0; for (size_t i = 0; i < v.size(); i++) n += v[i] != NULL; return n; }

// One world of reference objects: two frames of the same scene (world = camera of view A; view B = a camera moved by the image shift at
// depth Z0), a MapPoint behind three quarters of the features of each, two key frames made of the frames.  Built twice -- one world for the
// reference's matcher, one for the product's -- because Fuse changes key frames and map points.
struct World {
    ref::Map map;
    ref::KeyFrameDatabase db;
    Frame FA, FB;
    std::vector<MapPoint*> pa, pb;            // the point of feature i, NULL where it has none
    KeyFrame *KA, *KB;
    void* kfStore;
    std::map<const MapPoint*, int> index;
    World() : KA(NULL), KB(NULL), kfStore(NULL) {}
    ~World()
    {
        if (KA) KA->~KeyFrame();
        if (KB) KB->~KeyFrame();
        ::operator delete(kfStore);
        for (size_t i = 0; i < pa.size(); i++) delete pa[i];
        for (size_t i = 0; i < pb.size(); i++) delete pb[i];
    }
    // MapPoint(Pos, pMap, pFrame, idxF) (MapPoint.cc:52-80): normal, distance range and descriptor from the frame
    void make_points(std::vector<MapPoint*>& store, Frame& F, int base)
    {
        store.assign(F.N, static_cast<MapPoint*>(NULL));
        const cv::Mat Ow = F.GetCameraCenter();
        for (int i = 0; i < F.N; i++) {
            if (i % 4 == 3) continue;
            cv::Mat Pos(3, 1, CV_32F);                                             // depth Z0 in the frame's camera, then into the world (R = I)
            Pos.at<float>(0) = (F.mvKeysUn[i].pt.x - Frame::cx) * Z0 / Frame::fx + Ow.at<float>(0);
            Pos.at<float>(1) = (F.mvKeysUn[i].pt.y - Frame::cy) * Z0 / Frame::fy + Ow.at<float>(1);
            Pos.at<float>(2) = Z0 + Ow.at<float>(2);
            MapPoint* p = new MapPoint(Pos, &map, &F, i);
            map.AddMapPoint(p);
            store[i] = p;
            F.mvpMapPoints[i] = p;
            index[p] = base + i;
        }
    }
    void build(const Frame& FA0, const Frame& FB0, float dx, float dy)
    {
        FA = FA0; FB = FB0;
        FA.SetPose(pose(0, 0, 0));
        FB.SetPose(pose(dx * Z0 / FX, dy * Z0 / FY, 0));
        make_points(pa, FA, 0);
        make_points(pb, FB, 1 << 20);
        // both key frames in one block: their address order -- the order std::map<KeyFrame*, size_t> walks observations in -- is the same in every world
        kfStore = ::operator new(2 * sizeof(KeyFrame));
        KA = new (kfStore) KeyFrame(FA, &map, &db);
        KB = new (static_cast<char*>(kfStore) + sizeof(KeyFrame)) KeyFrame(FB, &map, &db);
        for (int set = 0; set < 2; set++) {
            std::vector<MapPoint*>& v = set ? pb : pa;
            KeyFrame* K = set ? KB : KA;
            for (size_t i = 0; i < v.size(); i++) {
                if (!v[i]) continue;
                v[i]->AddObservation(K, i);
                if (i % 5 == 0) v[i]->nObs = 0;                                    // seen by nobody yet (Observations() == 0)
            }
            for (size_t i = 16; i < v.size(); i += 17) if (v[i]) v[i]->SetBadFlag();      // also leaves the key frame (MapPoint.cc:161-178); the frames keep the pointer
        }
    }
    int id(const MapPoint* p) const
    {
        if (!p) return -1;
        std::map<const MapPoint*, int>::const_iterator it = index.find(p);
        return it == index.end() ? -2 : it->second;
    }
    std::vector<int> ids(const std::vector<MapPoint*>& v) const { std::vector<int> o(v.size()); for (size_t i = 0; i < v.size(); i++) o[i] = id(v[i]); return o; }
    // everything a search can change in the key frames and the map points
    std::vector<int> state() const
    {
        std::vector<int> o = ids(KA->GetMapPointMatches()), b = ids(KB->GetMapPointMatches());
        o.insert(o.end(), b.begin(), b.end());
        for (int set = 0; set < 2; set++) {
            const std::vector<MapPoint*>& v = set ? pb : pa;
            for (size_t i = 0; i < v.size(); i++) {
                if (!v[i]) continue;
                MapPoint* p = v[i];
                o.push_back(p->isBad()); o.push_back(p->Observations()); o.push_back(id(p->GetReplaced()));
                o.push_back(p->GetIndexInKeyFrame(KA)); o.push_back(p->GetIndexInKeyFrame(KB)); o.push_back(p->GetFound());
                const cv::Mat d = p->GetDescriptor();                              // Replace recomputes it (MapPoint.cc:222)
                unsigned hsh = 0;
                for (int k = 0; k < 32; k++) hsh = hsh * 31u + d.ptr<uint8_t>(0)[k];
                o.push_back((int)(hsh & 0x7FFFFFFFu));
            }
        }
        return o;
    }
};

static cv::Mat sim3(const cv::Mat& T, float s)          // [s R | s t; 0 1]
{
    cv::Mat S = T.clone();
    for (int r = 0; r < 3; r++) for (int c = 0; c < 4; c++) S.at<float>(r, c) *= s;
    return S;
}

static void tracking_searches(World& W, const Frame& FBbare, float dx, float dy, ref::ORBmatcher& theirs, ORB_SLAM2::ORBmatcher& mine, int co)
{
    const Frame& FA = W.FA;
    // ---- SearchByProjection(Frame&, vpMapPoints, th)  (ORBmatcher.cc:45-129, Tracking.cc:1329-1336)
    {
        std::vector<MapPoint*> local;
        for (int i = 0; i < FA.N; i++) {
            MapPoint* p = FA.mvpMapPoints[i];
            if (!p) continue;
            p->mbTrackInView = (i % 11 != 10);
            p->mTrackProjX = FA.mvKeysUn[i].pt.x + dx + ((i % 3) - 1) * 0.75f;
            p->mTrackProjY = FA.mvKeysUn[i].pt.y + dy + ((i % 5) - 2) * 0.5f;
            p->mTrackProjXR = p->mTrackProjX - BF / Z0;
            p->mnTrackScaleLevel = FA.mvKeysUn[i].octave;
            p->mTrackViewCos = (i % 2) ? 0.9995f : 0.99f;
            local.push_back(p);
        }
        for (int pass = 0; pass < 2; pass++) {
            Frame R = pass ? W.FB : FBbare, M = R;
            const float th = pass ? 3.f : 5.f;
            const int nr = theirs.SearchByProjection(R, local, th);
            const int nm = mine.SearchByProjection(M, local, th);
            CHECK(nr == nm && R.mvpMapPoints == M.mvpMapPoints, "SearchByProjection(Frame, MapPoints)");
            printf("SearchByProjection(F, MPs) checkOri %d pass %d: %d matches, %d slots\n", co, pass, nr, assigned(R.mvpMapPoints));
            CHECK(nr > FA.N / 8, "SearchByProjection(Frame, MapPoints): too few matches for a meaningful comparison");
        }
    }
    // ---- SearchByProjection(CurrentFrame, LastFrame, th, bMono)  (:1330-1472, Tracking.cc:1012-1034)
    for (int variant = 0; variant < 4; variant++) {
        const float tz = variant == 2 ? -0.3f : (variant == 3 ? 0.3f : 0.f);       // variants 2 / 3: motion along z (forward / backward)
        const bool bMono = variant == 1;
        Frame R = FBbare;
        R.SetPose(pose(dx * Z0 / FX, dy * Z0 / FY, tz));
        Frame M = R;
        Frame Last = FA;
        Last.mvbOutlier.assign(FA.N, false);
        for (int i = 0; i < FA.N; i += 13) Last.mvbOutlier[i] = true;
        const float th = bMono ? 15.f : 7.f;
        const int nr = theirs.SearchByProjection(R, Last, th, bMono);
        const int nm = mine.SearchByProjection(M, Last, th, bMono);
        CHECK(nr == nm && R.mvpMapPoints == M.mvpMapPoints, "SearchByProjection(CurrentFrame, LastFrame)");
        printf("SearchByProjection(Cur, Last) checkOri %d variant %d: %d matches, %d slots\n", co, variant, nr, assigned(R.mvpMapPoints));
        if (variant < 2) CHECK(nr > FA.N / 8, "SearchByProjection(CurrentFrame, LastFrame): too few matches");
    }
    // ---- SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)  (:159-288, Tracking.cc:907-910)
    {
        Frame R = FBbare, M = FBbare;
        std::vector<MapPoint*> vr, vm;
        const int nr = theirs.SearchByBoW(W.KA, R, vr);
        const int nm = mine.SearchByBoW(W.KA, M, vm);
        CHECK(nr == nm && vr == vm, "SearchByBoW(KeyFrame, Frame)");
        printf("SearchByBoW(KF, F) checkOri %d: %d matches\n", co, nr);
        CHECK(nr > FA.N / 10, "SearchByBoW(KeyFrame, Frame): too few matches");
    }
    // ---- SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)  (:522-655, LoopClosing.cc:249-275)
    {
        std::vector<MapPoint*> vr, vm;
        const int nr = theirs.SearchByBoW(W.KA, W.KB, vr);
        const int nm = mine.SearchByBoW(W.KA, W.KB, vm);
        CHECK(nr == nm && vr == vm, "SearchByBoW(KeyFrame, KeyFrame)");
        printf("SearchByBoW(KF, KF) checkOri %d: %d matches\n", co, nr);
        CHECK(nr > FA.N / 10, "SearchByBoW(KeyFrame, KeyFrame): too few matches");
    }
    // ---- SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)  (:405-520, Tracking.cc:742-743)
    {
        std::vector<cv::Point2f> pr(FA.N), pm;
        for (int i = 0; i < FA.N; i++) pr[i] = FA.mvKeysUn[i].pt;
        pm = pr;
        std::vector<int> mr, mm;
        Frame A1 = FA, A2 = FA, B1 = W.FB, B2 = W.FB;
        const int nr = theirs.SearchForInitialization(A1, B1, pr, mr, 100);
        const int nm = mine.SearchForInitialization(A2, B2, pm, mm, 100);
        bool same = nr == nm && mr == mm && pr.size() == pm.size();
        for (size_t i = 0; same && i < pr.size(); i++) same = pr[i].x == pm[i].x && pr[i].y == pm[i].y;
        CHECK(same, "SearchForInitialization");
        printf("SearchForInitialization checkOri %d: %d matches\n", co, nr);
        CHECK(nr > FA.N / 20, "SearchForInitialization: too few matches");
    }
}

static void map_side_searches(const Frame& FA0, const Frame& FB0, int w, int h, float dx, float dy, ref::ORBmatcher& theirs, ORB_SLAM2::ORBmatcher& mine)
{
    // ---- Fuse(pKF, vpMapPoints, th)  (ORBmatcher.cc:827-977, LocalMapping.cc:497, :522)
    {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        CHECK(R.state() == M.state(), "the two worlds start out equal");
        std::vector<MapPoint*> vr = R.FA.mvpMapPoints, vm = M.FA.mvpMapPoints;          // A's points (with NULLs and bad ones) into key frame B
        for (size_t i = 0; i < R.FA.mvpMapPoints.size(); i += 7) { vr.push_back(R.FA.mvpMapPoints[i]); vm.push_back(M.FA.mvpMapPoints[i]); }   // some listed twice
        const float th = 3.f;
        const int nr = theirs.Fuse(R.KB, vr, th);
        const int nm = mine.Fuse(M.KB, vm, th);
        CHECK(nr == nm && R.state() == M.state(), "Fuse(KeyFrame, MapPoints)");
        printf("Fuse(KF, MPs): %d fused of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "Fuse(KeyFrame, MapPoints): too few fused");
    }
    // ---- Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)  (:979-1102, LoopClosing.cc:608)
    {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        std::vector<MapPoint*> vr, vm;
        for (size_t i = 0; i < R.pa.size(); i++) if (R.pa[i]) { vr.push_back(R.pa[i]); vm.push_back(M.pa[i]); }
        const float th = 4.f;
        std::vector<MapPoint*> repR(vr.size(), static_cast<MapPoint*>(NULL)), repM(vm.size(), static_cast<MapPoint*>(NULL));
        const int nr = theirs.Fuse(R.KB, sim3(R.KB->GetPose(), 1.25f), vr, th, repR);
        const int nm = mine.Fuse(M.KB, sim3(M.KB->GetPose(), 1.25f), vm, th, repM);
        CHECK(nr == nm && R.state() == M.state() && R.ids(repR) == M.ids(repM), "Fuse(KeyFrame, Scw, MapPoints)");
        printf("Fuse(KF, Scw, MPs): %d fused of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "Fuse(KeyFrame, Scw, MapPoints): too few fused");
    }
    // ---- SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)  (:290-403, LoopClosing.cc:385)
    {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        std::vector<MapPoint*> vr, vm;
        for (size_t i = 0; i < R.pa.size(); i++) if (R.pa[i]) { vr.push_back(R.pa[i]); vm.push_back(M.pa[i]); }
        // already matched: every sixth feature of B holds its own point, and a few of A's points are among them
        std::vector<MapPoint*> mr(R.KB->N, static_cast<MapPoint*>(NULL)), mm(M.KB->N, static_cast<MapPoint*>(NULL));
        for (int i = 0; i < R.KB->N; i += 6) { mr[i] = R.KB->GetMapPoint(i); mm[i] = M.KB->GetMapPoint(i); }
        for (size_t i = 1; i < vr.size() && 7 * i + 1 < mr.size(); i += 9) { mr[7 * i + 1] = vr[i]; mm[7 * i + 1] = vm[i]; }
        const int th = 10;
        const int nr = theirs.SearchByProjection(R.KB, sim3(R.KB->GetPose(), 0.8f), vr, mr, th);
        const int nm = mine.SearchByProjection(M.KB, sim3(M.KB->GetPose(), 0.8f), vm, mm, th);
        CHECK(nr == nm && R.ids(mr) == M.ids(mm), "SearchByProjection(KeyFrame, Scw, MapPoints)");
        printf("SearchByProjection(KF, Scw, MPs): %d matches of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "SearchByProjection(KeyFrame, Scw, MapPoints): too few matches");
    }
    // ---- SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)  (:1474-1601, Tracking.cc:1597, :1611)
    for (int variant = 0; variant < 2; variant++) {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        std::set<MapPoint*> foundR, foundM;
        for (size_t i = 0; i < R.pa.size(); i += 5) if (R.pa[i]) { foundR.insert(R.pa[i]); foundM.insert(M.pa[i]); }
        // the current frame keeps a third of its own points (occupied slots), the rest is free
        for (int i = 0; i < R.FB.N; i++) if (i % 3) { R.FB.mvpMapPoints[i] = NULL; M.FB.mvpMapPoints[i] = NULL; }
        const float th = variant ? 3.f : 10.f;
        const int ORBdist = variant ? 64 : 100;
        const int nr = theirs.SearchByProjection(R.FB, R.KA, foundR, th, ORBdist);
        const int nm = mine.SearchByProjection(M.FB, M.KA, foundM, th, ORBdist);
        CHECK(nr == nm && R.ids(R.FB.mvpMapPoints) == M.ids(M.FB.mvpMapPoints), "SearchByProjection(Frame, KeyFrame)");
        printf("SearchByProjection(F, KF) variant %d: %d matches\n", variant, nr);
        CHECK(nr > R.FA.N / 12, "SearchByProjection(Frame, KeyFrame): too few matches");
    }
    // ---- SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)  (:1104-1328, LoopClosing.cc:333)
    for (int variant = 0; variant < 2; variant++) {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        const float s12 = 1.f, th = variant ? 7.5f : 10.f;
        cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);                    // camera 2 -> camera 1: p1 = s12 R12 p2 + t12 = p2 + (t1w - t2w)
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R12.at<float>(r, c) = r == c ? 1.f : 0.f;
        const cv::Mat TA = R.KA->GetPose(), TB = R.KB->GetPose();
        for (int k = 0; k < 3; k++) t12.at<float>(k) = TA.at<float>(k, 3) - TB.at<float>(k, 3);
        if (variant) t12.at<float>(0) += 0.004f;
        std::vector<MapPoint*> mr(R.KA->N, static_cast<MapPoint*>(NULL)), mm(M.KA->N, static_cast<MapPoint*>(NULL));
        for (int i = 2; i < R.KA->N && i < R.KB->N; i += 10)              // a few pairs known beforehand
            if (R.KB->GetMapPoint(i)) { mr[i] = R.KB->GetMapPoint(i); mm[i] = M.KB->GetMapPoint(i); }
        const int nr = theirs.SearchBySim3(R.KA, R.KB, mr, s12, R12, t12, th);
        const int nm = mine.SearchBySim3(M.KA, M.KB, mm, s12, R12, t12, th);
        CHECK(nr == nm && R.ids(mr) == M.ids(mm), "SearchBySim3");
        printf("SearchBySim3 variant %d: %d found\n", variant, nr);
        CHECK(nr > R.KA->N / 12, "SearchBySim3: too few found");
    }
    // ---- SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)  (:657-825, LocalMapping.cc:276)
    for (int onlyStereo = 0; onlyStereo < 2; onlyStereo++) {
        World R, M;
        R.build(FA0, FB0, dx, dy);
        M.build(FA0, FB0, dx, dy);
        // only features WITHOUT a map point take part: free two thirds of them in both key frames
        for (int u = 0; u < 2; u++) {
            World& U = u ? M : R;
            for (int i = 0; i < U.KA->N; i++) if (i % 3) U.KA->EraseMapPointMatch(i);
            for (int i = 0; i < U.KB->N; i++) if (i % 3 != 1) U.KB->EraseMapPointMatch(i);
        }
        // fundamental matrix of a translation t = C2 - C1 between identical cameras: F12 = K^-T [t]x K^-1
        cv::Mat TB = R.KB->GetPose();
        const float tx = TB.at<float>(0, 3), ty = TB.at<float>(1, 3), tz = 0.02f;
        TB.at<float>(2, 3) = tz;
        R.KB->SetPose(TB);
        M.KB->SetPose(TB);
        cv::Mat F12(3, 3, CV_32F);
        {
            const double ifx = 1.0 / FX, ify = 1.0 / FY, cx = w * 0.5, cy = h * 0.5;
            const double Kinv[3][3] = {{ifx, 0, -cx * ifx}, {0, ify, -cy * ify}, {0, 0, 1}};
            const double tX[3][3] = {{0, -tz, ty}, {tz, 0, -tx}, {-ty, tx, 0}};
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++) {
                    double a = 0;
                    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) a += Kinv[i][r] * tX[i][j] * Kinv[j][c];
                    F12.at<float>(r, c) = (float)a;
                }
        }
        std::vector<std::pair<size_t, size_t> > pr, pm;
        const int nr = theirs.SearchForTriangulation(R.KA, R.KB, F12, pr, onlyStereo != 0);
        const int nm = mine.SearchForTriangulation(M.KA, M.KB, F12, pm, onlyStereo != 0);
        CHECK(nr == nm && pr == pm, "SearchForTriangulation");
        printf("SearchForTriangulation onlyStereo %d: %d pairs\n", onlyStereo, nr);
        if (!onlyStereo) CHECK(nr > R.KA->N / 20, "SearchForTriangulation: too few pairs");
    }
}

// The reference's own call lines, verbatim (argument types included: `int th`, integer literals, a `float` scale), against the product's class:
// compiled, never run.  Overload resolution must pick the reference-signature templates.
static void call_sites_compile(Frame& mCurrentFrame, Frame& mLastFrame, Frame& mInitialFrame, KeyFrame* mpReferenceKF, KeyFrame* mpCurrentKF, KeyFrame* pKF,
                               KeyFrame* pKF2, KeyFrame* mpCurrentKeyFrame, KeyFrame* pKFi, std::vector<MapPoint*>& mvpLocalMapPoints,
                               std::vector<KeyFrame*>& vpCandidateKFs, bool bMonocular)
{
    std::vector<MapPoint*> vpMapPointMatches, vpFuseCandidates, mvpLoopMapPoints, mvpCurrentMatchedPoints, vpReplacePoints;
    std::vector<std::vector<MapPoint*> > vvpMapPointMatches(1);
    std::vector<cv::Point2f> mvbPrevMatched;
    std::vector<int> mvIniMatches;
    std::vector<std::pair<size_t, size_t> > vMatchedIndices;
    std::set<MapPoint*> sFound;
    cv::Mat F12, mScw, cvScw, R, t;
    const float s = 1.f;
    int i = 0;
    {
        ORB_SLAM2::ORBmatcher matcher(0.9,true);                                                                                    // Tracking.cc:742
        int nmatches = matcher.SearchForInitialization(mInitialFrame,mCurrentFrame,mvbPrevMatched,mvIniMatches,100);                // :743
        (void)nmatches;
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.7,true);                                                                                    // :907
        int nmatches = matcher.SearchByBoW(mpReferenceKF,mCurrentFrame,vpMapPointMatches);                                          // :910
        (void)nmatches;
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.9,true);                                                                                    // :1012
        int th = 15;                                                                                                                // :1023-1027
        int nmatches = matcher.SearchByProjection(mCurrentFrame,mLastFrame,th,bMonocular);                                          // :1028
        nmatches = matcher.SearchByProjection(mCurrentFrame,mLastFrame,2*th,bMonocular);                                            // :1034
        (void)nmatches;
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.8);                                                                                         // :1329
        int th = 1;                                                                                                                 // :1330
        matcher.SearchByProjection(mCurrentFrame,mvpLocalMapPoints,th);                                                             // :1336
    }
    {
        ORB_SLAM2::ORBmatcher matcher2(0.9,true);                                                                                   // :1541
        int nadditional =matcher2.SearchByProjection(mCurrentFrame,vpCandidateKFs[i],sFound,10,100);                                // :1597
        nadditional =matcher2.SearchByProjection(mCurrentFrame,vpCandidateKFs[i],sFound,3,64);                                      // :1611
        (void)nadditional;
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.6,false);                                                                                   // LocalMapping.cc:223
        matcher.SearchForTriangulation(mpCurrentKeyFrame,pKF2,F12,vMatchedIndices,false);                                           // :276
    }
    {
        ORB_SLAM2::ORBmatcher matcher;                                                                                              // :491
        matcher.Fuse(pKFi,vpMapPointMatches);                                                                                       // :497
        matcher.Fuse(mpCurrentKeyFrame,vpFuseCandidates);                                                                           // :522
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.75,true);                                                                                   // LoopClosing.cc:249
        int nmatches = matcher.SearchByBoW(mpCurrentKF,pKF,vvpMapPointMatches[i]);                                                  // :275
        (void)nmatches;
        matcher.SearchBySim3(mpCurrentKF,pKF,vpMapPointMatches,s,R,t,7.5);                                                          // :333
        matcher.SearchByProjection(mpCurrentKF, mScw, mvpLoopMapPoints, mvpCurrentMatchedPoints,10);                                // :385
    }
    {
        ORB_SLAM2::ORBmatcher matcher(0.8);                                                                                         // :598
        matcher.Fuse(pKF,cvScw,mvpLoopMapPoints,4,vpReplacePoints);                                                                 // :608
    }
}

static int run(char** argv)
{
    (void)&call_sites_compile;
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    const float dx = (float)atof(argv[6]), dy = (float)atof(argv[7]);
    std::vector<uint8_t> ia = read_raw(argv[1], (size_t)w * h), ib = read_raw(argv[2], (size_t)w * h);
    ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
    // what the first Frame constructor computes once (Frame.cc:205-221)
    Frame::fx = FX; Frame::fy = FY; Frame::cx = w * 0.5f; Frame::cy = h * 0.5f; Frame::invfx = 1.f / FX; Frame::invfy = 1.f / FY;
    Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
    Frame::mfGridElementWidthInv = (float)FRAME_GRID_COLS / (Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = (float)FRAME_GRID_ROWS / (Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    Frame FA, FB;
    fill_frame(FA, ex, ia, w, h);
    fill_frame(FB, ex, ib, w, h);
    printf("features %d %d\n", FA.N, FB.N);
    for (int co = 0; co < 2; co++) {
        const bool checkOri = co != 0;
        const float nn = co ? 0.9f : 0.7f;
        ref::ORBmatcher theirs(nn, checkOri);
        ORB_SLAM2::ORBmatcher mine(nn, checkOri);
        {
            World W;
            W.build(FA, FB, dx, dy);
            Frame FBbare = FB;                               // view B, posed, without map points
            FBbare.SetPose(W.FB.mTcw);
            tracking_searches(W, FBbare, dx, dy, theirs, mine, co);
        }
        map_side_searches(FA, FB, w, h, dx, dy, theirs, mine);
    }
    // ---- DescriptorDistance (:1649-1665)
    for (int i = 0; i + 1 < FA.N; i += 7)
        CHECK(ref::ORBmatcher::DescriptorDistance(FA.mDescriptors.row(i), FB.mDescriptors.row(i % FB.N)) ==
                  ORB_SLAM2::ORBmatcher::DescriptorDistance(FA.mDescriptors.row(i), FB.mDescriptors.row(i % FB.N)),
              "DescriptorDistance");
    if (!fails) printf("ok\n");
    return fails ? 4 : 0;
}

int main(int argc, char** argv)
{
    if (argc < 8) return 2;
    try {
        return run(argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
