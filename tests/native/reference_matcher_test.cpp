// The product's ORBmatcher (include/ORBmatcher.h, CUDA) against the UNMODIFIED reference src/ORBmatcher.cc (oracle/_ref/libmatcher_ref.so,
// see oracle/ref_shim/slam/matcher_ref.h) on the SAME Frame / KeyFrame / MapPoint objects: every search must leave the objects in the
// same state and return the same count.  Built with -DORBX_USE_OPENCV -I oracle/ref_shim (the cv types of the reference build).
// usage: reference_matcher_test a.raw b.raw w h nfeatures dx dy      (b = a moved by (dx, dy) pixels)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <vector>
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "matcher_ref.h"

typedef refm::Frame Frame;
typedef refm::KeyFrame KeyFrame;
typedef refm::MapPoint MapPoint;

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

static void fill_frame(Frame& F, ORB_SLAM2::ORBextractor& ex, std::vector<uint8_t>& img, int w, int h)
{
    cv::Mat im(h, w, CV_8UC1, img.data());
    ex(im, cv::Mat(), F.mvKeys, F.mDescriptors);
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
    F.mTcw = pose(0, 0, 0);
    F.mFeatVec.clear();
    for (int i = 0; i < F.N; i++) F.mFeatVec.addFeature(F.mDescriptors.ptr<uint8_t>(i)[0] >> 3, (unsigned)i);       // a stand-in vocabulary: 32 nodes
    F.grid.build(F.mvKeysUn, Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY);
}

static void fill_keyframe(KeyFrame& K, const Frame& F, int w, int h)
{
    K.N = F.N; K.fx = Frame::fx; K.fy = Frame::fy; K.cx = Frame::cx; K.cy = Frame::cy; K.invfx = 1.f / K.fx; K.invfy = 1.f / K.fy; K.mbf = F.mbf; K.mb = F.mb;
    K.mvKeys = F.mvKeys; K.mvKeysUn = F.mvKeysUn; K.mvuRight = F.mvuRight; K.mvDepth = F.mvDepth; K.mDescriptors = F.mDescriptors; K.mFeatVec = F.mFeatVec;
    K.mnScaleLevels = F.mnScaleLevels; K.mfScaleFactor = F.mfScaleFactor; K.mfLogScaleFactor = F.mfLogScaleFactor;
    K.mvScaleFactors = F.mvScaleFactors; K.mvLevelSigma2 = F.mvLevelSigma2; K.mvInvLevelSigma2 = F.mvInvLevelSigma2;
    K.mnMinX = 0; K.mnMinY = 0; K.mnMaxX = w; K.mnMaxY = h;
    K.mvpMapPoints = F.mvpMapPoints;
    K.Tcw = F.mTcw.clone();
    K.Ow = cv::Mat(3, 1, CV_32F);
    for (int c = 0; c < 3; c++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)K.Tcw.at<float>(k, c) * (double)K.Tcw.at<float>(k, 3);
        K.Ow.at<float>(c) = (float)(-s);
    }
    K.grid.build(K.mvKeysUn, 0.f, (float)w, 0.f, (float)h);
}

// map points behind the keypoints of F (every fourth keypoint has none): depth z, world = camera of F (pose identity)
static void make_points(std::vector<MapPoint>& store, Frame& F, float z)
{
    store.assign(F.N, MapPoint());
    for (int i = 0; i < F.N; i++) {
        if (i % 4 == 3) continue;
        MapPoint& p = store[i];
        p.mnId = i;
        p.mDescriptor = F.mDescriptors.row(i).clone();
        p.mWorldPos = cv::Mat(3, 1, CV_32F);
        p.mWorldPos.at<float>(0) = (F.mvKeysUn[i].pt.x - Frame::cx) * z / Frame::fx;
        p.mWorldPos.at<float>(1) = (F.mvKeysUn[i].pt.y - Frame::cy) * z / Frame::fy;
        p.mWorldPos.at<float>(2) = z;
        p.mNormalVector = cv::Mat(3, 1, CV_32F);
        const float n = std::sqrt(p.mWorldPos.at<float>(0) * p.mWorldPos.at<float>(0) + p.mWorldPos.at<float>(1) * p.mWorldPos.at<float>(1) + z * z);
        for (int k = 0; k < 3; k++) p.mNormalVector.at<float>(k) = p.mWorldPos.at<float>(k) / n;
        p.mfMaxDistance = n * F.mvScaleFactors[F.mvKeysUn[i].octave];
        p.mfMinDistance = p.mfMaxDistance / F.mvScaleFactors[F.mnScaleLevels - 1];
        p.nObs = (i % 5 == 0) ? 0 : 2;
        p.bad = (i % 17 == 16);
        F.mvpMapPoints[i] = &p;
    }
}

static bool same_points(const std::vector<MapPoint*>& a, const std::vector<MapPoint*>& b)
{
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); i++) if (a[i] != b[i]) return false;
    return true;
}
static int assigned(const std::vector<MapPoint*>& v) { int n = 0; for (size_t i = 0; i < v.size(); i++) n += v[i] != NULL; return n; }

// ---- two identical universes of objects for the searches that MUTATE map points and key frames (Fuse): one for the reference, one
// for the product.  World = camera of view A; view B is the same scene from a camera moved by the image shift at depth Z0.
struct Universe {
    Frame FA, FB;
    std::vector<MapPoint> pa, pb;
    KeyFrame KA, KB;
    void build(const Frame& FA0, const Frame& FB0, int w, int h, float dx, float dy)
    {
        FA = FA0; FB = FB0;
        FA.mTcw = pose(0, 0, 0);
        FB.mTcw = pose(dx * Z0 / FX, dy * Z0 / FY, 0);
        make_points(pa, FA, Z0);
        make_points(pb, FB, Z0);
        for (size_t i = 0; i < pb.size(); i++)           // B's points from B's camera into the world: X_w = X_c - t (R = I)
            if (FB.mvpMapPoints[i])
                for (int k = 0; k < 2; k++) pb[i].mWorldPos.at<float>(k) -= FB.mTcw.at<float>(k, 3);
        fill_keyframe(KA, FA, w, h);
        fill_keyframe(KB, FB, w, h);
        for (size_t i = 0; i < pa.size(); i++) if (FA.mvpMapPoints[i]) pa[i].mObservations[&KA] = i;
        for (size_t i = 0; i < pb.size(); i++) if (FB.mvpMapPoints[i]) pb[i].mObservations[&KB] = i;
    }
    int id(const MapPoint* p) const
    {
        if (!p) return -1;
        if (!pa.empty() && p >= &pa[0] && p <= &pa[pa.size() - 1]) return (int)(p - &pa[0]);
        if (!pb.empty() && p >= &pb[0] && p <= &pb[pb.size() - 1]) return (1 << 20) + (int)(p - &pb[0]);
        return -2;
    }
    std::vector<int> ids(const std::vector<MapPoint*>& v) const { std::vector<int> o(v.size()); for (size_t i = 0; i < v.size(); i++) o[i] = id(v[i]); return o; }
    // everything Fuse can change
    std::vector<int> state() const
    {
        std::vector<int> o = ids(KA.mvpMapPoints), b = ids(KB.mvpMapPoints);
        o.insert(o.end(), b.begin(), b.end());
        for (int set = 0; set < 2; set++) {
            const std::vector<MapPoint>& v = set ? pb : pa;
            for (size_t i = 0; i < v.size(); i++) {
                o.push_back(v[i].bad); o.push_back(v[i].nObs); o.push_back(id(v[i].mpReplaced));
                o.push_back(v[i].mObservations.count(const_cast<KeyFrame*>(&KA)) ? (int)v[i].mObservations.find(const_cast<KeyFrame*>(&KA))->second : -1);
                o.push_back(v[i].mObservations.count(const_cast<KeyFrame*>(&KB)) ? (int)v[i].mObservations.find(const_cast<KeyFrame*>(&KB))->second : -1);
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

static void map_side_searches(const Frame& FA0, const Frame& FB0, int w, int h, float dx, float dy, float nn, bool checkOri, ORB_SLAM2::ORBmatcher& mine)
{
    // ---- Fuse(pKF, vpMapPoints, th)  (ORBmatcher.cc:827-977, LocalMapping::SearchInNeighbors)
    {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        std::vector<MapPoint*> vr = R.FA.mvpMapPoints, vm = M.FA.mvpMapPoints;          // A's points (with NULLs and bad ones) into key frame B
        for (size_t i = 0; i < R.FA.mvpMapPoints.size(); i += 7) { vr.push_back(R.FA.mvpMapPoints[i]); vm.push_back(M.FA.mvpMapPoints[i]); }   // some listed twice
        const float th = 3.f;
        const int nr = refm::Fuse(nn, checkOri, &R.KB, vr, th);
        const int nm = mine.Fuse(&M.KB, vm, th);                          // as at LocalMapping.cc:497 matcher.Fuse(pKFi,vpMapPointMatches) / :522
        CHECK(nr == nm && R.state() == M.state(), "Fuse(KeyFrame, MapPoints)");
        printf("Fuse(KF, MPs): %d fused of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "Fuse(KeyFrame, MapPoints): too few fused");
    }
    // ---- Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)  (:979-1102, LoopClosing::SearchAndFuse)
    {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        std::vector<MapPoint*> vr, vm;
        for (size_t i = 0; i < R.pa.size(); i++) if (R.FA.mvpMapPoints[i]) { vr.push_back(&R.pa[i]); vm.push_back(&M.pa[i]); }
        const float th = 4.f;
        std::vector<MapPoint*> repR(vr.size(), static_cast<MapPoint*>(NULL)), repM(vm.size(), static_cast<MapPoint*>(NULL));
        const int nr = refm::Fuse(nn, checkOri, &R.KB, sim3(R.KB.Tcw, 1.25f), vr, th, repR);
        const int nm = mine.Fuse(&M.KB, sim3(M.KB.Tcw, 1.25f), vm, th, repM);   // as at LoopClosing.cc:608 matcher.Fuse(pKF,cvScw,mvpLoopMapPoints,4,vpReplacePoints)
        CHECK(nr == nm && R.state() == M.state() && R.ids(repR) == M.ids(repM), "Fuse(KeyFrame, Scw, MapPoints)");
        printf("Fuse(KF, Scw, MPs): %d fused of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "Fuse(KeyFrame, Scw, MapPoints): too few fused");
    }
    // ---- SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)  (:290-403, LoopClosing::ComputeSim3)
    {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        std::vector<MapPoint*> vr, vm;
        for (size_t i = 0; i < R.pa.size(); i++) if (R.FA.mvpMapPoints[i]) { vr.push_back(&R.pa[i]); vm.push_back(&M.pa[i]); }
        // already matched: every sixth feature of B holds its own point, and a few of A's points are among them
        std::vector<MapPoint*> mr(R.KB.N, static_cast<MapPoint*>(NULL)), mm(M.KB.N, static_cast<MapPoint*>(NULL));
        for (int i = 0; i < R.KB.N; i += 6) { mr[i] = R.KB.mvpMapPoints[i]; mm[i] = M.KB.mvpMapPoints[i]; }
        for (size_t i = 1; i < vr.size() && 7 * i + 1 < mr.size(); i += 9) { mr[7 * i + 1] = vr[i]; mm[7 * i + 1] = vm[i]; }
        const int th = 10;
        const int nr = refm::SearchByProjection(nn, checkOri, &R.KB, sim3(R.KB.Tcw, 0.8f), vr, mr, th);
        const int nm = mine.SearchByProjection(&M.KB, sim3(M.KB.Tcw, 0.8f), vm, mm, th);   // as at LoopClosing.cc:385 matcher.SearchByProjection(mpCurrentKF, mScw, mvpLoopMapPoints, mvpCurrentMatchedPoints,10)
        CHECK(nr == nm && R.ids(mr) == M.ids(mm), "SearchByProjection(KeyFrame, Scw, MapPoints)");
        printf("SearchByProjection(KF, Scw, MPs): %d matches of %d candidates\n", nr, (int)vr.size());
        CHECK(nr > (int)vr.size() / 10, "SearchByProjection(KeyFrame, Scw, MapPoints): too few matches");
    }
    // ---- SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)  (:1474-1601, Tracking::Relocalization)
    for (int variant = 0; variant < 2; variant++) {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        std::set<MapPoint*> foundR, foundM;
        for (size_t i = 0; i < R.pa.size(); i += 5) if (R.FA.mvpMapPoints[i]) { foundR.insert(&R.pa[i]); foundM.insert(&M.pa[i]); }
        // the current frame keeps a third of its own points (occupied slots), the rest is free
        for (int i = 0; i < R.FB.N; i++) if (i % 3) { R.FB.mvpMapPoints[i] = NULL; M.FB.mvpMapPoints[i] = NULL; }
        const float th = variant ? 3.f : 10.f;
        const int ORBdist = variant ? 64 : 100;
        const int nr = refm::SearchByProjection(nn, checkOri, R.FB, &R.KA, foundR, th, ORBdist);
        const int nm = mine.SearchByProjection(M.FB, &M.KA, foundM, th, ORBdist);   // as at Tracking.cc:1597 matcher2.SearchByProjection(mCurrentFrame,vpCandidateKFs[i],sFound,10,100) / :1611 (..,3,64)
        CHECK(nr == nm && R.ids(R.FB.mvpMapPoints) == M.ids(M.FB.mvpMapPoints), "SearchByProjection(Frame, KeyFrame)");
        printf("SearchByProjection(F, KF) variant %d: %d matches\n", variant, nr);
        CHECK(nr > R.FA.N / 12, "SearchByProjection(Frame, KeyFrame): too few matches");
    }
}


static void loop_and_triangulation_searches(const Frame& FA0, const Frame& FB0, int w, int h, float dx, float dy, float nn, bool checkOri,
                                            ORB_SLAM2::ORBmatcher& mine)
{
    // ---- SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)  (ORBmatcher.cc:1104-1328, LoopClosing::ComputeSim3)
    for (int variant = 0; variant < 2; variant++) {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        const float s12 = 1.f, th = variant ? 7.5f : 10.f;
        cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);                    // camera 2 -> camera 1: p1 = s12 R12 p2 + t12 = p2 + (t1w - t2w)
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R12.at<float>(r, c) = r == c ? 1.f : 0.f;
        for (int k = 0; k < 3; k++) t12.at<float>(k) = R.KA.Tcw.at<float>(k, 3) - R.KB.Tcw.at<float>(k, 3);
        if (variant) t12.at<float>(0) += 0.004f;
        std::vector<MapPoint*> mr(R.KA.N, static_cast<MapPoint*>(NULL)), mm(M.KA.N, static_cast<MapPoint*>(NULL));
        for (int i = 2; i < R.KA.N && i < R.KB.N; i += 10)                // a few pairs known beforehand
            if (R.KB.mvpMapPoints[i]) { mr[i] = R.KB.mvpMapPoints[i]; mm[i] = M.KB.mvpMapPoints[i]; }
        const int nr = refm::SearchBySim3(nn, checkOri, &R.KA, &R.KB, mr, s12, R12, t12, th);
        const int nm = mine.SearchBySim3(&M.KA, &M.KB, mm, s12, R12, t12, th);   // as at LoopClosing.cc:333 matcher.SearchBySim3(mpCurrentKF,pKF,vpMapPointMatches,s,R,t,7.5)
        CHECK(nr == nm && R.ids(mr) == M.ids(mm), "SearchBySim3");
        printf("SearchBySim3 variant %d: %d found\n", variant, nr);
        CHECK(nr > R.KA.N / 12, "SearchBySim3: too few found");
    }
    // ---- SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)  (:657-825, LocalMapping::CreateNewMapPoints)
    for (int onlyStereo = 0; onlyStereo < 2; onlyStereo++) {
        Universe R, M;
        R.build(FA0, FB0, w, h, dx, dy);
        M.build(FA0, FB0, w, h, dx, dy);
        // only features WITHOUT a map point take part: free two thirds of them in both key frames
        for (int i = 0; i < R.KA.N; i++) if (i % 3) { R.KA.mvpMapPoints[i] = NULL; M.KA.mvpMapPoints[i] = NULL; }
        for (int i = 0; i < R.KB.N; i++) if (i % 3 != 1) { R.KB.mvpMapPoints[i] = NULL; M.KB.mvpMapPoints[i] = NULL; }
        // fundamental matrix of a translation t = C2 - C1 between identical cameras: F12 = K^-T [t]x K^-1
        const float tx = R.KB.Tcw.at<float>(0, 3), ty = R.KB.Tcw.at<float>(1, 3), tz = 0.02f;
        for (int u = 0; u < 2; u++) { Universe& U = u ? M : R; U.KB.Tcw.at<float>(2, 3) = tz; }
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
        const int nr = refm::SearchForTriangulation(nn, checkOri, &R.KA, &R.KB, F12, pr, onlyStereo != 0);
        const int nm = mine.SearchForTriangulation(&M.KA, &M.KB, F12, pm, onlyStereo != 0);   // as at LocalMapping.cc:276 matcher.SearchForTriangulation(mpCurrentKeyFrame,pKF2,F12,vMatchedIndices,false)
        CHECK(nr == nm && pr == pm, "SearchForTriangulation");
        printf("SearchForTriangulation onlyStereo %d: %d pairs\n", onlyStereo, nr);
        if (!onlyStereo) CHECK(nr > R.KA.N / 20, "SearchForTriangulation: too few pairs");
    }
}


static int run(char** argv)
{
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    const float dx = (float)atof(argv[6]), dy = (float)atof(argv[7]);
    std::vector<uint8_t> ia = read_raw(argv[1], (size_t)w * h), ib = read_raw(argv[2], (size_t)w * h);
    ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
    Frame::fx = FX; Frame::fy = FY; Frame::cx = w * 0.5f; Frame::cy = h * 0.5f; Frame::invfx = 1.f / FX; Frame::invfy = 1.f / FY;
    Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
    Frame FA, FB;
    fill_frame(FA, ex, ia, w, h);
    fill_frame(FB, ex, ib, w, h);
    std::vector<MapPoint> ptsA, ptsB;
    make_points(ptsA, FA, Z0);
    Frame FBbare = FB;                                  // B without map points
    make_points(ptsB, FB, Z0);
    printf("features %d %d\n", FA.N, FB.N);

    for (int co = 0; co < 2; co++) {
        const bool checkOri = co != 0;
        const float nn = co ? 0.9f : 0.7f;
        ORB_SLAM2::ORBmatcher mine(nn, checkOri);

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
                Frame R = pass ? FB : FBbare, M = R;
                const float th = pass ? 3.f : 5.f;
                const int nr = refm::SearchByProjection(nn, checkOri, R, local, th);
                const int nm = mine.SearchByProjection(M, local, th);
                CHECK(nr == nm && same_points(R.mvpMapPoints, M.mvpMapPoints), "SearchByProjection(Frame, MapPoints)");
                printf("SearchByProjection(F, MPs) checkOri %d pass %d: %d matches, %d slots\n", co, pass, nr, assigned(R.mvpMapPoints));
                CHECK(nr > FA.N / 8, "SearchByProjection(Frame, MapPoints): too few matches for a meaningful comparison");
            }
        }
        // ---- SearchByProjection(CurrentFrame, LastFrame, th, bMono)  (:1330-1472, Tracking.cc:1012-1034)
        for (int variant = 0; variant < 4; variant++) {
            // the image shift as a camera translation at the common depth Z0; variants 2 / 3 add motion along z (forward / backward)
            const float tz = variant == 2 ? -0.3f : (variant == 3 ? 0.3f : 0.f);
            const bool bMono = variant == 1;
            Frame R = FBbare, M;
            R.mTcw = pose(dx * Z0 / FX, dy * Z0 / FY, tz);
            M = R;
            M.mTcw = R.mTcw.clone();
            Frame Last = FA;
            Last.mvbOutlier.assign(FA.N, false);
            for (int i = 0; i < FA.N; i += 13) Last.mvbOutlier[i] = true;
            const float th = bMono ? 15.f : 7.f;
            const int nr = refm::SearchByProjection(nn, checkOri, R, Last, th, bMono);
            const int nm = mine.SearchByProjection(M, Last, th, bMono);
            CHECK(nr == nm && same_points(R.mvpMapPoints, M.mvpMapPoints), "SearchByProjection(CurrentFrame, LastFrame)");
            printf("SearchByProjection(Cur, Last) checkOri %d variant %d: %d matches, %d slots\n", co, variant, nr, assigned(R.mvpMapPoints));
            if (variant < 2) CHECK(nr > FA.N / 8, "SearchByProjection(CurrentFrame, LastFrame): too few matches");
        }
        // ---- SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)  (:159-288, Tracking.cc:907-910)
        KeyFrame KA, KB;
        fill_keyframe(KA, FA, w, h);
        fill_keyframe(KB, FB, w, h);
        {
            Frame R = FBbare, M = FBbare;
            std::vector<MapPoint*> vr, vm;
            const int nr = refm::SearchByBoW(nn, checkOri, &KA, R, vr);
            const int nm = mine.SearchByBoW(&KA, M, vm);
            CHECK(nr == nm && same_points(vr, vm), "SearchByBoW(KeyFrame, Frame)");
            printf("SearchByBoW(KF, F) checkOri %d: %d matches\n", co, nr);
            CHECK(nr > FA.N / 10, "SearchByBoW(KeyFrame, Frame): too few matches");
        }
        // ---- SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)  (:522-655, LoopClosing.cc:249-275)
        {
            std::vector<MapPoint*> vr, vm;
            const int nr = refm::SearchByBoW(nn, checkOri, &KA, &KB, vr);
            const int nm = mine.SearchByBoW(&KA, &KB, vm);
            CHECK(nr == nm && same_points(vr, vm), "SearchByBoW(KeyFrame, KeyFrame)");
            printf("SearchByBoW(KF, KF) checkOri %d: %d matches\n", co, nr);
            CHECK(nr > FA.N / 10, "SearchByBoW(KeyFrame, KeyFrame): too few matches");
        }
        // ---- SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)  (:405-520, Tracking.cc:742-743)
        {
            std::vector<cv::Point2f> pr(FA.N), pm;
            for (int i = 0; i < FA.N; i++) pr[i] = FA.mvKeysUn[i].pt;
            pm = pr;
            std::vector<int> mr, mm;
            Frame A1 = FA, A2 = FA, B1 = FB, B2 = FB;
            const int nr = refm::SearchForInitialization(nn, checkOri, A1, B1, pr, mr, 100);
            const int nm = mine.SearchForInitialization(A2, B2, pm, mm, 100);
            bool same = nr == nm && mr == mm && pr.size() == pm.size();
            for (size_t i = 0; same && i < pr.size(); i++) same = pr[i].x == pm[i].x && pr[i].y == pm[i].y;
            CHECK(same, "SearchForInitialization");
            printf("SearchForInitialization checkOri %d: %d matches\n", co, nr);
            CHECK(nr > FA.N / 20, "SearchForInitialization: too few matches");
        }
        map_side_searches(FA, FBbare, w, h, dx, dy, nn, checkOri, mine);
        loop_and_triangulation_searches(FA, FBbare, w, h, dx, dy, nn, checkOri, mine);
    }
    // ---- DescriptorDistance (:1649-1665)
    for (int i = 0; i + 1 < FA.N; i += 7)
        CHECK(refm::DescriptorDistance(FA.mDescriptors.row(i), FB.mDescriptors.row(i % FB.N)) ==
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
