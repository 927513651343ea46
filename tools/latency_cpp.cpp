// latency_cpp.cpp -- latency of the drop-in C++ classes used the way the reference uses them (built and run by tools/bench_latency.py):
//   * ORB_SLAM2::ORBextractor::operator() on one frame per call (Frame.cc:247-253), with and without the mvImagePyramid host mirror;
//   * an ORBmatcher built on the stack per call, as at Tracking.cc:742,907,1012 -- against one that persists;
//   * the tracking searches with host arrays (FrameView) and with the frame resident on the device (orbm_frame).
// usage: latency_cpp a.raw b.raw w h nfeatures   -> one JSON object per line
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <vector>
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;
typedef std::chrono::steady_clock Clock;

static std::vector<uint8_t> read_raw(const char* path, size_t n)
{
    std::vector<uint8_t> v(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(3); }
    fclose(f);
    return v;
}

static void report(const char* what, int reps, const std::function<void()>& fn, const char* extra = "")
{
    for (int i = 0; i < 10; i++) fn();
    std::vector<double> t(reps);
    for (int i = 0; i < reps; i++) {
        const Clock::time_point t0 = Clock::now();
        fn();
        t[i] = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
    }
    std::sort(t.begin(), t.end());
    printf("{\"op\": \"%s\", \"ms_median\": %.4f, \"ms_p99\": %.4f, \"reps\": %d%s}\n", what, t[reps / 2], t[(int)(reps * 0.99)], reps, extra);
    fflush(stdout);
}

int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    std::vector<uint8_t> ia = read_raw(argv[1], (size_t)w * h), ib = read_raw(argv[2], (size_t)w * h);
    try {
        ORBextractor ex(nf, 1.2f, 8, 20, 7);
        cv::Mat imA(h, w, CV_8UC1, ia.data()), imB(h, w, CV_8UC1, ib.data());
        std::vector<cv::KeyPoint> kA, kB;
        cv::Mat dA, dB;
        report("ORBextractor::operator(), mvImagePyramid mirrored (default)", 300, [&] { ex(imA, cv::Mat(), kA, dA); });
        ex.setPyramidDownload(false);
        report("ORBextractor::operator(), setPyramidDownload(false)", 300, [&] { ex(imA, cv::Mat(), kA, dA); });
        ex.setPyramidDownload(true);
        ex(imA, cv::Mat(), kA, dA);
        ex(imB, cv::Mat(), kB, dB);
        char extra[160];
        snprintf(extra, sizeof extra, ", \"features\": [%d, %d]", (int)kA.size(), (int)kB.size());
        FrameView VA = {&kA, &dA, 0.f, (float)w, 0.f, (float)h}, VB = {&kB, &dB, 0.f, (float)w, 0.f, (float)h};
        std::vector<cv::Point2f> prev0(kA.size()), prev;
        for (size_t i = 0; i < kA.size(); i++) prev0[i] = kA[i].pt;
        std::vector<int> m12;
        int nm = 0;
        {
            ORBmatcher persistent(0.9, true);
            report("SearchForInitialization, persistent ORBmatcher", 200, [&] { prev = prev0; nm = persistent.SearchForInitialization(VA, VB, prev, m12, 100); }, extra);
        }
        report("SearchForInitialization, ORBmatcher built on the stack per call (Tracking.cc:742)", 200,
               [&] { ORBmatcher matcher(0.9, true); prev = prev0; nm = matcher.SearchForInitialization(VA, VB, prev, m12, 100); }, extra);
        report("ORBmatcher constructor + destructor alone", 2000, [&] { ORBmatcher matcher(0.9, true); });
        // the frame resident on the device
        orbx_matcher* m = nullptr;
        if (orbm_create(-1, &m) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        std::vector<orbx_kp> pa(kA.size()), pb(kB.size());
        for (size_t i = 0; i < kA.size(); i++) { pa[i].x = kA[i].pt.x; pa[i].y = kA[i].pt.y; pa[i].size = kA[i].size; pa[i].angle = kA[i].angle; pa[i].response = kA[i].response; pa[i].octave = kA[i].octave; pa[i].class_id = -1; }
        for (size_t i = 0; i < kB.size(); i++) { pb[i].x = kB[i].pt.x; pb[i].y = kB[i].pt.y; pb[i].size = kB[i].size; pb[i].angle = kB[i].angle; pb[i].response = kB[i].response; pb[i].octave = kB[i].octave; pb[i].class_id = -1; }
        orbm_frame *FA = nullptr, *FB = nullptr;
        report("orbm_frame_create (upload + device grid), per frame", 200, [&] {
            orbm_frame_destroy(FB);
            if (orbm_frame_create(m, pb.data(), dB.ptr<uint8_t>(0), (int)pb.size(), 0.f, (float)w, 0.f, (float)h, &FB) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        }, extra);
        if (orbm_frame_create(m, pa.data(), dA.ptr<uint8_t>(0), (int)pa.size(), 0.f, (float)w, 0.f, (float)h, &FA) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        std::vector<float> prevf(2 * pa.size());
        std::vector<int> m12d(pa.size());
        int nmd = 0;
        report("SearchForInitialization, device frames (orbm_search_for_initialization_dev)", 300, [&] {
            for (size_t i = 0; i < pa.size(); i++) { prevf[2 * i] = pa[i].x; prevf[2 * i + 1] = pa[i].y; }
            if (orbm_search_for_initialization_dev(m, FA, FB, prevf.data(), m12d.data(), 100, 0.9f, 1, &nmd) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        }, extra);
        if (nmd != nm || m12d != m12) { fprintf(stderr, "device-frame search differs from the FrameView search (%d vs %d)\n", nmd, nm); return 4; }
        // SearchByProjection(Frame, MapPoints): every feature of A re-projected near its place into frame B
        const int np = (int)pa.size();
        std::vector<float> sf = ex.GetScaleFactors(), xy(2 * np), xr(np, -1.f), vc(np, 0.9995f), uR(pb.size(), -1.f);
        std::vector<int> lvl(np), fmA(pb.size()), fmB(pb.size());
        std::vector<uint8_t> has(np, 1), occ(pb.size(), 0), md((size_t)np * 32);
        for (int i = 0; i < np; i++) { xy[2 * i] = pa[i].x + 1.5f; xy[2 * i + 1] = pa[i].y - 1.f; lvl[i] = pa[i].octave; memcpy(&md[(size_t)i * 32], dA.ptr<uint8_t>(i), 32); }
        int n1 = 0, n2 = 0;
        snprintf(extra, sizeof extra, ", \"points\": %d, \"features\": %d", np, (int)pb.size());
        report("SearchByProjection(Frame, MapPoints), host arrays (orbm_search_by_projection)", 300, [&] {
            if (orbm_search_by_projection(m, pb.data(), dB.ptr<uint8_t>(0), uR.data(), occ.data(), (int)pb.size(), 0.f, (float)w, 0.f, (float)h, sf.data(), 8, xy.data(),
                                          xr.data(), lvl.data(), vc.data(), md.data(), has.data(), np, 3.f, 0.8f, fmA.data(), &n1) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        }, extra);
        report("SearchByProjection(Frame, MapPoints), device frame (orbm_search_by_projection_dev)", 300, [&] {
            if (orbm_search_by_projection_dev(m, FB, uR.data(), occ.data(), sf.data(), 8, xy.data(), xr.data(), lvl.data(), vc.data(), md.data(), has.data(), np, 3.f, 0.8f,
                                              fmB.data(), &n2) != ORBX_OK) throw std::runtime_error(orbx_last_error());
        }, extra);
        if (n1 != n2 || fmA != fmB) { fprintf(stderr, "device-frame projection search differs (%d vs %d)\n", n2, n1); return 4; }
        orbm_frame_destroy(FA);
        orbm_frame_destroy(FB);
        orbm_destroy(m);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
