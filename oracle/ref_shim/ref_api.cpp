// ref_api.cpp -- C ABI over the UNMODIFIED reference extractor (test infrastructure, not product).
//
// liborb_ref.so = /root/reference/src/ORBextractor.cc compiled where it lies (never copied) with the
// reference's flags (CMakeLists.txt:26-27,36) + this file + cv_shim.cpp + oracle/orb_primitives.cpp.
// It is the object the oracle restatement and the CUDA path are checked against.
//
// Node allocator.  DistributeOctTree sorts pair<int, ExtractorNode*> (ORBextractor.cc:684), so ties on
// the key count are broken by the HEAP ADDRESS of std::list nodes: the reference's keypoint set depends
// on the allocator and its history.  The source is not touched; what this file controls is the
// environment it runs in, through the replaceable global operator new / delete:
//   allocator 0: malloc / free, i.e. exactly libstdc++'s default (what a stock build does on this libc;
//                results depend on the heap's history, counted in BASELINE.md section 5)
//   allocator 1: a monotone per-thread arena, reset at the start of every call: addresses grow in
//                allocation order and are never re-used inside a call, which makes the reference's
//                own comparison deterministic (a later-created node compares greater).  This is the
//                definition the oracle's tie_mode 0 and the CUDA quadtree are held to.
#include <opencv2/core/core.hpp>
#include "ORBextractor.h"
#include "../orb_oracle.h"
#include <sys/mman.h>
#include <new>
#include <atomic>

namespace {

const size_t ARENA_BYTES = (size_t)2 << 30;     // virtual; only touched pages become resident
const int MAX_ARENAS = 512;
struct Arena { char* base; size_t used; };
std::atomic<int> g_narenas{0};
char* g_arena_base[MAX_ARENAS];
thread_local Arena t_arena = {nullptr, 0};
thread_local bool t_bump = false;               // inside a call that runs under allocator 1
thread_local int t_allocator = 0;
std::atomic<long> g_bump_allocs{0};

void* arena_alloc(size_t n)
{
    if (!t_arena.base) {
        void* p = mmap(nullptr, ARENA_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) throw std::bad_alloc();
        const int slot = g_narenas.fetch_add(1);
        if (slot >= MAX_ARENAS) throw std::bad_alloc();
        g_arena_base[slot] = (char*)p;
        t_arena.base = (char*)p;
        t_arena.used = 0;
    }
    n = (n + 15) & ~(size_t)15;
    if (t_arena.used + n > ARENA_BYTES) throw std::bad_alloc();
    void* r = t_arena.base + t_arena.used;
    t_arena.used += n;
    g_bump_allocs.fetch_add(1, std::memory_order_relaxed);
    return r;
}

bool in_arena(const void* p)
{
    const int n = std::min(g_narenas.load(std::memory_order_acquire), MAX_ARENAS);
    for (int i = 0; i < n; i++)
        if ((const char*)p >= g_arena_base[i] && (const char*)p < g_arena_base[i] + ARENA_BYTES) return true;
    return false;
}

struct BumpScope {
    bool on;
    explicit BumpScope(bool enable) : on(enable) { if (on) { t_arena.used = 0; t_bump = true; } }
    ~BumpScope() { if (on) t_bump = false; }
};

} // namespace

void* operator new(size_t n)
{
    if (t_bump) return arena_alloc(n);
    void* p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

namespace {
// the reference class, with its protected members reachable for stage checks
struct RefExtractor : public ORB_SLAM2::ORBextractor {
    RefExtractor(int nf, float sf, int nl, int ini, int mn) : ORB_SLAM2::ORBextractor(nf, sf, nl, ini, mn) {}
    using ORB_SLAM2::ORBextractor::DistributeOctTree;
    using ORB_SLAM2::ORBextractor::mnFeaturesPerLevel;
    using ORB_SLAM2::ORBextractor::umax;
    using ORB_SLAM2::ORBextractor::mvScaleFactor;
    using ORB_SLAM2::ORBextractor::mvInvScaleFactor;
    using ORB_SLAM2::ORBextractor::mvLevelSigma2;
    using ORB_SLAM2::ORBextractor::mvInvLevelSigma2;
    using ORB_SLAM2::ORBextractor::nlevels;
};
} // namespace

extern "C" {

typedef struct orbr_extractor { RefExtractor* ex; } orbr_extractor;

#ifndef ORBR_FLAGS
#define ORBR_FLAGS "unknown"
#endif
const char* orbr_build_info(void) { return "reference src/ORBextractor.cc, unmodified; flags: " ORBR_FLAGS; }

// allocator for the calling thread's later orbr_extract / orbr_distribute calls (see the header comment)
void orbr_set_allocator(int mode) { t_allocator = mode ? 1 : 0; }
long orbr_bump_allocations(void) { return g_bump_allocs.load(); }

orbr_extractor* orbr_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    orbr_extractor* h = new orbr_extractor;
    h->ex = new RefExtractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
    return h;
}
void orbr_destroy(orbr_extractor* h) { if (h) { delete h->ex; delete h; } }

// ORBextractor::operator() (ORBextractor.cc:1043-1105) on a w x h 8-bit image.  Returns the keypoint count
// (may exceed cap: then only cap are written); -1 if the reference threw.
int orbr_extract(orbr_extractor* h, const uint8_t* img, int w, int hh, size_t stride, orbo_kp* kps, uint8_t* desc, int cap)
{
    try {
        int n = 0;
        {
            BumpScope scope(t_allocator == 1);
            cv::Mat image(hh, w, CV_8UC1, (void*)img, stride);
            std::vector<cv::KeyPoint> keys;
            cv::Mat descriptors;
            (*h->ex)(image, cv::Mat(), keys, descriptors);
            n = (int)keys.size();
            for (int i = 0; i < n && i < cap; i++) {
                const cv::KeyPoint& k = keys[i];
                kps[i].x = k.pt.x; kps[i].y = k.pt.y; kps[i].size = k.size; kps[i].angle = k.angle; kps[i].response = k.response;
                kps[i].octave = k.octave; kps[i].class_id = k.class_id;
                memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
            }
        }
        return n;
    } catch (...) { return -1; }
}

int orbr_levels(orbr_extractor* h) { return h->ex->nlevels; }
void orbr_scale_tables(orbr_extractor* h, float* sf, float* isf, float* sig2, float* isig2)
{
    for (int i = 0; i < h->ex->nlevels; i++) {
        sf[i] = h->ex->mvScaleFactor[i]; isf[i] = h->ex->mvInvScaleFactor[i];
        sig2[i] = h->ex->mvLevelSigma2[i]; isig2[i] = h->ex->mvInvLevelSigma2[i];
    }
}
void orbr_accessor_tables(orbr_extractor* h, float* sf, float* isf, float* sig2, float* isig2, int* levels, float* scaleFactor)
{
    // through the public accessors (ORBextractor.h:63-83), as Frame.cc:69-75 reads them
    std::vector<float> a = h->ex->GetScaleFactors(), b = h->ex->GetInverseScaleFactors(), c = h->ex->GetScaleSigmaSquares(),
                       d = h->ex->GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < a.size(); i++) { sf[i] = a[i]; isf[i] = b[i]; sig2[i] = c[i]; isig2[i] = d[i]; }
    *levels = h->ex->GetLevels();
    *scaleFactor = h->ex->GetScaleFactor();
}
void orbr_features_per_level(orbr_extractor* h, int* out) { for (int i = 0; i < h->ex->nlevels; i++) out[i] = h->ex->mnFeaturesPerLevel[i]; }
void orbr_umax(orbr_extractor* h, int* out16) { for (int i = 0; i < 16; i++) out16[i] = h->ex->umax[i]; }

// mvImagePyramid of the last call (ORBextractor.h:85): size of a level and its padded (w+38) x (h+38) buffer.
void orbr_level_size(orbr_extractor* h, int level, int* w, int* hh) { *w = h->ex->mvImagePyramid[level].cols; *hh = h->ex->mvImagePyramid[level].rows; }
void orbr_level_padded(orbr_extractor* h, int level, uint8_t* dst)
{
    const cv::Mat& m = h->ex->mvImagePyramid[level];
    const size_t step = m.step;
    const uint8_t* origin = m.data - 19 * step - 19;
    for (int y = 0; y < m.rows + 38; y++) memcpy(dst + (size_t)y * (m.cols + 38), origin + (size_t)y * step, m.cols + 38);
}

// ORBextractor::DistributeOctTree (ORBextractor.cc:539-763) alone: candidates (x, y, response) -> indices of the kept ones
// in output order.  The index travels in KeyPoint::class_id, which the reference copies along untouched.
int orbr_distribute(orbr_extractor* h, const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* out_idx, int cap)
{
    try {
        int m = 0;
        {
            BumpScope scope(t_allocator == 1);
            std::vector<cv::KeyPoint> in;
            in.reserve(n);
            for (int i = 0; i < n; i++) in.push_back(cv::KeyPoint((float)xyr[3 * i], (float)xyr[3 * i + 1], 7.f, -1.f, (float)xyr[3 * i + 2], 0, i));
            std::vector<cv::KeyPoint> out = h->ex->DistributeOctTree(in, minX, maxX, minY, maxY, N, 0);
            m = (int)out.size();
            for (int i = 0; i < m && i < cap; i++) out_idx[i] = out[i].class_id;
        }
        return m;
    } catch (...) { return -1; }
}

} // extern "C"
