// orbx_extractor.cu -- host side of the extractor half of the C ABI (include/orbx.h): parameter tables, the per-shape
// plan, workspace management and the launch sequence.  No CPU compute path exists: without a CUDA device every entry
// point fails with ORBX_ERR_CUDA.
#include "extract_kernels.cuh"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>

namespace orbx {

static thread_local std::string g_err;

void set_error(const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
}

bool cuda_ok(cudaError_t e, const char* what, const char* file, int line)
{
    if (e == cudaSuccess) return true;
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return false;
}

bool tma_encode_planes(CUtensorMap* map, const void* base, int w, int h, int frames, size_t pitch, size_t frameStride, int boxW, int boxH)
{
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static const EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess) fn = nullptr;
        cudaGetLastError();
        return (EncodeFn)fn;
    }();
    if (!encode || !tma_aligned(base, pitch, frameStride) || boxW > 256 || boxH > 256 || (boxW & 15) || frames < 1) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frameStride};
    const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1u}, estr[3] = {1u, 1u, 1u};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static inline int round_half_even(float v) { return (int)lrintf(v); }      // cvRound
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static const uint32_t kPatternPacked[256] = {
#include "rbrief_pattern.inc"
};

}  // namespace orbx

using namespace orbx;

struct orbx_extractor {
    // reference members (include/ORBextractor.h:93-110)
    int nfeatures = 0;
    double scaleFactor = 0;
    int nlevels = 0, iniThFAST = 0, minThFAST = 0;
    std::vector<int> mnFeaturesPerLevel, umax;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

    int device = 0, maxBatch = 1, fmaMode = 0;
    cudaStream_t stream = nullptr;
    float4* d_patT = nullptr;          // rBRIEF pattern as floats: patT[bit * 32 + byte] = (x0, y0, x1, y1)
    uint2* d_icTab = nullptr;          // IC_Angle weight words (see extract_kernels.cuh)
    UMaxTable um{};

    // current shape
    int W = 0, H = 0, capB = 0;
    Plan plan{};
    FrameBufs bufs{};
    int capS = 0, sortCap = 0;
    size_t fastSmem = 0, octSmem = 0;
    int octForm = 0;                   // orbx_set_octree_form: 0 = automatic
    std::vector<int> xtabOff, ytabOff;
    int* d_xtab = nullptr;            // 8 ints per group of 4 output columns, all levels
    int4* d_ytab = nullptr;
    uint8_t* d_lv0 = nullptr;          // own level-0 planes (host-input path)
    size_t lv0Pitch = 0, lv0Frame = 0;
    void* d_pool = nullptr;            // everything else, one allocation
    uint8_t* h_in = nullptr;           // pinned staging
    uint8_t* h_out = nullptr;
    size_t h_in_bytes = 0, h_out_bytes = 0;
    uint8_t* d_color = nullptr;        // colour staging of the host-input path (orbx_extract_color_batch)
    size_t colorBytes = 0;
    uint8_t* d_scratch = nullptr;      // copy-out scratch (padded level)
    size_t scratchBytes = 0;
    // host mirror of frame 0's pyramid (orbx_set_pyramid_mirror): level 0 at offset 0 with pitch lv0Pitch, level l > 0 at
    // lv0Frame + planeOff[l] with the device pitch; page-locked, filled by two copies that run beside FAST + quadtree
    bool mirror = false;
    uint8_t* h_pyr = nullptr;

    // TMA maps (tma.cuh): unblurred levels with the 48 x 31 box of IC_Angle, blurred levels with the 64 x 37 descriptor box.
    // Levels >= 1 and the blurred planes are the handle's own buffers (encoded once per shape); level 0 follows bufs.lv0.
    LevelMaps mapsImg{}, mapsBlur{}, mapsFast{};      // mapsFast: unblurred levels with the FAST cell box, fastPitch x (hCell + 6)
    bool tmaShape = false, tmaLv0 = false;
    const void* mapLv0 = nullptr;
    unsigned long long mapLv0Pitch = 0, mapLv0Frame = 0;
    int mapLv0B = 0;

    int lastB = 0;                     // frames of the last call (0 = none)
    int lastLaunches = 0;
    bool profile = false;              // record an event after every stage of enqueue()

    // Small batches (the reference's one frame per call) are launch-bound: the 11 launches of a call are captured into a CUDA
    // graph the second time the same (batch, level-0 buffer, knobs) comes by and replayed from then on.
    struct GraphKey {
        int B = 0, form = 0, fma = 0, mirror = 0;
        const void* lv0 = nullptr;
        unsigned long long pitch = 0, frame = 0;
        bool operator==(const GraphKey& o) const
        {
            return B == o.B && form == o.form && fma == o.fma && mirror == o.mirror && lv0 == o.lv0 && pitch == o.pitch && frame == o.frame;
        }
    };
    GraphKey graphKey, seenKey;
    cudaStream_t stream2 = nullptr;    // small batches: the blur runs beside FAST + quadtree (it only needs the pyramid)
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    cudaGraphExec_t graphExec = nullptr;
    bool graphsOff = false;            // capture failed once on this handle, or ORBX_GRAPHS=0
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

namespace orbx {

static void drop_graph(orbx_extractor* ex)
{
    if (ex->graphExec) cudaGraphExecDestroy(ex->graphExec);
    ex->graphExec = nullptr;
    ex->graphKey = ex->seenKey = orbx_extractor::GraphKey();
}

static void free_shape(orbx_extractor* ex)
{
    drop_graph(ex);
    cudaFree(ex->d_xtab); ex->d_xtab = nullptr;
    cudaFree(ex->d_ytab); ex->d_ytab = nullptr;
    cudaFree(ex->d_lv0); ex->d_lv0 = nullptr;
    cudaFree(ex->d_color); ex->d_color = nullptr; ex->colorBytes = 0;
    cudaFree(ex->d_pool); ex->d_pool = nullptr;
    cudaFree(ex->d_scratch); ex->d_scratch = nullptr; ex->scratchBytes = 0;
    if (ex->h_in) cudaFreeHost(ex->h_in);
    if (ex->h_out) cudaFreeHost(ex->h_out);
    if (ex->h_pyr) cudaFreeHost(ex->h_pyr);
    ex->h_in = ex->h_out = ex->h_pyr = nullptr;
    ex->W = ex->H = ex->capB = 0;
    ex->lastB = 0;
}

// Level geometry, cell grid, quadtree capacities (ComputePyramid :1107-1115, ComputeKeyPointsOctTree :773-787,
// DistributeOctTree :543-545).
static int build_plan(orbx_extractor* ex, int w, int h)
{
    Plan& P = ex->plan;
    memset(&P, 0, sizeof P);
    P.nlevels = ex->nlevels;
    P.iniTh = std::min(std::max(ex->iniThFAST, 0), 255);      // cv::FAST clamps its threshold
    P.minTh = std::min(std::max(ex->minThFAST, 0), 255);
    P.fmaMode = ex->fmaMode;
    size_t pyr = 0, blur = 0, slots = 0, keys = 0;
    int cells = 0, nodes = 0, tiles = 0, capS = 0;
    for (int l = 0; l < ex->nlevels; l++) {
        LevelGeom& G = P.lv[l];
        const float scale = ex->mvInvScaleFactor[l];
        G.w = round_half_even((float)w * scale);
        G.h = round_half_even((float)h * scale);
        if (G.w > MAX_DIM || G.h > MAX_DIM) { set_error("level %d is %dx%d: larger than %d", l, G.w, G.h, MAX_DIM); return ORBX_ERR_ARG; }
        G.pitch = (int)align_up((size_t)G.w, 32);
        G.maxBX = G.w - EDGE + 3;
        G.maxBY = G.h - EDGE + 3;
        const float width = (float)(G.maxBX - MIN_BORDER), height = (float)(G.maxBY - MIN_BORDER);
        G.nCols = (int)(width / 30.f);
        G.nRows = (int)(height / 30.f);
        if (G.nCols <= 0 || G.nRows <= 0) {
            set_error("level %d (%dx%d) is smaller than one 30-px FAST cell: the reference divides by zero here", l, G.w, G.h);
            return ORBX_ERR_ARG;
        }
        G.wCell = (int)ceilf(width / (float)G.nCols);
        G.hCell = (int)ceilf(height / (float)G.nRows);
        G.cellBase = cells;
        cells += G.nCols * G.nRows;
        G.slotCap = ((G.wCell + 1) / 2) * ((G.hCell + 1) / 2);      // 3x3 strict NMS keeps an independent set of the king graph
        G.quota = ex->mnFeaturesPerLevel[l];
        G.nIni = (int)roundf(width / height);                        // :543 (float / int difference of ints)
        if (G.nIni <= 0) { set_error("level %d: height > 2*width, the reference divides by zero (nIni = 0)", l); return ORBX_ERR_ARG; }
        G.hX = width / (float)G.nIni;
        G.nodeCap = std::max(G.quota, 4 * G.nIni) + 4;
        capS = std::max(capS, std::max(G.quota + 4, 4 * G.nIni) + 4);
        G.nodeBase = nodes;
        nodes += G.nodeCap;
        // k_blur: a lane owns one word (4 columns) of one 60-row strip; the (strip, word) items of a level are dealt to warps 32 at a time
        G.blurTilesX = (G.w + 3) / 4;            // words per row
        G.blurTileBase = tiles;
        tiles += (G.blurTilesX * ((G.h + 59) / 60) + 31) / 32;
        G.scaledPatch = (int)(31 * ex->mvScaleFactor[l]);            // PATCH_SIZE*mvScaleFactor[level] (:845)
        G.scale = ex->mvScaleFactor[l];
        G.planeOff = pyr;
        if (l > 0) pyr += align_up((size_t)G.pitch * G.h, 256);
        G.blurOff = blur;
        blur += align_up((size_t)G.pitch * G.h, 256);
        G.slotOff = slots;
        G.keyOff = slots;
        slots += (size_t)G.nCols * G.nRows * G.slotCap;
    }
    keys = slots;
    if (capS > MAX_NODES) { set_error("per-level feature quota too large for the quadtree kernel (%d > %d nodes)", capS, MAX_NODES); return ORBX_ERR_ARG; }
    P.totalCells = cells;
    P.totalNodes = nodes;
    P.totalBlurTiles = tiles;
    P.pyrSlab = align_up(pyr, 256);
    P.blurSlab = align_up(blur, 256);
    P.slotSlab = slots;
    P.keySlab = keys;
    ex->capS = capS;
    int sc = 1;
    while (sc < capS) sc <<= 1;
    ex->sortCap = sc;
    P.fastPitch = fast_tile_pitch(P);
    if (!P.fastPitch) { set_error("a FAST cell is wider than the 128-byte tile of the cell kernel"); return ORBX_ERR_ARG; }
    ex->fastSmem = fast_smem_bytes(P);
    ex->octSmem = octree_smem_bytes(capS, sc);
    if (ex->octSmem > 227 * 1024) { set_error("quadtree needs %zu bytes of shared memory (> 227 KB): nfeatures too large", ex->octSmem); return ORBX_ERR_ARG; }
    return ORBX_OK;
}

// cv::resize INTER_LINEAR coefficient tables (8-bit fixed-point path): 11-bit weights from float fx.
// Column table: one record of 8 ints per group of 4 output columns (k_resize):
//   [0] word index of the aligned word pair holding the bytes (sx, sx+1) of columns 0 and 1, [1] same for columns 2, 3,
//   [2] sel01, [3] sel23 : PRMT selectors that build (L0, R0, L1, R1) resp. (L2, R2, L3, R3) from that word pair
//       (R = sx+1 clamped to the last column),
//   [4..7] a0 | a1 << 16 of the four columns.
static bool build_resize_tables(int sw, int sh, int dw, int dh, std::vector<int>& xt, std::vector<int4>& yt)
{
    const double sxs = (double)sw / dw, sys = (double)sh / dh;
    const int ngroups = (dw + 3) / 4;
    std::vector<int> sxv(ngroups * 4), sx1v(ngroups * 4), a0v(ngroups * 4), a1v(ngroups * 4);
    for (int dx = 0; dx < ngroups * 4; dx++) {
        const int d = std::min(dx, dw - 1);          // columns past dw (pitch padding) repeat the last one
        float fx = (float)((d + 0.5) * sxs - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0.f; sx = 0; }
        if (sx + 1 >= sw) { fx = 0.f; sx = sw - 1; }
        sxv[dx] = sx;
        sx1v[dx] = std::min(sx + 1, sw - 1);
        a0v[dx] = (short)round_half_even((1.f - fx) * 2048.f);
        a1v[dx] = (short)round_half_even(fx * 2048.f);
    }
    xt.assign((size_t)ngroups * 8, 0);
    const int lastWord = (sw - 1) >> 2;
    if (lastWord < 1) return false;
    for (int g = 0; g < ngroups; g++) {
        int* r = &xt[(size_t)g * 8];
        for (int pr = 0; pr < 2; pr++) {
            const int k0 = 4 * g + 2 * pr;
            const int word = std::min(sxv[k0] >> 2, lastWord - 1);       // the pair (word, word + 1) stays inside the row
            const int off[4] = {sxv[k0] - 4 * word, sx1v[k0] - 4 * word, sxv[k0 + 1] - 4 * word, sx1v[k0 + 1] - 4 * word};
            int sel = 0;
            for (int k = 0; k < 4; k++) {
                if (off[k] < 0 || off[k] > 7) return false;             // span wider than two words: level ratio above 2
                sel |= off[k] << (4 * k);
            }
            r[pr] = word;
            r[2 + pr] = sel;
        }
        for (int k = 0; k < 4; k++) r[4 + k] = (a0v[4 * g + k] & 0xFFFF) | (a1v[4 * g + k] << 16);
    }
    yt.resize(dh);
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * sys - 0.5);
        const int sy = (int)floorf(fy);
        fy -= sy;
        const int b0 = (short)round_half_even((1.f - fy) * 2048.f), b1 = (short)round_half_even(fy * 2048.f);
        yt[dy] = make_int4(std::min(std::max(sy, 0), sh - 1), std::min(std::max(sy + 1, 0), sh - 1), b0, b1);
    }
    return true;
}

static int ensure_shape(orbx_extractor* ex, int w, int h, int B)
{
    if (ex->W == w && ex->H == h && B <= ex->capB) return ORBX_OK;
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    free_shape(ex);
    int rc = build_plan(ex, w, h);
    if (rc) return rc;
    const Plan& P = ex->plan;
    const int cap = std::max(B, ex->maxBatch);

    // resize tables
    std::vector<int> xall;
    std::vector<int4> yall;
    ex->xtabOff.assign(ex->nlevels, 0);
    ex->ytabOff.assign(ex->nlevels, 0);
    for (int l = 1; l < ex->nlevels; l++) {
        std::vector<int> xt;
        std::vector<int4> yt;
        if (!build_resize_tables(P.lv[l - 1].w, P.lv[l - 1].h, P.lv[l].w, P.lv[l].h, xt, yt)) {
            set_error("scaleFactor too large for the resize kernel (level ratio above 2)");
            return ORBX_ERR_ARG;
        }
        ex->xtabOff[l] = (int)xall.size();
        ex->ytabOff[l] = (int)yall.size();
        xall.insert(xall.end(), xt.begin(), xt.end());
        yall.insert(yall.end(), yt.begin(), yt.end());
    }
    ORBX_CUDA(cudaMalloc(&ex->d_xtab, std::max<size_t>(xall.size(), 8) * sizeof(int)));
    ORBX_CUDA(cudaMalloc(&ex->d_ytab, std::max<size_t>(yall.size(), 1) * sizeof(int4)));
    if (!xall.empty()) {
        ORBX_CUDA(cudaMemcpyAsync(ex->d_xtab, xall.data(), xall.size() * sizeof(int), cudaMemcpyHostToDevice, ex->stream));
        ORBX_CUDA(cudaMemcpyAsync(ex->d_ytab, yall.data(), yall.size() * sizeof(int4), cudaMemcpyHostToDevice, ex->stream));
        ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    }

    ex->lv0Pitch = align_up((size_t)w, 32);
    ex->lv0Frame = align_up(ex->lv0Pitch * h, 256);
    ORBX_CUDA(cudaMalloc(&ex->d_lv0, ex->lv0Frame * cap));

    // one pool for the rest
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
    const size_t oPyr = take(P.pyrSlab * cap), oBlur = take(P.blurSlab * cap);
    const size_t oCnt = take(sizeof(int) * P.totalCells * cap);
    const size_t oSlots = take(sizeof(uint32_t) * P.slotSlab * cap), oKeys = take(sizeof(uint32_t) * P.keySlab * cap);
    const size_t oKnode = take(sizeof(uint16_t) * P.keySlab * cap);
    const size_t oLK = take(sizeof(uint32_t) * P.totalNodes * cap), oLC = take(sizeof(int) * MAX_LEVELS * cap);
    const size_t oCT = take(sizeof(uint4) * FAST_CELL_REC * P.totalCells);
    // outputs contiguous: counts | kps | desc  (one D2H copy)
    const size_t oOut = off;
    const size_t oCounts = take(sizeof(int) * cap), oKps = take(sizeof(orbx_kp) * P.totalNodes * cap);
    const size_t oDesc = take((size_t)32 * P.totalNodes * cap);
    const size_t outBytes = off - oOut;
    ORBX_CUDA(cudaMalloc(&ex->d_pool, off));
    ORBX_CUDA(cudaMemsetAsync(ex->d_pool, 0, off, ex->stream));
    uint8_t* base = (uint8_t*)ex->d_pool;
    FrameBufs& F = ex->bufs;
    F.pyr = base + oPyr;
    F.blur = base + oBlur;
    F.cellCount = (int*)(base + oCnt);
    F.slots = (uint32_t*)(base + oSlots);
    F.keys = (uint32_t*)(base + oKeys);
    F.knode = (uint16_t*)(base + oKnode);
    F.levelKeys = (uint32_t*)(base + oLK);
    F.levelCount = (int*)(base + oLC);
    F.counts = (int*)(base + oCounts);
    F.cellTab = (const uint4*)(base + oCT);
    {   // FAST cell table (ComputeKeyPointsOctTree :789-806)
        std::vector<uint4> tab(FAST_CELL_REC * (size_t)P.totalCells);
        for (int l = 0; l < ex->nlevels; l++) {
            const LevelGeom& G = P.lv[l];
            for (int i = 0; i < G.nRows; i++)
                for (int j = 0; j < G.nCols; j++) fast_cell_record(G, l, i, j, P.fastPitch, &tab[FAST_CELL_REC * (size_t)(G.cellBase + i * G.nCols + j)]);
        }
        ORBX_CUDA(cudaMemcpyAsync(base + oCT, tab.data(), sizeof(uint4) * tab.size(), cudaMemcpyHostToDevice, ex->stream));
        ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    }
    F.kps = (orbx_kp*)(base + oKps);
    F.desc = base + oDesc;
    ex->h_out_bytes = outBytes;
    ex->h_in_bytes = ex->lv0Frame * cap;
    ORBX_CUDA(cudaMallocHost(&ex->h_in, ex->h_in_bytes));
    ORBX_CUDA(cudaMallocHost(&ex->h_out, ex->h_out_bytes));
    if (ex->mirror) ORBX_CUDA(cudaMallocHost(&ex->h_pyr, ex->lv0Frame + P.pyrSlab));
    {   // tensor maps of the handle's own planes; any failure (old driver) leaves the LDGSTS kernels in charge
        static const bool tmaEnv = !(getenv("ORBX_TMA") && atoi(getenv("ORBX_TMA")) == 0);
        bool ok = tmaEnv && ex->nlevels <= TMA_LEVELS;
        for (int l = 0; l < ex->nlevels && ok; l++) {
            const LevelGeom& G = P.lv[l];
            ok = tma_encode_planes(&ex->mapsBlur.m[l], F.blur + G.blurOff, G.w, G.h, cap, (size_t)G.pitch, (size_t)P.blurSlab, 64, 37);
            if (ok && l > 0)
                ok = tma_encode_planes(&ex->mapsImg.m[l], F.pyr + G.planeOff, G.w, G.h, cap, (size_t)G.pitch, (size_t)P.pyrSlab, 48, 31) &&
                     tma_encode_planes(&ex->mapsFast.m[l], F.pyr + G.planeOff, G.w, G.h, cap, (size_t)G.pitch, (size_t)P.pyrSlab, P.fastPitch, G.hCell + 6);
        }
        ex->tmaShape = ok;
        ex->tmaLv0 = false;
        ex->mapLv0 = nullptr;
    }
    ex->W = w; ex->H = h; ex->capB = cap;
    return ORBX_OK;
}

constexpr int GRAPH_MAX_BATCH = 8;

// frame 0's pyramid -> the page-locked host mirror (what the C++ class shows as mvImagePyramid)
static cudaError_t mirror_pyramid(orbx_extractor* ex, cudaStream_t s)
{
    const Plan& P = ex->plan;
    cudaError_t e = cudaMemcpy2DAsync(ex->h_pyr, ex->lv0Pitch, ex->bufs.lv0, ex->bufs.lv0Pitch, (size_t)ex->W, (size_t)ex->H, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && ex->nlevels > 1)
        e = cudaMemcpyAsync(ex->h_pyr + ex->lv0Frame, ex->bufs.pyr, P.pyrSlab, cudaMemcpyDeviceToHost, s);
    return e;
}

// The kernel launches of one call.  Up to GRAPH_MAX_BATCH frames: 11 launches, the blur on the second stream beside FAST + quadtree.
// From SPLIT_MIN_PIXELS on: the batch runs as two halves, one per stream, so that the tail of every kernel of one half (SMs
// draining) is filled by the other half's kernels (measured: 262 k -> 273 k frames/s at 512 VGA frames); 22 launches.
constexpr size_t SPLIT_MIN_PIXELS = (size_t)128 * 640 * 480;      // level-0 pixels of a call from which the split pays (and at least 16 frames)

static int launch_chain(orbx_extractor* ex, cudaStream_t s, int f0, int nb, bool prof, bool forkBlur)
{
    const Plan& P = ex->plan;
    FrameBufs F = ex->bufs;
    F.frame0 = f0;
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[0], s));
    for (int l = 1; l < ex->nlevels; l++) {
        const uint8_t* src = l == 1 ? F.lv0 : F.pyr + P.lv[l - 1].planeOff;
        const int spitch = l == 1 ? (int)F.lv0Pitch : P.lv[l - 1].pitch;
        const size_t sframe = l == 1 ? (size_t)F.lv0Frame : (size_t)P.pyrSlab;
        launch_resize(s, src + (size_t)f0 * sframe, spitch, sframe, F.pyr + P.lv[l].planeOff + (size_t)f0 * P.pyrSlab, P.lv[l].w, P.lv[l].h, P.lv[l].pitch,
                      P.pyrSlab, ex->d_xtab + ex->xtabOff[l], ex->d_ytab + ex->ytabOff[l], nb, P.lv[l - 1].h);
    }
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[1], s));
    // a few frames leave most of the GPU idle: the blur (which needs the pyramid only) runs on a second stream beside
    // FAST + quadtree and joins before the descriptors; inside a captured graph this becomes a fork / join of nodes
    if (forkBlur) {
        ORBX_CUDA(cudaEventRecord(ex->evFork, s));
        ORBX_CUDA(cudaStreamWaitEvent(ex->stream2, ex->evFork, 0));
        launch_blur(ex->stream2, P, F, nb);
        if (ex->mirror) ORBX_CUDA(mirror_pyramid(ex, ex->stream2));      // the copy engine works while FAST + quadtree run
        ORBX_CUDA(cudaEventRecord(ex->evJoin, ex->stream2));
    }
    launch_fast_cells(s, P, F, nb, ex->fastSmem, ex->tmaLv0 ? &ex->mapsFast : nullptr);
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[2], s));
    ORBX_CUDA(launch_octree(s, P, F, nb, ex->capS, ex->sortCap, ex->octForm));
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[3], s));
    if (forkBlur) ORBX_CUDA(cudaStreamWaitEvent(s, ex->evJoin, 0));
    else launch_blur(s, P, F, nb);
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[4], s));
    launch_orient_describe(s, P, F, ex->d_icTab, ex->d_patT, nb, ex->tmaLv0 ? &ex->mapsImg : nullptr, ex->tmaLv0 ? &ex->mapsBlur : nullptr);
    if (prof) ORBX_CUDA(cudaEventRecord(ex->ev[5], s));
    if (ex->mirror && !forkBlur && f0 == 0) ORBX_CUDA(mirror_pyramid(ex, s));
    return ORBX_OK;
}

static int launch_all(orbx_extractor* ex, int B, bool prof)
{
    static const bool splitEnv = !(getenv("ORBX_SPLIT") && atoi(getenv("ORBX_SPLIT")) == 0);
    cudaStream_t s = ex->stream;
    const int perChain = (ex->nlevels - 1) + 4;
    if (!prof && splitEnv && B >= 16 && (size_t)B * ex->W * ex->H >= SPLIT_MIN_PIXELS && ex->stream2 != nullptr) {
        const int h1 = B / 2;
        ORBX_CUDA(cudaEventRecord(ex->evFork, s));
        ORBX_CUDA(cudaStreamWaitEvent(ex->stream2, ex->evFork, 0));
        int rc = launch_chain(ex, s, 0, h1, false, false);
        if (!rc) rc = launch_chain(ex, ex->stream2, h1, B - h1, false, false);
        if (rc) return rc;
        ORBX_CUDA(cudaEventRecord(ex->evJoin, ex->stream2));
        ORBX_CUDA(cudaStreamWaitEvent(s, ex->evJoin, 0));
        ex->lastLaunches = 2 * perChain;
        return ORBX_OK;
    }
    const int rc = launch_chain(ex, s, 0, B, prof, !prof && B <= GRAPH_MAX_BATCH && ex->stream2 != nullptr);
    ex->lastLaunches = perChain;
    return rc;
}

// The whole extraction of B frames on the handle's stream: ComputePyramid -> cell FAST -> DistributeOctTree ->
// GaussianBlur -> orientation + descriptors (operator(), ORBextractor.cc:1043-1105).
// level-0 tensor maps follow the buffer the frames are in (the handle's own planes or the caller's device images)
static void refresh_lv0_maps(orbx_extractor* ex, int B)
{
    const FrameBufs& F = ex->bufs;
    if (!ex->tmaShape) { ex->tmaLv0 = false; return; }
    const int frames = F.lv0 == ex->d_lv0 ? ex->capB : B;
    if (ex->mapLv0 == F.lv0 && ex->mapLv0Pitch == F.lv0Pitch && ex->mapLv0Frame == F.lv0Frame && ex->mapLv0B >= B && ex->mapLv0B == frames) return;
    ex->tmaLv0 = tma_encode_planes(&ex->mapsImg.m[0], F.lv0, ex->plan.lv[0].w, ex->plan.lv[0].h, frames, (size_t)F.lv0Pitch, (size_t)F.lv0Frame, 48, 31) &&
                 tma_encode_planes(&ex->mapsFast.m[0], F.lv0, ex->plan.lv[0].w, ex->plan.lv[0].h, frames, (size_t)F.lv0Pitch, (size_t)F.lv0Frame,
                                   ex->plan.fastPitch, ex->plan.lv[0].hCell + 6);
    ex->mapLv0 = F.lv0; ex->mapLv0Pitch = F.lv0Pitch; ex->mapLv0Frame = F.lv0Frame; ex->mapLv0B = frames;
}

static int enqueue(orbx_extractor* ex, int B)
{
    refresh_lv0_maps(ex, B);
    static const bool graphsEnv = !(getenv("ORBX_GRAPHS") && atoi(getenv("ORBX_GRAPHS")) == 0);
    cudaStream_t s = ex->stream;
    ex->lastB = B;
    if (!ex->profile && graphsEnv && !ex->graphsOff && B <= GRAPH_MAX_BATCH) {
        orbx_extractor::GraphKey key;
        key.B = B; key.form = ex->octForm; key.fma = ex->plan.fmaMode; key.mirror = ex->mirror + 2 * (int)ex->tmaLv0;
        key.lv0 = ex->bufs.lv0; key.pitch = ex->bufs.lv0Pitch; key.frame = ex->bufs.lv0Frame;
        if (ex->graphExec && key == ex->graphKey) {
            ORBX_CUDA(cudaGraphLaunch(ex->graphExec, s));
            return ORBX_OK;                                   // lastLaunches: the kernels inside the graph, as captured
        }
        if (key == ex->seenKey) {                             // second call in a row with these arguments: capture it
            // (the first call ran plainly, so every cudaFuncSetAttribute of the launchers has already happened on this thread)
            cudaGraph_t g = nullptr;
            cudaGraphExec_t exec = nullptr;
            bool ok = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) {
                const int rc = launch_all(ex, B, false);
                ok = cudaStreamEndCapture(s, &g) == cudaSuccess && rc == ORBX_OK && g != nullptr;
            }
            if (ok) ok = cudaGraphInstantiate(&exec, g, 0) == cudaSuccess;
            if (g) cudaGraphDestroy(g);
            if (ok) {
                if (ex->graphExec) cudaGraphExecDestroy(ex->graphExec);
                ex->graphExec = exec;
                ex->graphKey = key;
                ORBX_CUDA(cudaGraphLaunch(ex->graphExec, s));
                return ORBX_OK;
            }
            cudaGetLastError();                               // capture is a shortcut, never a requirement
            ex->graphsOff = true;
        }
        ex->seenKey = key;
    }
    int rc = launch_all(ex, B, ex->profile);
    if (rc) return rc;
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx

// ================================================================================================ C ABI
extern "C" {

const char* orbx_last_error(void) { return g_err.c_str(); }
const char* orbx_version(void) { return "orbx 0.1 sm_100a"; }

int orbx_alloc_pinned(size_t bytes, int write_combined, void** out)
{
    if (!out || bytes == 0) { set_error("orbx_alloc_pinned: bad argument"); return ORBX_ERR_ARG; }
    *out = nullptr;
    ORBX_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocPortable | (write_combined ? cudaHostAllocWriteCombined : 0)));
    return ORBX_OK;
}

int orbx_free_pinned(void* p)
{
    if (p) ORBX_CUDA(cudaFreeHost(p));
    return ORBX_OK;
}

int orbx_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device, int max_batch,
                orbx_extractor** out)
{
    if (!out) { set_error("orbx_create: out is null"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (nfeatures < 0 || nlevels < 1 || nlevels > MAX_LEVELS || !(scaleFactor > 1.0f) || max_batch < 1) {
        set_error("orbx_create: need nfeatures >= 0, 1 <= nlevels <= %d, scaleFactor > 1, max_batch >= 1", MAX_LEVELS);
        return ORBX_ERR_ARG;
    }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { set_error("orbx_create: device %d not present (%d CUDA devices)", device, ndev); return ORBX_ERR_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));

    orbx_extractor* ex = new orbx_extractor;
    ex->nfeatures = nfeatures;
    ex->scaleFactor = scaleFactor;
    ex->nlevels = nlevels;
    ex->iniThFAST = iniThFAST;
    ex->minThFAST = minThFAST;
    ex->device = device;
    ex->maxBatch = max_batch;

    // scale tables: float chain, the member scaleFactor is a double (ORBextractor.cc:415-431)
    ex->mvScaleFactor.resize(nlevels);
    ex->mvLevelSigma2.resize(nlevels);
    ex->mvInvScaleFactor.resize(nlevels);
    ex->mvInvLevelSigma2.resize(nlevels);
    ex->mvScaleFactor[0] = 1.0f;
    ex->mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        ex->mvScaleFactor[i] = (float)(ex->mvScaleFactor[i - 1] * ex->scaleFactor);
        ex->mvLevelSigma2[i] = ex->mvScaleFactor[i] * ex->mvScaleFactor[i];
    }
    for (int i = 0; i < nlevels; i++) {
        ex->mvInvScaleFactor[i] = 1.0f / ex->mvScaleFactor[i];
        ex->mvInvLevelSigma2[i] = 1.0f / ex->mvLevelSigma2[i];
    }
    // per-level quotas: geometric series in float, remainder to the last level (:435-446)
    ex->mnFeaturesPerLevel.resize(nlevels);
    const float factor = (float)(1.0f / ex->scaleFactor);
    float desired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        ex->mnFeaturesPerLevel[l] = round_half_even(desired);
        sum += ex->mnFeaturesPerLevel[l];
        desired *= factor;
    }
    ex->mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);
    // umax: row half-widths of the radius-15 disc, made symmetric (:454-469)
    ex->umax.assign(HALF_PATCH + 1, 0);
    const int vmax = (int)floor(HALF_PATCH * sqrt(2.f) / 2 + 1);
    const int vmin = (int)ceil(HALF_PATCH * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH * HALF_PATCH;
    for (int v = 0; v <= vmax; ++v) ex->umax[v] = (int)lrint(sqrt(hp2 - v * v));
    for (int v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (ex->umax[v0] == ex->umax[v0 + 1]) ++v0;
        ex->umax[v] = v0;
        ++v0;
    }
    for (int i = 0; i < 16; i++) ex->um.v[i] = ex->umax[i];

    if (!cuda_ok(cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking), "cudaStreamCreate", __FILE__, __LINE__)) { delete ex; return ORBX_ERR_CUDA; }
    if (cudaStreamCreateWithFlags(&ex->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ex->evFork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ex->evJoin, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();              // the second stream is an optimisation: without it everything runs on `stream`
        if (ex->stream2) cudaStreamDestroy(ex->stream2);
        ex->stream2 = nullptr;
    }
    // rBRIEF pattern, transposed so that lane = descriptor byte reads consecutive entries, coordinates already as floats
    float4 patT[256];
    for (int byte = 0; byte < 32; byte++)
        for (int bit = 0; bit < 8; bit++) {
            const uint32_t w = kPatternPacked[byte * 8 + bit];
            patT[bit * 32 + byte] = make_float4((float)(int8_t)(w & 0xFF), (float)(int8_t)((w >> 8) & 0xFF),
                                                (float)(int8_t)((w >> 16) & 0xFF), (float)(int8_t)(w >> 24));
        }
    // IC_Angle weights (:77-104): pixel (u, v) of the patch counts iff |u| <= umax[|v|]; for the word `word` of row v + 15 and
    // alignment al, byte b is column u = 4 * word + b - al - 15
    std::vector<uint2> icTab(4 * IC_ITEMS_PAD, make_uint2(0u, 0u));
    for (int al = 0; al < 4; al++)
        for (int row = 0; row < 31; row++)
            for (int word = 0; word < 9; word++) {
                const int v = row - HALF_PATCH;
                uint32_t wu = 0, wv = 0;
                for (int b = 0; b < 4; b++) {
                    const int u = 4 * word + b - al - HALF_PATCH;
                    if (abs(u) <= ex->umax[abs(v)]) {
                        wu |= (uint32_t)(uint8_t)(int8_t)u << (8 * b);
                        wv |= (uint32_t)(uint8_t)(int8_t)v << (8 * b);
                    }
                }
                icTab[(size_t)al * IC_ITEMS_PAD + row * 9 + word] = make_uint2(wu, wv);
            }
    if (!cuda_ok(cudaMalloc(&ex->d_patT, sizeof patT), "cudaMalloc(pattern)", __FILE__, __LINE__) ||
        !cuda_ok(cudaMemcpy(ex->d_patT, patT, sizeof patT, cudaMemcpyHostToDevice), "cudaMemcpy(pattern)", __FILE__, __LINE__) ||
        !cuda_ok(cudaMalloc(&ex->d_icTab, icTab.size() * sizeof(uint2)), "cudaMalloc(icTab)", __FILE__, __LINE__) ||
        !cuda_ok(cudaMemcpy(ex->d_icTab, icTab.data(), icTab.size() * sizeof(uint2), cudaMemcpyHostToDevice), "cudaMemcpy(icTab)", __FILE__, __LINE__)) {
        cudaFree(ex->d_patT);
        cudaFree(ex->d_icTab);
        cudaStreamDestroy(ex->stream);
        delete ex;
        return ORBX_ERR_CUDA;
    }
    *out = ex;
    return ORBX_OK;
}

void orbx_destroy(orbx_extractor* ex)
{
    if (!ex) return;
    cudaSetDevice(ex->device);
    if (ex->stream) cudaStreamSynchronize(ex->stream);
    free_shape(ex);
    cudaFree(ex->d_patT);
    cudaFree(ex->d_icTab);
    for (int i = 0; i < 6; i++)
        if (ex->ev[i]) cudaEventDestroy(ex->ev[i]);
    if (ex->stream2) cudaStreamDestroy(ex->stream2);
    if (ex->evFork) cudaEventDestroy(ex->evFork);
    if (ex->evJoin) cudaEventDestroy(ex->evJoin);
    if (ex->stream) cudaStreamDestroy(ex->stream);
    delete ex;
}

int orbx_set_fma_mode(orbx_extractor* ex, int fma_mode)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    ex->fmaMode = fma_mode ? 1 : 0;
    ex->plan.fmaMode = ex->fmaMode;
    return ORBX_OK;
}

int orbx_set_pyramid_mirror(orbx_extractor* ex, int on)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    if (on && !ex->h_pyr && ex->W) {
        ORBX_CUDA(cudaStreamSynchronize(ex->stream));
        ORBX_CUDA(cudaMallocHost(&ex->h_pyr, ex->lv0Frame + ex->plan.pyrSlab));
    }
    ex->mirror = on != 0;
    return ORBX_OK;
}

int orbx_pyramid_mirror(orbx_extractor* ex, int level, const uint8_t** ptr, size_t* pitch)
{
    if (!ex || !ptr || !pitch || level < 0 || level >= ex->nlevels) { set_error("orbx_pyramid_mirror: bad argument"); return ORBX_ERR_ARG; }
    if (!ex->mirror || !ex->h_pyr || ex->lastB == 0) { set_error("orbx_pyramid_mirror: the mirror is off or no extraction has run"); return ORBX_ERR_STATE; }
    if (level == 0) { *ptr = ex->h_pyr; *pitch = ex->lv0Pitch; }
    else { *ptr = ex->h_pyr + ex->lv0Frame + ex->plan.lv[level].planeOff; *pitch = (size_t)ex->plan.lv[level].pitch; }
    return ORBX_OK;
}

int orbx_set_octree_form(orbx_extractor* ex, int form)
{
    if (!ex || form < 0 || form > 3) { set_error("orbx_set_octree_form: form must be 0..3"); return ORBX_ERR_ARG; }
    ex->octForm = form;
    return ORBX_OK;
}

int orbx_set_profiling(orbx_extractor* ex, int on)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    if (on && !ex->ev[0])
        for (int i = 0; i < 6; i++) ORBX_CUDA(cudaEventCreate(&ex->ev[i]));
    ex->profile = on != 0;
    return ORBX_OK;
}

int orbx_stage_ms(orbx_extractor* ex, float* out5)
{
    if (!ex || !out5) { set_error("null argument"); return ORBX_ERR_ARG; }
    if (!ex->profile || ex->lastB == 0) { set_error("orbx_stage_ms: profiling is off or nothing ran"); return ORBX_ERR_STATE; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    ORBX_CUDA(cudaEventSynchronize(ex->ev[5]));
    for (int i = 0; i < 5; i++) ORBX_CUDA(cudaEventElapsedTime(&out5[i], ex->ev[i], ex->ev[i + 1]));
    return ORBX_OK;
}

int orbx_levels(const orbx_extractor* ex) { return ex ? ex->nlevels : 0; }

int orbx_scale_tables(const orbx_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    for (int i = 0; i < ex->nlevels; i++) {
        if (scale) scale[i] = ex->mvScaleFactor[i];
        if (inv_scale) inv_scale[i] = ex->mvInvScaleFactor[i];
        if (sigma2) sigma2[i] = ex->mvLevelSigma2[i];
        if (inv_sigma2) inv_sigma2[i] = ex->mvInvLevelSigma2[i];
    }
    return ORBX_OK;
}

int orbx_features_per_level(const orbx_extractor* ex, int* out)
{
    if (!ex || !out) { set_error("null argument"); return ORBX_ERR_ARG; }
    for (int i = 0; i < ex->nlevels; i++) out[i] = ex->mnFeaturesPerLevel[i];
    return ORBX_OK;
}

int orbx_umax(const orbx_extractor* ex, int* out16)
{
    if (!ex || !out16) { set_error("null argument"); return ORBX_ERR_ARG; }
    for (int i = 0; i < 16; i++) out16[i] = ex->umax[i];
    return ORBX_OK;
}

int orbx_max_keypoints(const orbx_extractor* ex)
{
    if (!ex) return 0;
    if (ex->W) return ex->plan.totalNodes;      // exact capacity for the current shape
    // before the first call: quota + 2 per level in the largest-first phase, or the 4*nIni children of the unconditional
    // first pass, assuming an aspect ratio <= 8.5 (nIni <= 8)
    int n = 0;
    for (int l = 0; l < ex->nlevels; l++) n += std::max(ex->mnFeaturesPerLevel[l], 4 * 8) + 4;
    return n;
}

void* orbx_stream(orbx_extractor* ex) { return ex ? (void*)ex->stream : nullptr; }
int orbx_last_launches(const orbx_extractor* ex) { return ex ? ex->lastLaunches : 0; }

int orbx_sync(orbx_extractor* ex)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    return ORBX_OK;
}

int orbx_extract_device(orbx_extractor* ex, const uint8_t* d_imgs, int B, int w, int h, size_t pitch, size_t frame_stride)
{
    if (!ex || !d_imgs || B < 1 || w <= 0 || h <= 0) { set_error("orbx_extract_device: bad argument"); return ORBX_ERR_ARG; }
    if ((pitch & 3) || (frame_stride & 3) || ((uintptr_t)d_imgs & 3) || pitch < (size_t)w) {
        set_error("orbx_extract_device: pointer, pitch and frame stride must be multiples of 4 and pitch >= w");
        return ORBX_ERR_ARG;
    }
    ORBX_CUDA(cudaSetDevice(ex->device));
    int rc = ensure_shape(ex, w, h, B);
    if (rc) return rc;
    ex->bufs.lv0 = d_imgs;
    ex->bufs.lv0Pitch = pitch;
    ex->bufs.lv0Frame = frame_stride;
    return enqueue(ex, B);
}

static int fetch_impl(orbx_extractor* ex, orbx_kp* kps, uint8_t* desc, int cap, int* n);

static int color_args_ok(const char* who, int channels)
{
    if (channels != 3 && channels != 4) { set_error("%s: channels must be 3 or 4", who); return 0; }
    return 1;
}

int orbx_extract_color_device(orbx_extractor* ex, const uint8_t* d_imgs, int B, int w, int h, size_t pitch, size_t frame_stride,
                              int channels, int rgb)
{
    if (!ex || !d_imgs || B < 1 || w <= 0 || h <= 0) { set_error("orbx_extract_color_device: bad argument"); return ORBX_ERR_ARG; }
    if (!color_args_ok("orbx_extract_color_device", channels)) return ORBX_ERR_ARG;
    const size_t al = channels == 4 ? 16 : 4;
    if ((pitch % al) || (frame_stride % al) || ((uintptr_t)d_imgs % al) || pitch < (size_t)w * channels) {
        set_error("orbx_extract_color_device: pointer, pitch and frame stride must be multiples of %d and pitch >= w * channels", (int)al);
        return ORBX_ERR_ARG;
    }
    ORBX_CUDA(cudaSetDevice(ex->device));
    int rc = ensure_shape(ex, w, h, B);
    if (rc) return rc;
    launch_gray(ex->stream, d_imgs, pitch, frame_stride, ex->d_lv0, w, h, ex->lv0Pitch, ex->lv0Frame, channels, rgb, B);
    ex->bufs.lv0 = ex->d_lv0;
    ex->bufs.lv0Pitch = ex->lv0Pitch;
    ex->bufs.lv0Frame = ex->lv0Frame;
    rc = enqueue(ex, B);
    if (rc) return rc;
    ex->lastLaunches += 1;
    return ORBX_OK;
}

int orbx_extract_color_batch(orbx_extractor* ex, const uint8_t* const* imgs, int B, int w, int h, size_t stride, int channels, int rgb,
                             orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    if (!ex || !imgs || B < 1 || cap < 0) { set_error("orbx_extract_color_batch: bad argument"); return ORBX_ERR_ARG; }
    if (!color_args_ok("orbx_extract_color_batch", channels)) return ORBX_ERR_ARG;
    if (w <= 0 || h <= 0) {
        if (n) for (int b = 0; b < B; b++) n[b] = 0;
        return ORBX_OK;
    }
    if (stride < (size_t)w * channels) { set_error("orbx_extract_color_batch: stride < w * channels"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    int rc = ensure_shape(ex, w, h, B);
    if (rc) return rc;
    const size_t cpitch = align_up((size_t)w * channels, 32), cframe = align_up(cpitch * h, 256);
    if (ex->colorBytes < cframe * B) {
        ORBX_CUDA(cudaStreamSynchronize(ex->stream));
        cudaFree(ex->d_color);
        ex->d_color = nullptr;
        ex->colorBytes = 0;
        ORBX_CUDA(cudaMalloc(&ex->d_color, cframe * std::max(B, ex->maxBatch)));
        ex->colorBytes = cframe * std::max(B, ex->maxBatch);
    }
    for (int b = 0; b < B; b++) {
        if (!imgs[b]) { set_error("orbx_extract_color_batch: imgs[%d] is null", b); return ORBX_ERR_ARG; }
        ORBX_CUDA(cudaMemcpy2DAsync(ex->d_color + (size_t)b * cframe, cpitch, imgs[b], stride, (size_t)w * channels, h, cudaMemcpyHostToDevice, ex->stream));
    }
    rc = orbx_extract_color_device(ex, ex->d_color, B, w, h, cpitch, cframe, channels, rgb);
    if (rc) return rc;
    return fetch_impl(ex, kps, desc, cap, n);
}

// copies the packed output block (counts | kps | desc) of the last call to the pinned buffer and unpacks it
static bool is_pinned_host(const void* p)
{
    cudaPointerAttributes attr;
    const bool ok = p && cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return ok;
}

static int fetch_impl(orbx_extractor* ex, orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    const int B = ex->lastB;
    const Plan& P = ex->plan;
    uint8_t* base = (uint8_t*)ex->bufs.counts;
    const size_t kOff = (uint8_t*)ex->bufs.kps - base, dOff = ex->bufs.desc - base;
    // Fast path: the caller's arrays are page-locked and laid out like the device results (cap == per-frame stride):
    // the keypoints and descriptors go straight into them, with no staging buffer and no host-side unpacking.
    const bool direct = kps && desc && cap == P.totalNodes && is_pinned_host(kps) && is_pinned_host(desc);
    ORBX_CUDA(cudaMemcpyAsync(ex->h_out, base, sizeof(int) * B, cudaMemcpyDeviceToHost, ex->stream));
    if (direct) {
        ORBX_CUDA(cudaMemcpyAsync(kps, base + kOff, sizeof(orbx_kp) * P.totalNodes * B, cudaMemcpyDeviceToHost, ex->stream));
        ORBX_CUDA(cudaMemcpyAsync(desc, base + dOff, (size_t)32 * P.totalNodes * B, cudaMemcpyDeviceToHost, ex->stream));
    } else {
        ORBX_CUDA(cudaMemcpyAsync(ex->h_out + kOff, base + kOff, sizeof(orbx_kp) * P.totalNodes * B, cudaMemcpyDeviceToHost, ex->stream));
        ORBX_CUDA(cudaMemcpyAsync(ex->h_out + dOff, base + dOff, (size_t)32 * P.totalNodes * B, cudaMemcpyDeviceToHost, ex->stream));
    }
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    int rc = ORBX_OK;
    const int* hc = (const int*)ex->h_out;
    const orbx_kp* hk = (const orbx_kp*)(ex->h_out + kOff);
    const uint8_t* hd = ex->h_out + dOff;
    for (int b = 0; b < B; b++) {
        const int cnt = hc[b];
        if (n) n[b] = cnt;
        if (direct) continue;
        int m = cnt;
        if (m > cap) { m = cap; rc = ORBX_ERR_CAPACITY; set_error("frame %d produced %d keypoints, capacity %d", b, cnt, cap); }
        if (kps && m > 0) memcpy(kps + (size_t)b * cap, hk + (size_t)b * P.totalNodes, sizeof(orbx_kp) * m);
        if (desc && m > 0) memcpy(desc + (size_t)b * cap * 32, hd + (size_t)b * P.totalNodes * 32, (size_t)32 * m);
    }
    return rc;
}

int orbx_fetch(orbx_extractor* ex, orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    if (!ex || cap < 0) { set_error("orbx_fetch: bad argument"); return ORBX_ERR_ARG; }
    if (ex->lastB == 0) { set_error("orbx_fetch: no extraction has run on this handle"); return ORBX_ERR_STATE; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    return fetch_impl(ex, kps, desc, cap, n);
}

int orbx_extract_batch(orbx_extractor* ex, const uint8_t* const* imgs, int B, int w, int h, size_t stride,
                       orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    if (!ex || !imgs || B < 1 || cap < 0) { set_error("orbx_extract_batch: bad argument"); return ORBX_ERR_ARG; }
    if (w <= 0 || h <= 0) {           // empty image: the reference returns without touching its outputs (:1046-1047)
        if (n) for (int b = 0; b < B; b++) n[b] = 0;
        return ORBX_OK;
    }
    if (stride < (size_t)w) { set_error("orbx_extract_batch: stride < w"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    int rc = ensure_shape(ex, w, h, B);
    if (rc) return rc;
    // frames whose host rows are laid out exactly like the device slab (stride == device pitch) travel as one linear
    // copy per run of frames that are also contiguous in host memory; anything else as one 2-D copy per frame
    const bool linear = stride == ex->lv0Pitch;
    for (int b = 0; b < B;) {
        if (!imgs[b]) { set_error("orbx_extract_batch: imgs[%d] is null", b); return ORBX_ERR_ARG; }
        cudaPointerAttributes attr;
        const bool pinned = cudaPointerGetAttributes(&attr, imgs[b]) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        const uint8_t* src = imgs[b];
        size_t spitch = stride;
        if (!pinned) {                 // pageable caller memory: stage through the handle's pinned buffer
            uint8_t* st = ex->h_in + (size_t)b * ex->lv0Frame;
            for (int y = 0; y < h; y++) memcpy(st + (size_t)y * ex->lv0Pitch, imgs[b] + (size_t)y * stride, w);
            src = st;
            spitch = ex->lv0Pitch;
        }
        int run = 1;
        if (pinned && linear) {
            while (b + run < B && imgs[b + run] == imgs[b] + (size_t)run * ex->lv0Frame &&
                   cudaPointerGetAttributes(&attr, imgs[b + run]) == cudaSuccess && attr.type == cudaMemoryTypeHost)
                run++;
            cudaGetLastError();
            const size_t bytes = (size_t)(run - 1) * ex->lv0Frame + (size_t)(h - 1) * stride + w;
            ORBX_CUDA(cudaMemcpyAsync(ex->d_lv0 + (size_t)b * ex->lv0Frame, src, bytes, cudaMemcpyHostToDevice, ex->stream));
        } else {
            ORBX_CUDA(cudaMemcpy2DAsync(ex->d_lv0 + (size_t)b * ex->lv0Frame, ex->lv0Pitch, src, spitch, w, h, cudaMemcpyHostToDevice, ex->stream));
        }
        b += run;
    }
    ex->bufs.lv0 = ex->d_lv0;
    ex->bufs.lv0Pitch = ex->lv0Pitch;
    ex->bufs.lv0Frame = ex->lv0Frame;
    rc = enqueue(ex, B);
    if (rc) return rc;
    return fetch_impl(ex, kps, desc, cap, n);
}

int orbx_extract(orbx_extractor* ex, const uint8_t* img, int w, int h, size_t stride, orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    if (!img || w <= 0 || h <= 0) { if (n) *n = 0; return ORBX_OK; }
    const uint8_t* one[1] = {img};
    return orbx_extract_batch(ex, one, 1, w, h, stride, kps, desc, cap, n);
}

int orbx_results_device(orbx_extractor* ex, const orbx_kp** d_kps, const uint8_t** d_desc, const int** d_counts, int* stride_kp)
{
    if (!ex) { set_error("null handle"); return ORBX_ERR_ARG; }
    if (ex->lastB == 0) { set_error("no extraction has run on this handle"); return ORBX_ERR_STATE; }
    if (d_kps) *d_kps = ex->bufs.kps;
    if (d_desc) *d_desc = ex->bufs.desc;
    if (d_counts) *d_counts = ex->bufs.counts;
    if (stride_kp) *stride_kp = ex->plan.totalNodes;
    return ORBX_OK;
}

int orbx_level_size(const orbx_extractor* ex, int level, int* w, int* h)
{
    if (!ex || level < 0 || level >= ex->nlevels || ex->W == 0) { set_error("orbx_level_size: bad level or no shape yet"); return ORBX_ERR_ARG; }
    if (w) *w = ex->plan.lv[level].w;
    if (h) *h = ex->plan.lv[level].h;
    return ORBX_OK;
}

static int plane_of(orbx_extractor* ex, int frame, int level, const uint8_t** p, size_t* pitch)
{
    if (!ex || ex->lastB == 0) { set_error("no extraction has run on this handle"); return ORBX_ERR_STATE; }
    if (frame < 0 || frame >= ex->lastB || level < 0 || level >= ex->nlevels) { set_error("bad frame/level"); return ORBX_ERR_ARG; }
    if (level == 0) { *p = ex->bufs.lv0 + (size_t)frame * ex->bufs.lv0Frame; *pitch = ex->bufs.lv0Pitch; }
    else { *p = ex->bufs.pyr + (size_t)frame * ex->plan.pyrSlab + ex->plan.lv[level].planeOff; *pitch = ex->plan.lv[level].pitch; }
    return ORBX_OK;
}

int orbx_last_batch(const orbx_extractor* ex) { return ex ? ex->lastB : 0; }

int orbx_pyramid_layout(orbx_extractor* ex, int level, const uint8_t** d_base, size_t* pitch, size_t* frame_stride)
{
    if (!d_base || !pitch || !frame_stride) { set_error("null argument"); return ORBX_ERR_ARG; }
    int rc = plane_of(ex, 0, level, d_base, pitch);
    if (rc) return rc;
    *frame_stride = level == 0 ? (size_t)ex->bufs.lv0Frame : (size_t)ex->plan.pyrSlab;
    return ORBX_OK;
}

int orbx_pyramid_device(orbx_extractor* ex, int frame, int level, const uint8_t** d_ptr, size_t* pitch)
{
    if (!d_ptr || !pitch) { set_error("null argument"); return ORBX_ERR_ARG; }
    return plane_of(ex, frame, level, d_ptr, pitch);
}

static int copy_plane_out(orbx_extractor* ex, const uint8_t* src, size_t spitch, int w, int h, int border, uint8_t* dst, size_t dst_stride)
{
    const int ow = w + 2 * border, oh = h + 2 * border;
    if (dst_stride < (size_t)ow) { set_error("dst_stride too small"); return ORBX_ERR_ARG; }
    const size_t need = (size_t)ow * oh;
    if (need > ex->scratchBytes) {
        cudaFree(ex->d_scratch);
        ex->d_scratch = nullptr;
        ex->scratchBytes = 0;
        ORBX_CUDA(cudaMalloc(&ex->d_scratch, need));
        ex->scratchBytes = need;
    }
    launch_reflect_copy(ex->stream, src, w, h, (int)spitch, ex->d_scratch, border, ow);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpy2DAsync(dst, dst_stride, ex->d_scratch, ow, ow, oh, cudaMemcpyDeviceToHost, ex->stream));
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    return ORBX_OK;
}

int orbx_pyramid_level(orbx_extractor* ex, int frame, int level, int border, uint8_t* dst, size_t dst_stride)
{
    if (!dst || border < 0) { set_error("bad argument"); return ORBX_ERR_ARG; }
    const uint8_t* p;
    size_t pitch;
    int rc = plane_of(ex, frame, level, &p, &pitch);
    if (rc) return rc;
    ORBX_CUDA(cudaSetDevice(ex->device));
    return copy_plane_out(ex, p, pitch, ex->plan.lv[level].w, ex->plan.lv[level].h, border, dst, dst_stride);
}

int orbx_debug_blurred_level(orbx_extractor* ex, int frame, int level, uint8_t* dst, size_t dst_stride)
{
    const uint8_t* p;
    size_t pitch;
    int rc = plane_of(ex, frame, level, &p, &pitch);
    if (rc) return rc;
    if (!dst) { set_error("null dst"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    const LevelGeom& G = ex->plan.lv[level];
    return copy_plane_out(ex, ex->bufs.blur + (size_t)frame * ex->plan.blurSlab + G.blurOff, G.pitch, G.w, G.h, 0, dst, dst_stride);
}

int orbx_debug_level_count(orbx_extractor* ex, int frame, int level, int* n)
{
    const uint8_t* p;
    size_t pitch;
    int rc = plane_of(ex, frame, level, &p, &pitch);
    if (rc) return rc;
    ORBX_CUDA(cudaSetDevice(ex->device));
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    ORBX_CUDA(cudaMemcpy(n, ex->bufs.levelCount + (size_t)frame * MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

int orbx_debug_candidates(orbx_extractor* ex, int frame, int level, int* xyr, int cap, int* n)
{
    const uint8_t* p;
    size_t pitch;
    int rc = plane_of(ex, frame, level, &p, &pitch);
    if (rc) return rc;
    if (!n) { set_error("null n"); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(ex->device));
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    const Plan& P = ex->plan;
    const LevelGeom& G = P.lv[level];
    const int ncell = G.nCols * G.nRows;
    std::vector<int> cnt(ncell);
    ORBX_CUDA(cudaMemcpy(cnt.data(), ex->bufs.cellCount + (size_t)frame * P.totalCells + G.cellBase, sizeof(int) * ncell, cudaMemcpyDeviceToHost));
    int total = 0;
    for (int c : cnt) total += c;
    *n = total;
    if (!xyr) return ORBX_OK;
    // read the per-cell slots (not the quadtree's compacted copy) so that this checks the FAST stage alone
    std::vector<uint32_t> slot(G.slotCap);
    int o = 0;
    for (int c = 0; c < ncell; c++) {
        if (!cnt[c]) continue;
        ORBX_CUDA(cudaMemcpy(slot.data(), ex->bufs.slots + (size_t)frame * P.slotSlab + G.slotOff + (size_t)c * G.slotCap,
                             sizeof(uint32_t) * cnt[c], cudaMemcpyDeviceToHost));
        for (int e = 0; e < cnt[c]; e++, o++)
            if (o < cap) { xyr[3 * o] = key_x(slot[e]); xyr[3 * o + 1] = key_y(slot[e]); xyr[3 * o + 2] = key_r(slot[e]); }
    }
    return total > cap ? ORBX_ERR_CAPACITY : ORBX_OK;
}

}  // extern "C"
