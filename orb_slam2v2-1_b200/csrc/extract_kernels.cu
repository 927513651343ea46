// extract_kernels.cu -- sm_100a kernels of the ORB extractor (reference src/ORBextractor.cc).
//
// Data layout in HBM (per handle, frame-major slabs, see FrameBufs): every pyramid level is a border-less u8 plane with a
// 32-byte-aligned pitch; the reference's 19-px REFLECT_101 border (ORBextractor.cc:1113-1128) is never read by any
// extraction stage, so it is produced on demand by k_reflect_copy when a caller asks for mvImagePyramid.
// All arithmetic that decides a keypoint or a descriptor bit is integer or individually-rounded float
// (__fmul_rn/__fadd_rn/__fdiv_rn): nvcc must not contract anything here.
#include "extract_kernels.cuh"
#include "tma.cuh"
#include <algorithm>
#include "sincosf_glibc.h"
#include <float.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>

namespace orbx {

// Opt-in to more than 48 KB of dynamic shared memory.  The attribute belongs to the (kernel, device) pair, not to the host
// thread: one table entry per device and kernel, raised under a lock so that a smaller request can never overwrite a larger one.
struct SmemOptIn {
    std::atomic<size_t> granted[64];
    std::mutex lock;
    template <class K> void ensure(K kernel, size_t bytes, size_t floor = 48 * 1024)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        std::atomic<size_t>& g = granted[dev & 63];
        if (bytes <= std::max(g.load(std::memory_order_acquire), floor)) return;
        std::lock_guard<std::mutex> hold(lock);
        if (bytes <= std::max(g.load(std::memory_order_acquire), floor)) return;
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        g.store(bytes, std::memory_order_release);
    }
};

__device__ __forceinline__ const uint8_t* level_plane(const Plan& P, const FrameBufs& F, int f, int l, int& pitch)
{
    if (l == 0) { pitch = (int)F.lv0Pitch; return F.lv0 + (size_t)f * F.lv0Frame; }
    pitch = P.lv[l].pitch;
    return F.pyr + (size_t)f * P.pyrSlab + P.lv[l].planeOff;
}

__device__ __forceinline__ int level_of_cell(const Plan& P, int cell)
{
    int l = 0;
    while (l + 1 < P.nlevels && cell >= P.lv[l + 1].cellBase) ++l;
    return l;
}

// ------------------------------------------------------------------------------------------------ pyramid
// cv::resize INTER_LINEAR, 8-bit: 11-bit coefficients (host tables), horizontal pass H = S[sx]*a0 + S[sx+1]*a1 in int32,
// vertical pass (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2  (always within 0..255: the weights sum to 2048).
// One CTA = a tile of 128 output columns x RS_TILE_H output rows (or 64 x 2*RS_TILE_H, see GX below), in two phases:
//   1. horizontal pass of every source row the tile touches, once, into shared memory (H >> 4 as u16, one uint2 per
//      group of 4 output columns: the kernel is bound by L1/shared-memory wavefronts, so the narrow format wins).  The (left, right) source bytes of two adjacent output columns span at most 4
//      bytes (level ratio <= 2), i.e. two aligned words: one PRMT (selector precomputed per column group on the host)
//      builds the word (L0, R0, L1, R1) and one IDP.2A per column (FMA pipe, the 16-bit weight pair a0 | a1 << 16 as the
//      other operand) finishes it.  Rows are independent, so the loop is unrolled and many loads are in flight per thread
//      (the level-0 source streams from HBM);
//   2. vertical pass from shared memory: two LDS.64 per 4 output pixels, products on the FMA pipe (IMAD), shift-adds as
//      LEA.HI, one 32-bit store.
constexpr int RS_TILE_H = 64;

struct ResizeCols { int4 a, b; };   // a = {word of pair 01, word of pair 23, sel01, sel23}; b = 4 x (a0 | a1 << 16)

// GX = column groups per tile row: 32 (a warp = one row of 128 output columns) or 16 (a warp = two rows of 64 columns, tiles
// twice as tall).  The host picks per level whichever leaves fewer lanes idle in the last tile column (widths such as 533 = 134
// groups waste 16 % of the lanes at GX = 32 and 7 % at GX = 16).
template <int GX>
__global__ void __launch_bounds__(256) k_resize(const uint8_t* __restrict__ src, int spitch, size_t sframe,
                                                uint8_t* __restrict__ dst, int dw, int dh, int dpitch, size_t dframe,
                                                const ResizeCols* __restrict__ xtab, const int4* __restrict__ ytab)
{
    constexpr int RPW = 32 / GX, RSTEP = 8 * RPW, TILE_H = RS_TILE_H * RPW;
    extern __shared__ uint2 Hs[];                                  // [source row of the tile][column group], 4 x u16
    const int f = blockIdx.z, lane = threadIdx.x & 31, tx = lane % GX, ty = (threadIdx.x >> 5) * RPW + lane / GX;
    const int g = min(blockIdx.x * GX + tx, (dw - 1) >> 2);        // column group (clamped: idle lanes repeat the last one)
    const int oy0 = blockIdx.y * TILE_H, oy1 = min(oy0 + TILE_H, dh);
    const int sFirst = __ldg(&ytab[oy0]).x, n = __ldg(&ytab[oy1 - 1]).y - sFirst + 1;
    const int4 ca = __ldg(&xtab[g].a), cb = __ldg(&xtab[g].b);
    const uint32_t* S = reinterpret_cast<const uint32_t*>(src + (size_t)f * sframe + (size_t)sFirst * spitch);
    const int wpitch = spitch >> 2;
    // pointer stepping: two running row pointers (pair 01 / pair 23)
    const uint32_t* p01 = S + (size_t)ty * wpitch + ca.x;
    const uint32_t* p23 = S + (size_t)ty * wpitch + ca.y;
    const size_t rstep = (size_t)wpitch * RSTEP;
    // unrolled by 2 (8 loads in flight per thread) and the vertical loop not at all: measured best, 4 / 2 was 7 % slower
#pragma unroll 2
    for (int r = ty; r < n; r += RSTEP, p01 += rstep, p23 += rstep) {
        const uint32_t wa = __ldg(p01), wb = __ldg(p01 + 1), wc = __ldg(p23), wd = __ldg(p23 + 1);
        const uint32_t W01 = __byte_perm(wa, wb, ca.z), W23 = __byte_perm(wc, wd, ca.w);
        // H >> 4 <= 32640 fits 16 bits: (H1 >> 4) << 16 | (H0 >> 4) == ((H1 << 12) & 0xFFFF0000) | (H0 >> 4), the merge on the FMA pipe
        const unsigned H0 = __dp2a_lo((unsigned)cb.x, W01, 0u), H1 = __dp2a_hi((unsigned)cb.y, W01, 0u);
        const unsigned H2 = __dp2a_lo((unsigned)cb.z, W23, 0u), H3 = __dp2a_hi((unsigned)cb.w, W23, 0u);
        Hs[r * GX + tx] = make_uint2((H1 >> 4) * 65536u + (H0 >> 4), (H3 >> 4) * 65536u + (H2 >> 4));
    }
    __syncthreads();
    if (blockIdx.x * GX + tx > ((dw - 1) >> 2)) return;
    uint8_t* D = dst + (size_t)f * dframe + (size_t)g * 4;
#pragma unroll 1
    for (int dy = oy0 + ty; dy < oy1; dy += RSTEP) {
        const int4 yy = __ldg(&ytab[dy]);                          // {sy0, sy1 (clamped), b0, b1}
        const uint2 h0 = Hs[(yy.x - sFirst) * GX + tx], h1 = Hs[(yy.y - sFirst) * GX + tx];
        const unsigned b0 = (unsigned)yy.z, b1 = (unsigned)yy.w;
        // v = ((b0 * H0) >> 16) + ((b1 * H1) >> 16) + 2 <= 1022: the upper halves of two products are gathered by one PRMT
        // each and both pixels of a pair are summed as u16x2 in one add; (v >> 2) then sits in bytes 0 and 2 of the shifted word
        const unsigned A01 = __byte_perm(b0 * (h0.x & 0xFFFFu), b0 * (h0.x >> 16), 0x7632);
        const unsigned B01 = __byte_perm(b1 * (h1.x & 0xFFFFu), b1 * (h1.x >> 16), 0x7632);
        const unsigned A23 = __byte_perm(b0 * (h0.y & 0xFFFFu), b0 * (h0.y >> 16), 0x7632);
        const unsigned B23 = __byte_perm(b1 * (h1.y & 0xFFFFu), b1 * (h1.y >> 16), 0x7632);
        const unsigned p01v = (A01 + B01 + 0x00020002u) >> 2, p23v = (A23 + B23 + 0x00020002u) >> 2;
        *reinterpret_cast<uint32_t*>(D + (size_t)dy * dpitch) = __byte_perm(p01v, p23v, 0x6420);     // pitch is a multiple of 32: the last word may pad
    }
}

// shared memory for a level pair: source rows one tile can touch
static size_t resize_smem_bytes(int sh, int dh, int gx)
{
    const int tileH = RS_TILE_H * (32 / gx);
    const int rows = (int)((double)tileH * sh / dh) + 4;
    return (size_t)rows * gx * sizeof(uint2);
}

template <int GX>
static void launch_resize_gx(cudaStream_t s, const uint8_t* src, int spitch, size_t sframe, uint8_t* dst, int dw, int dh, int dpitch,
                             size_t dframe, const void* xtab, const int4* ytab, int B, int sh)
{
    const size_t smem = resize_smem_bytes(sh, dh, GX);
    static SmemOptIn optin;
    optin.ensure(k_resize<GX>, smem);
    const int tileH = RS_TILE_H * (32 / GX);
    dim3 grd(((dw + 3) / 4 + GX - 1) / GX, (dh + tileH - 1) / tileH, B);
    k_resize<GX><<<grd, 256, smem, s>>>(src, spitch, sframe, dst, dw, dh, dpitch, dframe, (const ResizeCols*)xtab, ytab);
}

void launch_resize(cudaStream_t s, const uint8_t* src, int spitch, size_t sframe, uint8_t* dst, int dw, int dh, int dpitch,
                   size_t dframe, const void* xtab, const int4* ytab, int B, int sh)
{
    const int groups = (dw + 3) / 4;
    const double u32 = (double)groups / (((groups + 31) / 32) * 32), u16 = (double)groups / (((groups + 15) / 16) * 16);
    if (u16 > u32 + 0.03) launch_resize_gx<16>(s, src, spitch, sframe, dst, dw, dh, dpitch, dframe, xtab, ytab, B, sh);
    else launch_resize_gx<32>(s, src, spitch, sframe, dst, dw, dh, dpitch, dframe, xtab, ytab, B, sh);
}

// ------------------------------------------------------------------------------------------------ FAST per cell
// One WARP per cell and one cell per CTA (a CTA of several cells holds its slots until its slowest cell, e.g. one that needs
// the minThFAST pass, is done: 0.83 -> 0.785 ms); everything below is warp-synchronous, no block barrier, no shared-memory atomic.
//   1. the cell ROI is staged in shared memory with a compile-time row pitch TP (64 bytes, 128 for very wide cells), its
//      origin on a 16-byte boundary of the level row, so that a lane copies 16 bytes per LDGSTS;
//   2. pre-filter on 8 pixels per lane (two words): a 9-arc always contains two adjacent compass pixels (ring positions
//      0, 4, 8, 12), so a corner needs |p - centre| > t on such a pair; |.| comes from the native VABSDIFF4.U8 and the
//      threshold test is SWAR on the bytes; the 8 result bits of a lane go into a per-row bit mask (one multiply gathers
//      them), and the survivor list is built in row-major order with one lane per row (shuffle scan of the row counts,
//      then each lane expands its row's set bits);
//   3. exact score of the survivors (packed s16x2 min/max, VIMNMX) into a u8 score map (0 = not a corner); the survivors
//      that are corners are compacted in place, in order;
//   4. NMS + emission from that list, which is already in cv::FAST's output order.
// The cell is processed at iniThFAST first and, only if that yields no keypoint, again at minThFAST
// (ORBextractor.cc:809-816): kept(t) = raw > t and raw > every neighbour's raw, so each pass is exact on its own.
#ifndef FAST_WARPS_N
#define FAST_WARPS_N 1
#endif
constexpr int FAST_WARPS = FAST_WARPS_N;

// raw = max over the 16 arcs of 9 contiguous ring pixels of min(ring - centre) resp. min(centre - ring)
// (OpenCV cornerScore<16>; corner iff raw > threshold, response = raw - 1).  min/max commute with the subtraction of the
// centre, so the arcs are evaluated on the raw ring values, and max over an arc = 255 - min over the arc of (255 - ring), so ONE
// max-of-min network serves both polarities: E[j] = (ring[j], 255 - ring[j]) as s16x2 (one IMAD per ring pixel, i.e. on the FMA
// pipe), nine-element minima from shared triples T[j] = min3(E[j..j+2]) and W[k] = min3(T[k], T[k+3], T[k+6]) (indices mod 16),
// then the maximum of the 16 W -- 40 three-input VIMNMX3 per pixel (the kernel binds on the ALU pipe, where they run).
__device__ __forceinline__ unsigned vmin3s2(unsigned a, unsigned b, unsigned c) { return __vmins2(__vmins2(a, b), c); }
__device__ __forceinline__ unsigned vmax3s2(unsigned a, unsigned b, unsigned c) { return __vmaxs2(__vmaxs2(a, b), c); }

template <int pitch>
__device__ __forceinline__ int fast_raw_score(const uint8_t* c)
{
    const int v = c[0];
    unsigned r[16];
    r[0] = c[3 * pitch]; r[1] = c[3 * pitch + 1]; r[2] = c[2 * pitch + 2]; r[3] = c[pitch + 3];
    r[4] = c[3]; r[5] = c[-pitch + 3]; r[6] = c[-2 * pitch + 2]; r[7] = c[-3 * pitch + 1];
    r[8] = c[-3 * pitch]; r[9] = c[-3 * pitch - 1]; r[10] = c[-2 * pitch - 2]; r[11] = c[-pitch - 3];
    r[12] = c[-3]; r[13] = c[pitch - 3]; r[14] = c[2 * pitch - 2]; r[15] = c[3 * pitch - 1];
    unsigned E[16];
#pragma unroll
    for (int j = 0; j < 16; j++) E[j] = r[j] * 0xFFFF0001u + 0x00FF0000u;          // r + ((255 - r) << 16): no borrow, r <= 255
    unsigned T[16], W[16];
#pragma unroll
    for (int j = 0; j < 16; j++) T[j] = vmin3s2(E[j], E[(j + 1) & 15], E[(j + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) W[k] = vmin3s2(T[k], T[(k + 3) & 15], T[(k + 6) & 15]);
    const unsigned m = vmax3s2(vmax3s2(vmax3s2(W[0], W[1], W[2]), vmax3s2(W[3], W[4], W[5]), vmax3s2(W[6], W[7], W[8])),
                               vmax3s2(vmax3s2(W[9], W[10], W[11]), vmax3s2(W[12], W[13], W[14]), W[15]), W[15]);
    // low half: max over arcs of the arc minimum (bright); high half: 255 - (min over arcs of the arc maximum) (dark)
    const int b = (int)(m & 0xFFFF) - v;
    const int d = v - 255 + (int)(m >> 16);
    return max(b, d);
}

// shared memory of one warp: tile (rows of TP bytes) | score map | survivor list
__host__ __device__ inline int fast_warp_smem(int wCell, int hCell, int tp)
{
    const int tile = (hCell + 6) * tp;
    const int sc = ((hCell + 2) * ((wCell + 2 + 3) & ~3) + 15) & ~15;
    const int list = ((wCell * hCell * 2) + 15) & ~15;
    return tile + sc + list;
}

size_t fast_smem_bytes(const Plan& P)
{
    int m = 0;
    for (int l = 0; l < P.nlevels; l++) m = max(m, fast_warp_smem(P.lv[l].wCell, P.lv[l].hCell, P.fastPitch));
    return (size_t)((m + 127) & ~127) * FAST_WARPS;        // every warp's tile starts on a 128-byte boundary (TMA destination)
}

// Cell table (FAST_CELL_REC uint4 per cell, per image shape; built by fast_cell_record on the host).  Everything a warp needs
// that does not depend on the frame or the lane is a ready-made word here, so the kernel's set-up is loads, not arithmetic:
//   r0 = {tile origin x (iniX rounded down to 16) | iniY << 16, iw | ih << 16 (detection zone; 0 = skipped cell), level | th << 8,
//         slot index of the cell's output inside a frame's slot array}
//   r1 = {offset of the score map | offset of the survivor list << 16 (bytes from the warp's tile), scBytes | scp << 16, scBias,
//         ceil(ih / rpiZ)}
//   r2 = {recipZ | rpiZ << 20 | nd << 26, recipS | rpiS << 20 | nq << 26, xoff - dx0, dFirst | yoff << 8}
//   r3 = {bit-7 mask of the zone bytes in the FIRST 8-byte column (two words), the same for the LAST column (two words)}
// nq = 16-byte pieces per tile row, rpiS = 32 / nq tile rows staged per warp iteration (LDGSTS form); nd = 8-byte columns that
// cover the detection zone, rpiZ = 32 / nd zone rows filtered per warp iteration; recip* = 65536 / n + 1 (lane / n by
// multiply-shift); dx0 = iniX - tile origin, scp = pitch of the score map ((ih + 2) x scp bytes, 1-px zero frame), scBias: score-map
// index of tile pixel (row, col) = row * scp + col + scBias; dFirst = first 8-byte column of the zone.
void fast_cell_record(const LevelGeom& G, int level, int i, int j, int tp, uint4* rec)
{
    const int iniX = MIN_BORDER + j * G.wCell, iniY = MIN_BORDER + i * G.hCell;
    const int maxX = min(iniX + G.wCell + 6, G.maxBX), maxY = min(iniY + G.hCell + 6, G.maxBY);
    const int tw = maxX - iniX, th = maxY - iniY;
    const bool skip = iniX >= G.maxBX - 6 || iniY >= G.maxBY - 3 || tw < 7 || th < 7;
    for (int k = 0; k < FAST_CELL_REC; k++) rec[k] = make_uint4(0, 0, 0, 0);
    if (skip) return;
    const int dx0 = iniX & 15, nq = (dx0 + tw + 15) >> 4, rpiS = 32 / nq;
    const int iw = tw - 6, ih = th - 6;
    const int zx0 = dx0 + 3, zx1 = zx0 + iw, dFirst = zx0 >> 3, nd = ((zx1 - 1) >> 3) - dFirst + 1, rpiZ = 32 / nd;
    const int scp = (iw + 2 + 3) & ~3, scBytes = ((ih + 2) * scp + 15) & ~15;
    const int scOff = (G.hCell + 6) * tp;                       // the tile always has room for hCell + 6 rows (fast_warp_smem)
    auto zone_mask = [&](int col) {                             // bytes of 8-byte column `col` that lie inside [zx0, zx1)
        const int lo = min(max(zx0 - 8 * col, 0), 8), hi = min(max(zx1 - 8 * col, 0), 8);
        unsigned long long m = 0;
        for (int b = lo; b < hi; b++) m |= 0x80ull << (8 * b);
        return m;
    };
    const unsigned long long mf = zone_mask(dFirst), ml = zone_mask(dFirst + nd - 1);
    rec[0] = make_uint4((unsigned)(iniX - dx0) | ((unsigned)iniY << 16), (unsigned)iw | ((unsigned)ih << 16), (unsigned)level | ((unsigned)th << 8),
                        (unsigned)(G.slotOff + (unsigned long long)(i * G.nCols + j) * G.slotCap));
    rec[1] = make_uint4((unsigned)scOff | ((unsigned)(scOff + scBytes) << 16), (unsigned)scBytes | ((unsigned)scp << 16),
                        (unsigned)(1 - zx0 - 2 * scp), (unsigned)((ih + rpiZ - 1) / rpiZ));
    rec[2] = make_uint4((unsigned)(65536 / nd + 1) | ((unsigned)rpiZ << 20) | ((unsigned)nd << 26),       // recip <= 65537: 17 bits
                        (unsigned)(65536 / nq + 1) | ((unsigned)rpiS << 20) | ((unsigned)nq << 26), (unsigned)(j * G.wCell - dx0),
                        (unsigned)dFirst | ((unsigned)(i * G.hCell) << 8));
    rec[3] = make_uint4((unsigned)mf, (unsigned)(mf >> 32), (unsigned)ml, (unsigned)(ml >> 32));
}

// tile pitch that fits every cell of the plan: 64 bytes unless a cell row (from its 16-byte-aligned origin) is longer
int fast_tile_pitch(const Plan& P)
{
    int need = 0;
    for (int l = 0; l < P.nlevels; l++) {
        const LevelGeom& G = P.lv[l];
        for (int j = 0; j < G.nCols; j++) {
            const int iniX = MIN_BORDER + j * G.wCell, maxX = min(iniX + G.wCell + 6, G.maxBX);
            if (iniX >= G.maxBX - 6 || maxX - iniX < 7) continue;
            need = max(need, (((iniX & 15) + maxX - iniX + 15) >> 4) << 4);
        }
    }
    return need <= 64 ? 64 : (need <= 128 ? 128 : 0);
}

__device__ __forceinline__ unsigned swar_gt(unsigned a, unsigned K) { return ((a & 0x7F7F7F7Fu) + K) | a; }    // bit 7 of a byte: byte > t

template <int TP, bool TMA>
__device__ __forceinline__ void fast_cell_body(const Plan& P, const FrameBufs& F, int smemPerWarp, int f, int cell, uint8_t* smem,
                                               const LevelMaps& M, uint64_t* bars)
{
    constexpr int TPW = TP / 4;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (cell >= P.totalCells) return;
    const unsigned FULL = 0xFFFFFFFFu, ltmask = (1u << lane) - 1;
    const uint4* rec = F.cellTab + (size_t)FAST_CELL_REC * cell;
    const uint4 r0 = __ldg(&rec[0]);
    int* cntOut = F.cellCount + (size_t)f * P.totalCells + cell;
    if (r0.y == 0) {                                      // skipped cell / ROI too small for cv::FAST
        if (lane == 0) *cntOut = 0;
        return;
    }
    const uint4 r1 = __ldg(&rec[1]), r2 = __ldg(&rec[2]), r3 = __ldg(&rec[3]);
    // cell ROI (ORBextractor.cc:791-806): [iniX, iniX + tw) x [iniY, iniY + th), overlap of 6 px with the next cell; detection zone
    // iw x ih: no corner within 3 px of the ROI edge
    const int tileX = r0.x & 0xFFFF, iniY = r0.x >> 16, ih = r0.y >> 16;
    const int l = r0.z & 0xFF;
    const int scp = r1.y >> 16, scBytes = r1.y & 0xFFFF, scBias = (int)r1.z, zIters = (int)r1.w;
    const int xAdj = (int)r2.z, dFirst = r2.w & 0xFF, yoff = r2.w >> 8;
    uint8_t* tile = smem + (size_t)wid * smemPerWarp;
    uint8_t* sc = tile + (r1.x & 0xFFFF);                 // score map, behind the hCell + 6 rows of the tile
    uint16_t* list = reinterpret_cast<uint16_t*>(tile + (r1.x >> 16));
    const uint32_t* tileW = reinterpret_cast<const uint32_t*>(tile);

    const unsigned bar = (unsigned)__cvta_generic_to_shared(&bars[wid]);
    if constexpr (TMA) {
        // the whole ROI as ONE box of TP x (hCell + 6) bytes from the 16-byte boundary at or below iniX (rows below the ROI of a
        // bottom cell come along or are zero-filled outside the plane; nothing reads them)
        if (lane == 0) {
            mbar_init(bar, 1);
            mbar_fence_init();
            mbar_expect_tx(bar, r1.x & 0xFFFFu);           // (hCell + 6) * TP bytes: the box ends where the score map begins
            tma_load_box((unsigned)__cvta_generic_to_shared(tile), &M.m[l], tileX, iniY, f, bar);
        }
        __syncwarp();
    } else {
    int pitch;
    const uint8_t* img = level_plane(P, F, f, l, pitch);
    {   // global -> shared without a register round trip (LDGSTS); completion is awaited after the set-up below.
        // lane = (row r0, 16-byte piece q0) of a block of rpiS whole tile rows: no index arithmetic inside the loop
        const int nq = r2.y >> 26, rpiS = (r2.y >> 20) & 0x3F, th = r0.z >> 8;
        const int sr0 = (int)(((unsigned)lane * (r2.y & 0xFFFFFu)) >> 16), q0 = lane - sr0 * nq;
        if (sr0 < rpiS) {
            const uint8_t* p = img + (size_t)(iniY + sr0) * pitch + tileX + 16 * q0;
            unsigned sdst = (unsigned)__cvta_generic_to_shared(tile) + (unsigned)(sr0 * TP + 16 * q0);
            const size_t pstep = (size_t)rpiS * pitch;
            const unsigned sstep = (unsigned)(rpiS * TP);
            // level 0 may be the caller's device image, only 4-byte aligned: same layout from 4-byte copies
            const bool wide = l > 0 || ((((size_t)F.lv0) | F.lv0Pitch | F.lv0Frame) & 15) == 0;
            if (wide) {
                // the usual cell (38 rows, 8 or 10 rows per iteration) needs 4 or 5 copies: straight-line code for those,
                // a loop for the rows of taller cells
                constexpr int FLAT = 5;
#pragma unroll
                for (int it = 0; it < FLAT; it++)
                    if (sr0 + it * rpiS < th)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sdst + (unsigned)it * sstep), "l"(p + (size_t)it * pstep));
                if (sr0 + FLAT * rpiS < th) {
                    p += FLAT * pstep;
                    sdst += FLAT * sstep;
                    for (int r = sr0 + FLAT * rpiS; r < th; r += rpiS, p += pstep, sdst += sstep)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sdst), "l"(p));
                }
            } else {
                for (int r = sr0; r < th; r += rpiS, p += pstep, sdst += sstep) {
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst), "l"(p));
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + 4), "l"(p + 4));
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + 8), "l"(p + 8));
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + 12), "l"(p + 12));
                }
            }
        }
    }
    }
    // pre-filter work items: lane = (row zr0, 8-byte column zc) of a block of rpiZ zone rows (zone: tile rows [3, 3 + ih))
    const int nd = r2.x >> 26, rpiZ = (r2.x >> 20) & 0x3F;
    const int zr0 = (int)(((unsigned)lane * (r2.x & 0xFFFFFu)) >> 16), zc = lane - zr0 * nd;
    const bool zactive = zr0 < rpiZ;
    // bytes of the item that lie inside the zone: only the first and the last column are partial
    const unsigned zm0 = zactive ? (zc == 0 ? r3.x : 0x80808080u) & (zc == nd - 1 ? r3.z : 0x80808080u) : 0u;
    const unsigned zm1 = zactive ? (zc == 0 ? r3.y : 0x80808080u) & (zc == nd - 1 ? r3.w : 0x80808080u) : 0u;
    uint32_t* out = F.slots + (size_t)f * P.slotSlab + r0.w;
    int total = 0;
    if constexpr (TMA) mbar_wait(bar, 0);
    else asm volatile("cp.async.wait_all;" ::: "memory");

    for (int pass = 0; pass < 2; pass++) {
        const int t = pass == 0 ? P.iniTh : P.minTh;
        if (pass == 1 && (total > 0 || P.minTh == P.iniTh)) break;
        // ---- pre-filter: every lane leaves the 8 survivor bits of its item in a per-row bit mask (kept in the score-map
        // region, which is not needed yet); the ordered list is then built with one lane per row
        constexpr int RMW_MAX = TP / 32;                  // mask words per row: 8 * nd bits, nd <= TP / 8
        const int rmw = (nd + 3) >> 2, rmb = rmw * 4;
        for (int i = lane; i < ((ih * rmb + 15) >> 4); i += 32) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
        __syncwarp();
        const unsigned K = (unsigned)(127 - min(t, 127)) * 0x01010101u;     // byte > t  <=>  bit 7 of ((byte & 0x7f) + 127 - t) | byte
        {
            const uint32_t* w = tileW + (3 + zr0) * TPW + 2 * (dFirst + zc);
            uint8_t* mk = sc + zr0 * rmb + zc;
            // zIters = ceil(ih / rpiZ) from the table: a warp-uniform trip count (no division by rpiZ for the unroller)
            // not unrolled: the 4-fold unrolled loop the compiler makes by itself is 2 % slower (registers, code size)
#pragma unroll 1
            for (int it = 0, row0 = zr0; it < zIters; it++, row0 += rpiZ, w += rpiZ * TPW, mk += rpiZ * rmb)
                if (zactive && row0 < ih) {
                    const uint2 C = *reinterpret_cast<const uint2*>(w);
                    const uint2 U = *reinterpret_cast<const uint2*>(w - 3 * TPW), D = *reinterpret_cast<const uint2*>(w + 3 * TPW);
                    const unsigned Lw = w[-1], Rw = w[2];
                    const unsigned L0 = __funnelshift_r(Lw, C.x, 8), R0 = __funnelshift_r(C.x, C.y, 24);     // x - 3, x + 3
                    const unsigned L1 = __funnelshift_r(C.x, C.y, 8), R1 = __funnelshift_r(C.y, Rw, 24);
                    const unsigned v0 = swar_gt(__vabsdiffu4(D.x, C.x), K) | swar_gt(__vabsdiffu4(U.x, C.x), K);
                    const unsigned h0 = swar_gt(__vabsdiffu4(L0, C.x), K) | swar_gt(__vabsdiffu4(R0, C.x), K);
                    const unsigned v1 = swar_gt(__vabsdiffu4(D.y, C.y), K) | swar_gt(__vabsdiffu4(U.y, C.y), K);
                    const unsigned h1 = swar_gt(__vabsdiffu4(L1, C.y), K) | swar_gt(__vabsdiffu4(R1, C.y), K);
                    // bit 7 of the 8 bytes -> 8 adjacent bits: bits 8i + 4k of t8 land on bits 24 + 4k + i of the product
                    const unsigned t8 = ((v0 & h0 & zm0) >> 7) | ((v1 & h1 & zm1) >> 3);
                    *mk = (uint8_t)((t8 * 0x01020408u) >> 24);
                }
        }
        __syncwarp();
        int n = 0;
        for (int r0 = 0; r0 < ih; r0 += 32) {
            const int r = r0 + lane;
            unsigned m[RMW_MAX];
            int cnt = 0;
#pragma unroll
            for (int j = 0; j < RMW_MAX; j++) {
                m[j] = (r < ih && j < rmw) ? reinterpret_cast<const unsigned*>(sc)[r * rmw + j] : 0u;
                cnt += __popc(m[j]);
            }
            int inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int up = __shfl_up_sync(FULL, inc, o);
                if (lane >= o) inc += up;
            }
            int pos = n + inc - cnt;
            n += __shfl_sync(FULL, inc, 31);
            const int e0 = (3 + r) * TP + 8 * dFirst;     // list entry = byte offset of the pixel in the tile
#pragma unroll
            for (int j = 0; j < RMW_MAX; j++)
                for (unsigned mm = m[j]; mm; mm &= mm - 1) list[pos++] = (uint16_t)(e0 + 32 * j + __ffs(mm) - 1);
        }
        __syncwarp();
        for (int i = lane; i < (scBytes >> 4); i += 32) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
        __syncwarp();
        // ---- exact score of the survivors
        // survivors that are corners (raw > t) are compacted in place (ballot ranks keep the order): the NMS below then
        // visits corners only, about half of the survivors
        int nc = 0;
        for (int k0 = 0; k0 < n; k0 += 32) {
            const int k = k0 + lane;
            unsigned e = 0;
            bool isc = false;
            if (k < n) {
                e = list[k];
                const int raw = fast_raw_score<TP>(tile + e);
                isc = raw > t;
                if (isc) sc[(int)((e / TP) * scp + (e & (TP - 1))) + scBias] = (uint8_t)raw;
            }
            const unsigned cb = __ballot_sync(FULL, isc);            // also orders this iteration's reads before its writes
            if (isc) list[nc + __popc(cb & ltmask)] = (uint16_t)e;   // nc <= k0: never ahead of the entries still to be read
            nc += __popc(cb);
        }
        n = nc;
        __syncwarp();
        // ---- NMS (strict >, neighbours outside the zone or not corners count as 0) + emission in list order
        for (int k0 = 0; k0 < n; k0 += 32) {
            const int k = k0 + lane;
            bool keep = false;
            unsigned e = 0;
            int v = 0;
            if (k < n) {
                e = list[k];
                const uint8_t* s = sc + (int)((e / TP) * scp + (e & (TP - 1))) + scBias;
                v = s[0];
                if (v) {
                    const int m1 = max(max((int)s[-1], (int)s[1]), max((int)s[-scp - 1], (int)s[-scp]));
                    const int m2 = max(max((int)s[-scp + 1], (int)s[scp - 1]), max((int)s[scp], (int)s[scp + 1]));
                    keep = v > max(m1, m2);
                }
            }
            const unsigned kb = __ballot_sync(FULL, keep);
            if (keep) {
                // coordinates relative to (minBorderX, minBorderY): ROI-relative + j*wCell (ORBextractor.cc:822-823)
                const int x = (int)(e & (TP - 1)) + xAdj, y = (int)(e / TP) + yoff;
                out[total + __popc(kb & ltmask)] = pack_key(x, y, v - 1);
            }
            total += __popc(kb);
        }
        __syncwarp();
    }
    if (lane == 0) *cntOut = total;
}

#ifndef FAST_MINB
#define FAST_MINB 32
#endif
template <int TP, bool TMA>
__global__ void __launch_bounds__(FAST_WARPS * 32, FAST_MINB) k_fast_cells(const Plan P, const FrameBufs F, int smemPerWarp,
                                                                          const __grid_constant__ LevelMaps M)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bars[FAST_WARPS];
    fast_cell_body<TP, TMA>(P, F, smemPerWarp, blockIdx.y + F.frame0, blockIdx.x * FAST_WARPS + (threadIdx.x >> 5), smem, M, bars);
}

template <int TP, bool TMA>
static void launch_fast_cells_tp(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, size_t smemBytes, const LevelMaps& M)
{
    static SmemOptIn optin;            // small levels can have cells of up to 61 x 61 pixels
    optin.ensure(k_fast_cells<TP, TMA>, smemBytes);
    dim3 grd((P.totalCells + FAST_WARPS - 1) / FAST_WARPS, B);
    k_fast_cells<TP, TMA><<<grd, FAST_WARPS * 32, smemBytes, s>>>(P, F, (int)(smemBytes / FAST_WARPS), M);
}

// maps: TMA maps of the unblurred levels with a box of fastPitch x (hCell + 6) bytes (null: LDGSTS staging)
void launch_fast_cells(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, size_t smemBytes, const LevelMaps* maps)
{
    LevelMaps M;
    if (maps) M = *maps;
    else memset(&M, 0, sizeof M);
    if (P.fastPitch == 64) {
        if (maps) launch_fast_cells_tp<64, true>(s, P, F, B, smemBytes, M);
        else launch_fast_cells_tp<64, false>(s, P, F, B, smemBytes, M);
    } else {
        if (maps) launch_fast_cells_tp<128, true>(s, P, F, B, smemBytes, M);
        else launch_fast_cells_tp<128, false>(s, P, F, B, smemBytes, M);
    }
}

// ------------------------------------------------------------------------------------------------ quadtree
// The kernel is bound by barrier and memory latency, so many small CTAs per SM hide each other's: for the usual quotas
// (a few hundred nodes per level) 128 threads x 10 CTAs per SM with 1536 candidates in shared memory (measured per 512
// 640x480 frames: 256 x 4 CTAs 0.236 ms, 128 x 10 CTAs 0.198 ms, 64 x 10 0.30 ms); levels with thousands of candidates
// keep the 256-thread form (512 threads for 1080p, nfeatures = 4000) with 4096 candidates in shared memory.
struct OctSmall { static constexpr int THREADS = 128, MINB = 10, KEYS = 1536; };
struct OctLarge { static constexpr int THREADS = 256, MINB = 4, KEYS = 4096; };
struct OctHuge { static constexpr int THREADS = 512, MINB = 2, KEYS = 4096; };

// Exclusive scan of val(0..n) by the whole CTA: out[i] = sum of val(j), j < i; returns the total.  A thread owns a run of
// consecutive elements; every warp scans the warp totals itself, so there are two barriers: one inside and one at the end
// (which publishes out[] and frees wsum for the next scan).  val is evaluated twice per element and must not read out[].
// wsum: T / 32 ints of shared memory.
template <int T, class V>                       // T compile-time: the per-thread run is a shift, not an integer division
__device__ int block_exclusive_scan(int n, int* out, int* wsum, V val)
{
    constexpr int NW = T >> 5;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + T - 1) / T;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; i++) s += val(i);
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    const int v = lane < NW ? wsum[lane] : 0;
    int vi = v;
#pragma unroll
    for (int o = 1; o < NW; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, vi, o);
        if (lane >= o) vi += t;
    }
    const int total = __shfl_sync(0xFFFFFFFFu, vi, NW - 1);
    int off = __shfl_sync(0xFFFFFFFFu, vi - v, wid) + inc - s;
    for (int i = b; i < e; i++) { const int t = val(i); out[i] = off; off += t; }
    __syncthreads();
    return total;
}

// small form when the quotas are small (nfeatures = 1000: 229 nodes at most per level) and the launch has about a wave of
// small CTAs; otherwise a (level, frame) problem is the critical path and the larger CTA finishes it sooner.  Measured per
// launch, small / large form: 640x480 nf=1000 x 512 frames 0.134 / 0.197 ms, x 128 frames 0.054 / 0.062 ms; 752x480 nf=1200
// x 192 frames 0.111 / 0.101 ms; 1241x376 nf=2000 x 128 frames 0.113 / 0.082 ms, x 512 frames 0.281 / 0.245 ms.
static bool octree_large(const Plan& P, int B, int capS)
{
    static const int thr = getenv("ORBX_OCT_THR") ? atoi(getenv("ORBX_OCT_THR")) : 250;       // tuning override
    return capS > thr || (long long)B * P.nlevels < 1000;
}

// bytes of the per-node arrays of k_octree (the candidate arrays follow them)
__host__ __device__ inline size_t octree_node_bytes(int capS, int sortCap)
{
    return ((size_t)capS * (8 + 8 + 4 + 4 + 16 + 4 + 4 + 2 + 2 + 2 + 8) + (size_t)sortCap * 4 + 15) & ~(size_t)15;
}

// shared memory of the larger form (the limit check at plan time); launch_octree sizes the launch itself
size_t octree_smem_bytes(int capS, int sortCap, int keysCap)
{
    const int OCT_KEYS_SMEM = keysCap > 0 ? keysCap : OctLarge::KEYS;
    // rectA, rectB (short4) | cntA, cntB | cc[4] | chBeg | tmp | skey[sortCap] | rankOf[2], mapS (u16) | childPos[4] (u16)
    // ... | keys (u32) and node positions (u16) of up to OCT_KEYS_SMEM candidates
    return octree_node_bytes(capS, sortCap) + (size_t)OCT_KEYS_SMEM * 6 + 64;
}

// DistributeOctTree.  The reference keeps nodes in a std::list built with push_front; a pass over the list splits
// nodes and the new list is  reverse(children in creation order) ++ (unsplit nodes in their old order).  Here the list
// is an array in list order, rebuilt every pass with prefix sums; every candidate carries the position of its node.
// Tie-break of the "largest first" phase (ORBextractor.cc:684 sorts by (size, heap address)): creation order, later
// first == smaller list position first (the portable definition of the oracle, tie_mode 0).
template <class CFG>
__global__ void __launch_bounds__(CFG::THREADS, CFG::MINB) k_octree(const Plan P, const FrameBufs F, int capS, int sortCap)
{
    constexpr int OCT_KEYS_SMEM = CFG::KEYS;
    extern __shared__ __align__(16) uint8_t smraw[];
    __shared__ int wsum[32];
    __shared__ int s_k[2], s_nexp[2], s_m;      // [pass parity]: a pass never resets what a slower warp of the previous pass still reads
    constexpr int T = CFG::THREADS;
    const int l = blockIdx.x, f = blockIdx.y + F.frame0, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const LevelGeom& G = P.lv[l];

    short4* rA = reinterpret_cast<short4*>(smraw);
    short4* rB = rA + capS;
    int* cA = reinterpret_cast<int*>(rB + capS);
    int* cB = cA + capS;
    int* cc = cB + capS;                   // child counts of the nodes being split, by processing rank
    int* chBeg = cc + 4 * capS;            // creation index of a rank's first child
    int* tmp = chBeg + capS;
    uint32_t* skey = reinterpret_cast<uint32_t*>(tmp + capS);   // processing order (positions) / sort keys
    uint16_t* rankOf2 = reinterpret_cast<uint16_t*>(skey + sortCap);      // processing rank of a list position, per pass parity
    uint16_t* mapS = rankOf2 + 2 * capS;
    uint16_t* childPos = mapS + capS;

    const int ncell = G.nCols * G.nRows;
    const int* cellCount = F.cellCount + (size_t)f * P.totalCells + G.cellBase;
    const uint32_t* slots = F.slots + (size_t)f * P.slotSlab + G.slotOff;
    uint32_t* keys = F.keys + (size_t)f * P.keySlab + G.keyOff;
    uint16_t* knode = F.knode + (size_t)f * P.keySlab + G.keyOff;
    uint32_t* keysS = reinterpret_cast<uint32_t*>(smraw + octree_node_bytes(capS, sortCap));
    uint16_t* knodeS = reinterpret_cast<uint16_t*>(keysS + OCT_KEYS_SMEM);
    uint32_t* outKeys = F.levelKeys + (size_t)f * P.totalNodes + G.nodeBase;
    int* outCount = F.levelCount + (size_t)f * MAX_LEVELS + l;

    // candidate order of the reference: cell row, cell column, then cv::FAST's row-major order inside the cell.  A thread owns
    // a run of consecutive cells: it sums their counts, a block scan of the sums gives the position of its first candidate,
    // and it copies its cells' candidates there (cells hold ~5 candidates each: a warp per cell left most lanes idle and the
    // dependent count -> candidates loads of 273 cells in a row on 4 warps were the longest phase of the kernel).
    const int nIni = G.nIni;
    int* rootCnt = cc;          // scratch: nIni <= capS
    for (int i = tid; i < nIni; i += T) rootCnt[i] = 0;
    const int per = (ncell + T - 1) / T;
    const int c0 = min(tid * per, ncell), c1 = min(c0 + per, ncell);
    int mine = 0;
    for (int c = c0; c < c1; c++) mine += __ldg(&cellCount[c]);
    int inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    int n, at;
    {   // every warp scans the warp totals itself
        constexpr int NW = T >> 5;
        const int v = lane < NW ? wsum[lane] : 0;
        int vi = v;
#pragma unroll
        for (int o = 1; o < NW; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, vi, o);
            if (lane >= o) vi += t;
        }
        n = __shfl_sync(0xFFFFFFFFu, vi, NW - 1);
        at = __shfl_sync(0xFFFFFFFFu, vi - v, wid) + inc - mine;
    }
    if (n == 0) {
        if (tid == 0) *outCount = 0;
        return;
    }
    if (n <= OCT_KEYS_SMEM) { keys = keysS; knode = knodeS; }
    for (int c = c0; c < c1; c++) {
        const int cnt = __ldg(&cellCount[c]);
        const uint32_t* src = slots + (size_t)c * G.slotCap;
        for (int e0 = 0; e0 < cnt; e0 += 4) {
            uint32_t kk[4];
#pragma unroll
            for (int j = 0; j < 4; j++) kk[j] = e0 + j < cnt ? __ldg(&src[e0 + j]) : 0u;       // four loads in flight
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (e0 + j < cnt) {
                    int root = 0;                                             // vpIniNodes[kp.pt.x/hX] (:569); one root: no division
                    if (nIni > 1) {
                        root = min((int)__fdiv_rn((float)key_x(kk[j]), G.hX), nIni - 1);
                        atomicAdd(&rootCnt[root], 1);
                    }
                    keys[at] = kk[j];
                    knode[at] = (uint16_t)root;
                    at++;
                }
        }
    }
    int m;
    if (nIni == 1) {                        // one root (every shape up to 1.5 : 1): it holds all n > 0 candidates
        if (tid == 0) {
            rA[0] = make_short4((short)0, (short)(int)G.hX, (short)0, (short)(G.maxBY - MIN_BORDER));
            cA[0] = n;
        }
        m = 1;
        __syncthreads();                    // wsum (the setup scan above) is reused by the scans below
    } else {
        __syncthreads();
        if (wid == 0) {                     // empty roots are erased (:574-585): a ballot per 32 roots ranks the others
            int m0 = 0;
            for (int i0 = 0; i0 < nIni; i0 += 32) {
                const int i = i0 + lane;
                const int rc = i < nIni ? rootCnt[i] : 0;
                const unsigned live = __ballot_sync(0xFFFFFFFFu, rc > 0);
                if (rc > 0) {
                    const int q = m0 + __popc(live & ((1u << lane) - 1));
                    rA[q] = make_short4((short)(int)__fmul_rn(G.hX, (float)i), (short)(int)__fmul_rn(G.hX, (float)(i + 1)),
                                        (short)0, (short)(G.maxBY - MIN_BORDER));
                    cA[q] = rc;
                    tmp[i] = q;
                }
                m0 += __popc(live);
            }
            if (lane == 0) s_m = m0;
        }
        __syncthreads();
        // the candidates were written in runs of cells, one run per thread: this pass and every later one index them tid, tid + T, ...
        for (int i = tid; i < n; i += T) knode[i] = (uint16_t)tmp[knode[i]];
        m = s_m;
        __syncthreads();                    // tmp is the output of the first scan below
    }
    const int N = G.quota;
    bool lf = false;

    // Barriers per pass: 9 (10 in the largest-first phase, plus the sort's).  knode[i] is only ever touched by thread i mod T.
    for (int pass = 0;; pass++) {
        const int par = pass & 1;
        uint16_t* rankOf = rankOf2 + par * capS;
        // ---- processing order P: all expandable nodes (more than one key)
        if (tid == 0) { s_k[par] = 0x7FFFFFFF; s_nexp[par] = 0; }
        const int np = block_exclusive_scan<T>(m, tmp, wsum, [&](int p) { rankOf[p] = 0xFFFF; return (int)(cA[p] > 1); });
        if (np == 0) break;                                   // nothing to split: list size cannot change
        if (!lf) {
            for (int p = tid; p < m; p += T)
                if (cA[p] > 1) { skey[tmp[p]] = (uint32_t)p; rankOf[p] = (uint16_t)tmp[p]; }      // list order (:606)
        } else {
            int sp = 1;
            while (sp < np) sp <<= 1;
            for (int p = tid; p < m; p += T) if (cA[p] > 1) skey[tmp[p]] = ((uint32_t)cA[p] << 12) | (uint32_t)(4095 - p);
            for (int r = np + tid; r < sp; r += T) skey[r] = 0;
            for (int size = 2; size <= sp; size <<= 1)
                for (int stride = size >> 1; stride > 0; stride >>= 1) {
                    __syncthreads();
                    for (int i = tid; i < sp; i += T) {
                        const int j = i ^ stride;
                        if (j > i) {
                            const uint32_t a = skey[i], b = skey[j];
                            const bool desc = (i & size) == 0;
                            if ((a < b) == desc) { skey[i] = b; skey[j] = a; }
                        }
                    }
                }
            __syncthreads();
            for (int r = tid; r < np; r += T) {               // thread r alone reads and writes skey[r]
                const uint32_t pos = 4095u - (skey[r] & 4095u);
                skey[r] = pos;
                rankOf[pos] = (uint16_t)r;
            }
        }
        for (int i = tid; i < 4 * np; i += T) cc[i] = 0;
        __syncthreads();

        // ---- quadrant of every key of an expandable node (DivideNode :481-537), child counts
        for (int i0 = 0; i0 < n; i0 += T) {
            const int i = i0 + tid;
            int code = -1;
            if (i < n) {
                const int pos = knode[i] & 0xFFF;
                const int r = rankOf[pos];
                if (r != 0xFFFF) {
                    const uint32_t key = keys[i];
                    const short4 rc = rA[pos];
                    const int midX = rc.x + ((rc.y - rc.x + 1) >> 1), midY = rc.z + ((rc.w - rc.z + 1) >> 1);
                    const int q = (key_x(key) < midX ? 0 : 1) + (key_y(key) < midY ? 0 : 2);
                    knode[i] = (uint16_t)(pos | (q << 12));
                    code = 4 * r + q;
                }
            }
            const unsigned peers = __match_any_sync(0xFFFFFFFFu, code);
            if (code >= 0 && lane == __ffs(peers) - 1) atomicAdd(&cc[code], __popc(peers));
        }
        __syncthreads();

        // ---- children per rank, creation indices, early stop of the largest-first phase (:730-731)
        const int totalCh = block_exclusive_scan<T>(np, chBeg, wsum, [&](int r) {
            return (int)(cc[4 * r] > 0) + (int)(cc[4 * r + 1] > 0) + (int)(cc[4 * r + 2] > 0) + (int)(cc[4 * r + 3] > 0);
        });
        int k = np;
        if (lf) {
            for (int r = tid; r < np; r += T) {
                const int end = (r + 1 < np) ? chBeg[r + 1] : totalCh;
                if (m + end - (r + 1) >= N) atomicMin(&s_k[par], r);
            }
            __syncthreads();
            if (s_k[par] != 0x7FFFFFFF) k = s_k[par] + 1;
        }
        const int C = (k == np) ? totalCh : chBeg[k];

        // ---- new list = reverse(children) ++ unsplit nodes
        block_exclusive_scan<T>(m, tmp, wsum, [&](int p) { return (int)((int)rankOf[p] >= k); });
        for (int p = tid; p < m; p += T)
            if ((int)rankOf[p] >= k) {
                const int q = C + tmp[p];
                mapS[p] = (uint16_t)q;
                rB[q] = rA[p];
                cB[q] = cA[p];
            }
        int nexp = 0;
        for (int r = tid; r < k; r += T) {
            const int pos = skey[r];
            const short4 rc = rA[pos];
            const int midX = rc.x + ((rc.y - rc.x + 1) >> 1), midY = rc.z + ((rc.w - rc.z + 1) >> 1);
            int c = chBeg[r];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int cq = cc[4 * r + q];
                if (cq > 0) {
                    const int at = C - 1 - c;
                    c++;
                    rB[at] = make_short4((q & 1) ? (short)midX : rc.x, (q & 1) ? rc.y : (short)midX,
                                         (q & 2) ? (short)midY : rc.z, (q & 2) ? rc.w : (short)midY);
                    cB[at] = cq;
                    childPos[4 * r + q] = (uint16_t)at;
                    nexp += cq > 1;
                }
            }
        }
        if (nexp) atomicAdd(&s_nexp[par], nexp);
        __syncthreads();
        for (int i = tid; i < n; i += T) {
            const int kn = knode[i];
            const int pos = kn & 0xFFF, q = kn >> 12;
            const int r = rankOf[pos];
            knode[i] = (r < k) ? childPos[4 * r + q] : mapS[pos];
        }
        const int m2 = C + (m - k);
        const bool done = (m2 >= N) || (m2 == m);
        if (!done && !lf && m2 + 3 * s_nexp[par] > N) lf = true;
        { short4* t = rA; rA = rB; rB = t; }
        { int* t = cA; cA = cB; cB = t; }
        m = m2;
        if (done) break;
        // no barrier here: the next pass writes the other rankOf / s_k / s_nexp, tmp, skey and cc, none of which the loop above
        // reads, and the arrays it reads first (cA, rA) were complete at the barrier before that loop
    }
    __syncthreads();

    // ---- best key of every node: highest response, first in candidate order (:742-760)
    unsigned* best = reinterpret_cast<unsigned*>(tmp);
    for (int p = tid; p < m; p += T) best[p] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += T)
        atomicMax(&best[knode[i] & 0xFFF], ((unsigned)key_r(keys[i]) << 20) | (unsigned)(0xFFFFF - i));
    __syncthreads();
    for (int p = tid; p < m; p += T) outKeys[p] = keys[0xFFFFF - (best[p] & 0xFFFFF)];
    if (tid == 0) *outCount = m;
}

template <class CFG>
static cudaError_t launch_octree_cfg(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, int capS, int sortCap)
{
    const size_t smemBytes = octree_smem_bytes(capS, sortCap, CFG::KEYS);
    static SmemOptIn optin;
    optin.ensure(k_octree<CFG>, smemBytes);
    if (cudaError_t e = cudaGetLastError()) return e;
    dim3 grd(P.nlevels, B);
    k_octree<CFG><<<grd, CFG::THREADS, smemBytes, s>>>(P, F, capS, sortCap);
    return cudaSuccess;
}

cudaError_t launch_octree(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, int capS, int sortCap, int form)
{
    // quotas of many hundreds of nodes (1080p, nfeatures = 4000): level 0 of one frame is the critical path of the launch, and
    // 512 threads finish it sooner (64 frames: 0.44 -> 0.27 ms; 1024 threads: no further gain)
    static const int hthr = getenv("ORBX_OCT_HUGE") ? atoi(getenv("ORBX_OCT_HUGE")) : 600;
    // a handful of frames (the reference's one frame per call): the grid is far smaller than the GPU and level 0 of a frame
    // is the critical path of the whole call, so the widest CTA wins (one 640x480 frame per call: 0.186 -> 0.165 ms)
    if (form == 0) form = (capS > hthr || B * P.nlevels <= 64) ? 3 : (octree_large(P, B, capS) ? 2 : 1);
    if (form == 3) return launch_octree_cfg<OctHuge>(s, P, F, B, capS, sortCap);
    if (form == 2) return launch_octree_cfg<OctLarge>(s, P, F, B, capS, sortCap);
    return launch_octree_cfg<OctSmall>(s, P, F, B, capS, sortCap);
}

// ------------------------------------------------------------------------------------------------ Gaussian blur
// cv::GaussianBlur(7x7, sigma 2) on u8, OpenCV 4.x fixed-point path: taps 18,34,48,56,48,34,18 / 256; the first pass is
// exact 8.8 (no rounding), the second 8.16, out = (acc + 32768) >> 16; BORDER_REFLECT_101 of the level itself (the
// reference blurs a border-less clone, ORBextractor.cc:1085).
// Both passes run on the dot-product instructions (IDP.4A / IDP.2A, 64 lanes/clk/SM beside the ALU pipe, see
// tools/int_pipe_bench.cu) instead of unpack + multiply-add chains.  A lane owns 4 adjacent columns (one 32-bit word) and
// walks down a strip of rows:
//   horizontal: the 7 taps of each of the 4 pixels lie in the words L | C | R (left neighbour, own, right neighbour, all
//               loaded straight from L1): 10 IDP.4A against constant tap words give h0..h3 (<= 65280, exact);
//   pairing:    P(r) = h(r-1) | h(r) << 16 per column (one IMAD each): a sliding window of the last 6 pairs per column
//               lives in registers;
//   vertical:   out(y) = 18 h(y-3) + 34 h(y-2) + 48 h(y-1) + 56 h(y) + 48 h(y+1) + 34 h(y+2) + 18 h(y+3) + 32768
//                      = IDP.2A(P(y-2),(18,34)) + IDP.2A(P(y),(48,56)) + IDP.2A(P(y+2),(48,34)) + IDP.2A(P(y+3),(0,18)),
//               and the result byte is bits 16..23 of the accumulator (one PRMT tree packs the 4 pixels).
constexpr int BLUR_ROWS = 60;        // rows per strip (+ 6 halo rows = 11 blocks of 6)
constexpr int BLUR_WARPS = 8;
#ifndef BLUR_MINB
#define BLUR_MINB 2
#endif

__device__ __forceinline__ int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// single reflection + clamp: exact for -len < p < 2*len - 1, which covers every coordinate whose value is actually used
__device__ __forceinline__ int reflect1(int p, int len)
{
    p = p < 0 ? -p : p;
    p = p >= len ? 2 * (len - 1) - p : p;
    return min(max(p, 0), len - 1);
}

#define TAPS4(a, b, c, d) ((unsigned)(a) | ((unsigned)(b) << 8) | ((unsigned)(c) << 16) | ((unsigned)(d) << 24))

// Border lanes (BORDER_REFLECT_101 in x).  Only bytes 1..3 of L (columns x-3..x-1) and bytes 0..2 of R (x+4..x+6) carry
// non-zero taps.  A lane loads L at pc-4 (not when x == 0), C at pc, and a third word at pc+4 -- or, for the last lane of a
// row whose width is 1 mod 4, at pc-8 (two of its reflected columns live there).  Border lanes then rebuild the three
// words with four PRMTs whose selectors are derived here, per lane, once:
//   FL = prmt(L, C, sL), FC = prmt(C, L, sC), X = prmt(L, C, s1), FR = prmt(X, R, s2).
struct BlurEdge { unsigned sL, sC, s1, s2; int dR; bool loadL, loadR, edge; };
__device__ __forceinline__ BlurEdge blur_edge(int x, int w)
{
    BlurEdge e;
    e.loadL = x > 0;
    e.dR = 4;
    e.loadR = x + 4 < w;
    e.sL = 0x3210u; e.sC = 0x3210u; e.s1 = 0x3210u; e.s2 = 0x7654u;     // identity: FL = L, FC = C, FR = R
    e.edge = false;
    if (x == 0) { e.sL = 0x5670u; e.edge = true; }                        // columns -3..-1 = pixels 3, 2, 1 = C bytes 3, 2, 1
    if (x + 8 > w) {                                                      // some of columns x .. x+6 lie beyond the row
        e.edge = true;
        if (x + 4 >= w && (w & 3) == 1) { e.dR = -8; e.loadR = true; }    // last lane, one valid pixel: third word = columns x-8..x-5
        unsigned sC = 0, s1 = 0, s2 = 0;
        for (int k = 0; k < 7; k++) {                                     // column x + k
            int c = x + k;
            const bool valid = c < w;
            if (!valid) c = 2 * (w - 1) - c;
            const int rel = c - x;                                        // byte position relative to the own word
            if (k < 4) {                                                  // FC byte k from C (rel 0..3) or L (rel -4..-1): prmt(C, L)
                sC |= (unsigned)(rel >= 0 ? rel : 8 + rel) << (4 * k);
            } else {                                                      // FR byte k-4
                const int b = k - 4;
                if (rel >= 4) s2 |= (unsigned)(4 + rel - 4) << (4 * b);                      // from the third word (= R) directly
                else if (rel >= -4) { s1 |= (unsigned)(rel >= 0 ? 4 + rel : 4 + rel) << (4 * b); s2 |= (unsigned)b << (4 * b); }
                else { s2 |= (unsigned)(4 + rel + 8) << (4 * b); }                           // columns x-8..x-5: third word loaded at pc-8
            }
        }
        // s1 gathers from prmt(L, C): L bytes are 0..3 (rel -4..-1 -> 4 + rel), C bytes are 4..7 (rel 0..3 -> 4 + rel)
        e.sC = sC; e.s1 = s1; e.s2 = s2 | 0x7000u;
    }
    return e;
}

// One strip of rows [y0, y1) of one lane's 4 columns.  INTERIOR: the strip needs no vertical reflection (rows y0-3 ..
// y1+2 all exist), so the source pointers simply advance by one pitch per row.
// PLAIN: no lane of the warp touches the left or right image border (every lane loads L, C, R as they are): the border
// selectors and their predicated PRMTs drop out.
template <bool INTERIOR, bool PLAIN>
__device__ __forceinline__ void blur_strip(const uint8_t* __restrict__ img, int pitch, int h, int x, int y0, int y1, const BlurEdge E,
                                           uint8_t* __restrict__ dst, int dpitch)
{
    unsigned Pw[6][4];                  // pairs P(r-5) .. P(r), slot = r mod 6 (statically indexed: the loop is unrolled by 6)
#pragma unroll
    for (int k = 0; k < 4; k++) Pw[5][k] = 0;
    int rr = INTERIOR ? y0 - 3 : reflect1(y0 - 3, h);
    const uint8_t* pc = img + (size_t)(unsigned)rr * (unsigned)pitch + x;
    const uint8_t* pr = pc + (PLAIN ? 4 : E.dR);
    // iteration i handles source row r = y0 - 3 + i and, from i = 6 on, emits output row y = r - 3
#define BLUR_ROW(J, EMIT)                                                                                             \
    {                                                                                                                 \
        unsigned L = 0, C, R = 0;                                                                                     \
        if (PLAIN || E.loadL) L = __ldg(reinterpret_cast<const unsigned*>(pc - 4));                                   \
        C = __ldg(reinterpret_cast<const unsigned*>(pc));                                                             \
        if (PLAIN || E.loadR) R = __ldg(reinterpret_cast<const unsigned*>(pr));                                       \
        if (!PLAIN && E.edge) {                                                                                       \
            const unsigned X = __byte_perm(L, C, E.s1), FL = __byte_perm(L, C, E.sL), FC = __byte_perm(C, L, E.sC);   \
            R = __byte_perm(X, R, E.s2); L = FL; C = FC;                                                              \
        }                                                                                                             \
        if (INTERIOR) { pc += pitch; pr += pitch; }                                                                   \
        else {                                                                                                        \
            const int nr = reflect1(rbase + (J) + 1, h);                                                              \
            pc += (long long)(nr - rr) * pitch; pr += (long long)(nr - rr) * pitch; rr = nr;                          \
        }                                                                                                             \
        const unsigned h0 = __dp4a(C, TAPS4(56, 48, 34, 18), __dp4a(L, TAPS4(0, 18, 34, 48), 0u));                    \
        const unsigned h1 = __dp4a(R, TAPS4(18, 0, 0, 0), __dp4a(C, TAPS4(48, 56, 48, 34), __dp4a(L, TAPS4(0, 0, 18, 34), 0u))); \
        const unsigned h2 = __dp4a(R, TAPS4(34, 18, 0, 0), __dp4a(C, TAPS4(34, 48, 56, 48), __dp4a(L, TAPS4(0, 0, 0, 18), 0u))); \
        const unsigned h3 = __dp4a(R, TAPS4(48, 34, 18, 0), __dp4a(C, TAPS4(18, 34, 48, 56), 0u));                    \
        /* slot J receives P(r) = (h(r-1), h(r)): low half = high half of P(r-1) (slot J-1), high half = h(r) */      \
        Pw[J][0] = __byte_perm(h0, Pw[((J) + 5) % 6][0], 0x1076);                                                     \
        Pw[J][1] = __byte_perm(h1, Pw[((J) + 5) % 6][1], 0x1076);                                                     \
        Pw[J][2] = __byte_perm(h2, Pw[((J) + 5) % 6][2], 0x1076);                                                     \
        Pw[J][3] = __byte_perm(h3, Pw[((J) + 5) % 6][3], 0x1076);                                                     \
        if (EMIT) {                                                                                                   \
            unsigned o[4];                                                                                            \
            _Pragma("unroll") for (int k = 0; k < 4; k++) {                                                           \
                /* P(y+3) = P(r) slot J, P(y+2) = P(r-1) slot J-1, P(y) = P(r-3) slot J-3, P(y-2) = P(r-5) slot J-5 */ \
                unsigned acc = __dp2a_lo(Pw[((J) + 1) % 6][k], TAPS4(18, 34, 0, 0), 32768u);                          \
                acc = __dp2a_lo(Pw[((J) + 3) % 6][k], TAPS4(48, 56, 0, 0), acc);                                      \
                acc = __dp2a_lo(Pw[((J) + 5) % 6][k], TAPS4(48, 34, 0, 0), acc);                                      \
                o[k] = __dp2a_lo(Pw[J][k], TAPS4(0, 18, 0, 0), acc);                                                  \
            }                                                                                                         \
            /* byte 2 of each accumulator */                                                                          \
            if (rbase + (J) - 3 < y1)                                                                                 \
                *reinterpret_cast<unsigned*>(dst) =                                                                   \
                    __byte_perm(__byte_perm(o[0], o[1], 0x0062), __byte_perm(o[2], o[3], 0x0062), 0x5410);            \
            dst += dpitch;                                                                                            \
        }                                                                                                             \
    }
    int rbase = y0 - 3;
    BLUR_ROW(0, false) BLUR_ROW(1, false) BLUR_ROW(2, false) BLUR_ROW(3, false) BLUR_ROW(4, false) BLUR_ROW(5, false)
    for (rbase += 6; rbase - 3 < y1; rbase += 6) {
        BLUR_ROW(0, true) BLUR_ROW(1, true) BLUR_ROW(2, true) BLUR_ROW(3, true) BLUR_ROW(4, true) BLUR_ROW(5, true)
    }
#undef BLUR_ROW
}

// A warp takes 32 consecutive (strip, word) items of a level, word fastest: every lane works whatever the level width (128-column
// tiles left 11 % of the lanes idle over the pyramid of a 640 x 480 frame); a warp whose lanes straddle two strips, or touch a
// border, takes the general form of the strip loop.
__device__ __forceinline__ void blur_tile_body(const Plan& P, const FrameBufs& F, int f, int tile)
{
    const int lane = threadIdx.x & 31;
    if (tile >= P.totalBlurTiles) return;
    int l = 0;
    while (l + 1 < P.nlevels && tile >= P.lv[l + 1].blurTileBase) ++l;
    const LevelGeom& G = P.lv[l];
    const int w = G.w, h = G.h, nw = G.blurTilesX;
    const int item = (tile - G.blurTileBase) * 32 + lane;
    const bool active = item < nw * ((h + BLUR_ROWS - 1) / BLUR_ROWS);
    const int strip = item / nw, x = (item - strip * nw) * 4;                  // first of this lane's 4 columns
    const int y0 = strip * BLUR_ROWS, y1 = min(y0 + BLUR_ROWS, h);
    // warp-uniform choice of the strip loop: all lanes interior in y / no lane on the left or right image border
    const bool interior = __all_sync(0xFFFFFFFFu, !active || (y0 >= 3 && y1 + 3 <= h));
    const bool plain = __all_sync(0xFFFFFFFFu, !active || (x > 0 && x + 8 <= w));
    if (!active) return;                                                        // no warp-wide operation below
    int pitch;
    const uint8_t* img = level_plane(P, F, f, l, pitch);
    const BlurEdge E = blur_edge(x, w);
    uint8_t* dst = F.blur + (size_t)f * P.blurSlab + G.blurOff + (size_t)(unsigned)y0 * (unsigned)G.pitch + x;
    if (interior) {
        if (plain) blur_strip<true, true>(img, pitch, h, x, y0, y1, E, dst, G.pitch);
        else blur_strip<true, false>(img, pitch, h, x, y0, y1, E, dst, G.pitch);
    } else blur_strip<false, false>(img, pitch, h, x, y0, y1, E, dst, G.pitch);
}

#undef TAPS4

__global__ void __launch_bounds__(BLUR_WARPS * 32, BLUR_MINB) k_blur(const Plan P, const FrameBufs F)
{
    blur_tile_body(P, F, blockIdx.y + F.frame0, blockIdx.x * BLUR_WARPS + (threadIdx.x >> 5));
}

void launch_blur(cudaStream_t s, const Plan& P, const FrameBufs& F, int B)
{
    dim3 grd((P.totalBlurTiles + BLUR_WARPS - 1) / BLUR_WARPS, B);
    k_blur<<<grd, BLUR_WARPS * 32, 0, s>>>(P, F);
}

// ------------------------------------------------------------------------------------------------ orientation + rBRIEF
// cv::fastAtan2 (degrees): float32 odd polynomial, every operation individually rounded.
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

constexpr int DESC_WARPS = 8;
constexpr int DESC_PATCH_ROWS = 37, DESC_PATCH_WPITCH = 12;      // blurred footprint in shared memory: 37 rows x 12 words (11 used)

// cvRound without the (quarter-rate) F2I: adding 1.5 * 2^23 rounds to nearest-even in the FADD itself and leaves the
// integer in the low mantissa bits; exact for |v| < 2^22.
__device__ __forceinline__ int round_magic_bits(float v) { return __float_as_int(__fadd_rn(v, 12582912.f)); }
constexpr int MAGIC_BITS = 0x4B400000;

// Shared memory of one warp in the TMA form.  A tiled TMA box must START on a 16-byte boundary of the row (an unaligned inner
// coordinate raises "illegal instruction"; measured with tools/tma_probe.cu), so each box begins at x rounded down to 16 and is wide
// enough for the slack: the 37-column descriptor footprint travels as 64 x 37, the 31-column orientation patch as 48 x 31.
constexpr int DESC_BOX_BLUR_W = 64, DESC_BOX_IMG_W = 48;
constexpr int DESC_BOX_BLUR_BYTES = DESC_BOX_BLUR_W * DESC_PATCH_ROWS, DESC_BOX_IMG_BYTES = DESC_BOX_IMG_W * 31;      // 2368, 1488
constexpr int DESC_BOX_IMG_OFF = 2432, DESC_BOX_STRIDE = 2432 + 1536;                                                  // 128-byte multiples

struct OrientMaps { LevelMaps img, blur; };

// TMA = true: both patches of a keypoint arrive by one cp.async.bulk.tensor each (one lane issues them, completion on two
// mbarriers), whatever their alignment; TMA = false: per-lane LDGSTS / LDG of aligned words, for a caller-owned level 0 whose base
// or pitch is not a multiple of 16 bytes.
template <bool TMA>
__global__ void __launch_bounds__(DESC_WARPS * 32) k_orient_describe(const Plan P, const FrameBufs F, const uint2* __restrict__ icTab,
                                                                    const float4* __restrict__ patT, const __grid_constant__ OrientMaps M)
{
    const int f = blockIdx.y + F.frame0, lane = threadIdx.x & 31;
    const int g = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);      // keypoint index inside the frame
    // levels are concatenated in ascending order (:1103): lane i holds the count of level i, a warp scan finds the level
    const int myc = lane < P.nlevels ? F.levelCount[(size_t)f * MAX_LEVELS + lane] : 0;
    int inc = myc;
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    const int total = __shfl_sync(0xFFFFFFFFu, inc, MAX_LEVELS - 1 < 31 ? 15 : 31);
    if (g == 0 && lane == 0) F.counts[f] = total;
    if (g >= total) return;
    const unsigned below = __ballot_sync(0xFFFFFFFFu, inc <= g);        // levels that end at or before g
    const int l = __popc(below & 0xFFFFu);
    const int idx = g - __shfl_sync(0xFFFFFFFFu, inc - myc, l);
    const LevelGeom& G = P.lv[l];
    const uint32_t key = F.levelKeys[(size_t)f * P.totalNodes + G.nodeBase + idx];
    const int cx = key_x(key) + MIN_BORDER, cy = key_y(key) + MIN_BORDER;      // level coordinates (:841-842)

    // The 37 x 37 footprint of the descriptor (rotated pattern reach <= 18) on the BLURRED level goes to shared memory as
    // 11 aligned words per row, copied by LDGSTS while the orientation is computed: the 512 test bytes are then shared-memory
    // reads instead of scattered global byte loads (which had the kernel bound by L1 wavefronts).
    __shared__ __align__(128) uint8_t s_box[TMA ? DESC_WARPS * DESC_BOX_STRIDE : 16];
    __shared__ uint64_t s_bar[DESC_WARPS][2];
    __shared__ uint32_t s_patch[TMA ? 1 : DESC_WARPS][TMA ? 1 : DESC_PATCH_ROWS * DESC_PATCH_WPITCH];
    const int bp = G.pitch;
    const int px0 = cx - 18, pal = px0 & (TMA ? 15 : 3);
    const unsigned boxBase = (unsigned)__cvta_generic_to_shared(s_box) + (unsigned)((threadIdx.x >> 5) * DESC_BOX_STRIDE);
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&s_bar[threadIdx.x >> 5][0]);
    if constexpr (TMA) {
        if (lane == 0) {
            mbar_init(bar0, 1);
            mbar_init(bar0 + 8, 1);
            mbar_fence_init();
            mbar_expect_tx(bar0, DESC_BOX_IMG_BYTES);
            tma_load_box(boxBase + DESC_BOX_IMG_OFF, &M.img.m[l], (cx - HALF_PATCH) & ~15, cy - HALF_PATCH, f, bar0);
            mbar_expect_tx(bar0 + 8, DESC_BOX_BLUR_BYTES);
            tma_load_box(boxBase, &M.blur.m[l], px0 - pal, cy - 18, f, bar0 + 8);
        }
        __syncwarp();
    } else {
        const uint8_t* src = F.blur + (size_t)f * P.blurSlab + G.blurOff + (size_t)(unsigned)((cy - 18) * bp + (px0 - pal));
        const unsigned sbase = (unsigned)__cvta_generic_to_shared(&s_patch[threadIdx.x >> 5][0]);
#pragma unroll
        for (int k = 0; k < 13; k++) {
            const int i = lane + 32 * k;
            if (k < 12 || i < DESC_PATCH_ROWS * 11) {
                const int row = (i * 373) >> 12, word = i - 11 * row;          // i / 11 for i < 416
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sbase + 4u * (unsigned)(row * DESC_PATCH_WPITCH + word)),
                             "l"(src + (size_t)(unsigned)(row * bp + 4 * word)));
            }
        }
    }

    // IC_Angle: first-order moments over the radius-15 disc of the UNBLURRED level.  The 31 x 31 patch is read as aligned
    // 32-bit words (9 per row whatever the alignment of cx - 15): 279 (row, word) items spread over the 32 lanes, and every
    // item is two IDP.4A against host-built weight words (u resp. v inside the disc, 0 outside) selected by the alignment.
    int m10 = 0, m01 = 0;
    if constexpr (TMA) {
        // the box starts at the 16-byte boundary at or below cx - 15: the patch is words (al >> 2) .. (al >> 2) + 8 of each 12-word
        // row, weighted by the table of its byte alignment al & 3 -- the same 279 (row, word) items as the LDG form, as LDS.32
        mbar_wait(bar0, 0);
        const int al = (cx - HALF_PATCH) & 15;
        const uint32_t* box = reinterpret_cast<const uint32_t*>(s_box + (threadIdx.x >> 5) * DESC_BOX_STRIDE + DESC_BOX_IMG_OFF) + (al >> 2);
        const uint2* tab = icTab + (al & 3) * IC_ITEMS_PAD;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            const int i = lane + 32 * k;
            if (k < 8 || i < IC_ITEMS) {
                const int row = (i * 57) >> 9, word = i - 9 * row;         // i / 9 for i < 288
                const uint2 wt = __ldg(&tab[i]);
                const unsigned px = box[row * (DESC_BOX_IMG_W / 4) + word];
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(px), "r"(wt.x));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m01) : "r"(px), "r"(wt.y));
            }
        }
    } else {
        int pitch;
        const uint8_t* img = level_plane(P, F, f, l, pitch);
        const int x0 = cx - HALF_PATCH, al = x0 & 3;
        const uint8_t* org = img + (size_t)(unsigned)((cy - HALF_PATCH) * pitch + (x0 - al));
        const uint2* tab = icTab + al * IC_ITEMS_PAD;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            const int i = lane + 32 * k;
            if (k < 8 || i < IC_ITEMS) {
                const int row = (i * 57) >> 9, word = i - 9 * row;         // i / 9 for i < 288
                const uint2 wt = __ldg(&tab[i]);
                const unsigned px = __ldg(reinterpret_cast<const unsigned*>(org + (unsigned)(row * pitch + 4 * word)));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(px), "r"(wt.x));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m01) : "r"(px), "r"(wt.y));
            }
        }
    }
    m10 = __reduce_add_sync(0xFFFFFFFFu, m10);             // REDUX.SUM: one instruction instead of five shuffle + add steps
    m01 = __reduce_add_sync(0xFFFFFFFFu, m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // rBRIEF on the blurred level: lane = descriptor byte, 8 test pairs each
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    orbx_sc::sincosf_glibc(__fmul_rn(angle, factorPI), &b, &a);       // a = cos, b = sin (:112-113)
    // shared-space byte address of the patch centre with both rounding biases folded in (32-bit arithmetic, wraps):
    // address = centre + (rbits - MAGIC) * pitch + (cbits - MAGIC)
    constexpr unsigned PB = TMA ? DESC_BOX_BLUR_W : DESC_PATCH_WPITCH * 4;
    const unsigned ctrb = (TMA ? boxBase : (unsigned)__cvta_generic_to_shared(&s_patch[threadIdx.x >> 5][0])) + 18u * PB + 18u + (unsigned)pal -
                          (unsigned)MAGIC_BITS * (PB + 1u);
    if constexpr (TMA) mbar_wait(bar0 + 8, 0);
    else {
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
    }
    unsigned val = 0;
#pragma unroll
    for (int bit = 0; bit < 8; bit++) {
        const float4 pt = __ldg(&patT[bit * 32 + lane]);
        int r0, c0, r1, c1;
        if (P.fmaMode) {     // contracted form g++ -O3 -march=native emits for x*b + y*a and x*a - y*b
            r0 = round_magic_bits(__fmaf_rn(pt.x, b, __fmul_rn(pt.y, a)));
            c0 = round_magic_bits(__fmaf_rn(pt.x, a, -__fmul_rn(pt.y, b)));
            r1 = round_magic_bits(__fmaf_rn(pt.z, b, __fmul_rn(pt.w, a)));
            c1 = round_magic_bits(__fmaf_rn(pt.z, a, -__fmul_rn(pt.w, b)));
        } else {
            r0 = round_magic_bits(__fadd_rn(__fmul_rn(pt.x, b), __fmul_rn(pt.y, a)));
            c0 = round_magic_bits(__fsub_rn(__fmul_rn(pt.x, a), __fmul_rn(pt.y, b)));
            r1 = round_magic_bits(__fadd_rn(__fmul_rn(pt.z, b), __fmul_rn(pt.w, a)));
            c1 = round_magic_bits(__fsub_rn(__fmul_rn(pt.z, a), __fmul_rn(pt.w, b)));
        }
        unsigned t0, t1;
        asm("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(ctrb + (unsigned)r0 * PB + (unsigned)c0));
        asm("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(ctrb + (unsigned)r1 * PB + (unsigned)c1));
        val |= (unsigned)(t0 < t1) << bit;
    }
    const size_t o = (size_t)f * P.totalNodes + g;
    F.desc[o * 32 + lane] = (uint8_t)val;
    // keypoint record: 7 words written by 7 lanes
    if (lane < 7) {
        const float kx = l ? __fmul_rn((float)cx, G.scale) : (float)cx;      // pt *= mvScaleFactor[level] for level != 0 (:1095-1101)
        const float ky = l ? __fmul_rn((float)cy, G.scale) : (float)cy;
        unsigned wv;
        switch (lane) {
            case 0: wv = __float_as_uint(kx); break;
            case 1: wv = __float_as_uint(ky); break;
            case 2: wv = __float_as_uint((float)G.scaledPatch); break;
            case 3: wv = __float_as_uint(angle); break;
            case 4: wv = __float_as_uint((float)key_r(key)); break;
            case 5: wv = (unsigned)l; break;
            default: wv = 0xFFFFFFFFu; break;                                  // class_id = -1
        }
        reinterpret_cast<unsigned*>(F.kps + o)[lane] = wv;
    }
}

// The same operator with KPW (16) keypoints per warp (TMA form only).  Everything that is one value per keypoint -- the level
// look-up, fastAtan2, the double-precision sincosf restatement (85 of the 614 warp instructions per keypoint of the kernel
// above), the 28-byte record -- runs once per WARP with lane k working for keypoint k; the two warp-cooperative parts, the
// moments (lane = patch words) and the descriptor (lane = descriptor byte), loop over the KPW keypoints with the next
// keypoint's box already in flight (two buffers per warp, two mbarriers per loop) and the 8 float4 pattern entries of the
// lane held in registers for all of them.
#ifndef OD32_WARPS_N
#define OD32_WARPS_N 8
#endif
#ifndef OD32_MINB
#define OD32_MINB 3
#endif
constexpr int OD32_WARPS = OD32_WARPS_N;
constexpr int OD32_BUF = 2432;            // one buffer takes either box (blurred 64 x 37 = 2368 B, unblurred 48 x 31 = 1488 B)

template <int KPW>                        // keypoints per warp (lanes >= KPW idle in the per-keypoint part): 16 by default, see launch_orient_describe
__global__ void __launch_bounds__(OD32_WARPS * 32, OD32_MINB) k_orient_describe32(const Plan P, const FrameBufs F, const uint2* __restrict__ icTab,
                                                                         const float4* __restrict__ patT,
                                                                         const __grid_constant__ OrientMaps M)
{
    constexpr unsigned FULL = 0xFFFFFFFFu;
    const int f = blockIdx.y + F.frame0, lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int g0 = (blockIdx.x * OD32_WARPS + wrp) * KPW;              // first keypoint of this warp inside the frame
    // levels are concatenated in ascending order (:1103): lane i holds the count of level i
    const int myc = lane < P.nlevels ? F.levelCount[(size_t)f * MAX_LEVELS + lane] : 0;
    int inc = myc;
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
        const int t = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += t;
    }
    const int total = __shfl_sync(FULL, inc, 15);
    if (g0 == 0 && lane == 0) F.counts[f] = total;
    if (g0 >= total) return;
    const int cnt = min(KPW, total - g0), g = g0 + lane;
    const bool valid = lane < cnt;

    // ---- lane k: level, key and centre of keypoint g0 + k
    int l = 0, idx = 0, nodeBase = 0, scaledPatch = 0;
    float scale = 1.f;
#pragma unroll
    for (int j = 0; j < MAX_LEVELS; j++) {
        if (j < P.nlevels) {
            const int hi = __shfl_sync(FULL, inc, j), lo = hi - __shfl_sync(FULL, myc, j);
            if (g >= lo && g < hi) { l = j; idx = g - lo; nodeBase = P.lv[j].nodeBase; scaledPatch = P.lv[j].scaledPatch; scale = P.lv[j].scale; }
        }
    }
    const uint32_t key = valid ? F.levelKeys[(size_t)f * P.totalNodes + nodeBase + idx] : 0u;
    const int cxl = key_x(key) + MIN_BORDER, cyl = key_y(key) + MIN_BORDER;      // level coordinates (:841-842)
    const unsigned packed = (unsigned)cxl | ((unsigned)cyl << 13) | ((unsigned)l << 26);

    __shared__ __align__(128) uint8_t s_buf[OD32_WARPS * 2 * OD32_BUF];
    __shared__ uint64_t s_bar[OD32_WARPS][4];
    const unsigned bufBase = (unsigned)__cvta_generic_to_shared(s_buf) + (unsigned)(wrp * 2 * OD32_BUF);
    const unsigned barA = (unsigned)__cvta_generic_to_shared(&s_bar[wrp][0]), barC = barA + 16;
    if (lane == 0) {
        mbar_init(barA, 1);
        mbar_init(barA + 8, 1);
        mbar_init(barC, 1);
        mbar_init(barC + 8, 1);
        mbar_fence_init();
    }
    __syncwarp();

    // ---- IC_Angle moments, one keypoint after the other: the unblurred 31 x 31 patch arrives as a 48 x 31 box from the
    // 16-byte boundary at or below cx - 15 (279 (row, word) items over the 32 lanes, two IDP.4A each, weights by alignment)
    if (lane == 0) {
        mbar_expect_tx(barA, DESC_BOX_IMG_BYTES);
        tma_load_box(bufBase, &M.img.m[l], (cxl - HALF_PATCH) & ~15, cyl - HALF_PATCH, f, barA);
    }
    int my10 = 0, my01 = 0;
    for (int k = 0; k < cnt; k++) {
        const unsigned pk = __shfl_sync(FULL, packed, k), pn = __shfl_sync(FULL, packed, (k + 1) & 31);
        if (lane == 0 && k + 1 < cnt) {          // the buffer of keypoint k - 1: every lane passed that iteration's REDUX
            const unsigned bn = barA + 8u * ((k + 1) & 1);
            mbar_expect_tx(bn, DESC_BOX_IMG_BYTES);
            tma_load_box(bufBase + (unsigned)(((k + 1) & 1) * OD32_BUF), &M.img.m[pn >> 26], (int)((pn & 0x1FFFu) - HALF_PATCH) & ~15,
                         (int)((pn >> 13) & 0x1FFFu) - HALF_PATCH, f, bn);
        }
        mbar_wait(barA + 8u * (k & 1), (k >> 1) & 1);
        const int al = ((int)(pk & 0x1FFFu) - HALF_PATCH) & 15;
        const uint32_t* box = reinterpret_cast<const uint32_t*>(s_buf + (wrp * 2 + (k & 1)) * OD32_BUF) + (al >> 2);
        const uint2* tab = icTab + (al & 3) * IC_ITEMS_PAD;
        int m10 = 0, m01 = 0;
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const int i = lane + 32 * it;
            if (it < 8 || i < IC_ITEMS) {
                const int row = (i * 57) >> 9, word = i - 9 * row;         // i / 9 for i < 288
                const uint2 wt = __ldg(&tab[i]);
                const unsigned px = box[row * (DESC_BOX_IMG_W / 4) + word];
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(px), "r"(wt.x));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m01) : "r"(px), "r"(wt.y));
            }
        }
        m10 = __reduce_add_sync(FULL, m10);
        m01 = __reduce_add_sync(FULL, m01);
        if (lane == k) { my10 = m10; my01 = m01; }
    }

    // ---- lane k: orientation, rotation and the record of keypoint k
    const float angle = fast_atan2_deg((float)my01, (float)my10);
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float ca, sb;
    orbx_sc::sincosf_glibc(__fmul_rn(angle, factorPI), &sb, &ca);      // ca = cos, sb = sin (:112-113)
    if (valid) {
        unsigned* rec = reinterpret_cast<unsigned*>(F.kps + ((size_t)f * P.totalNodes + g));
        rec[0] = __float_as_uint(l ? __fmul_rn((float)cxl, scale) : (float)cxl);      // pt *= mvScaleFactor[level] for level != 0 (:1095-1101)
        rec[1] = __float_as_uint(l ? __fmul_rn((float)cyl, scale) : (float)cyl);
        rec[2] = __float_as_uint((float)scaledPatch);
        rec[3] = __float_as_uint(angle);
        rec[4] = __float_as_uint((float)key_r(key));
        rec[5] = (unsigned)l;
        rec[6] = 0xFFFFFFFFu;                                                          // class_id = -1
    }

    // ---- rBRIEF on the blurred level, one keypoint after the other: lane = descriptor byte, its 8 test pairs in registers
    float4 pt[8];
#pragma unroll
    for (int bit = 0; bit < 8; bit++) pt[bit] = __ldg(&patT[bit * 32 + lane]);
    if (lane == 0) {                             // every lane passed the last REDUX above: both buffers are free
        mbar_expect_tx(barC, DESC_BOX_BLUR_BYTES);
        tma_load_box(bufBase, &M.blur.m[l], (cxl - 18) & ~15, cyl - 18, f, barC);
    }
    const bool fma = P.fmaMode != 0;
    uint8_t* dout = F.desc + ((size_t)f * P.totalNodes + g0) * 32 + lane;
    for (int k = 0; k < cnt; k++) {
        const unsigned pk = __shfl_sync(FULL, packed, k), pn = __shfl_sync(FULL, packed, (k + 1) & 31);
        const float a = __shfl_sync(FULL, ca, k), b = __shfl_sync(FULL, sb, k);
        if (lane == 0 && k + 1 < cnt) {          // the buffer of keypoint k - 1: freed by the __syncwarp that ended its iteration
            const unsigned bn = barC + 8u * ((k + 1) & 1);
            mbar_expect_tx(bn, DESC_BOX_BLUR_BYTES);
            tma_load_box(bufBase + (unsigned)(((k + 1) & 1) * OD32_BUF), &M.blur.m[pn >> 26], (int)((pn & 0x1FFFu) - 18) & ~15,
                         (int)((pn >> 13) & 0x1FFFu) - 18, f, bn);
        }
        // shared-space byte address of the patch centre with both rounding biases folded in (32-bit arithmetic, wraps):
        // address = centre + (rbits - MAGIC) * pitch + (cbits - MAGIC)
        const unsigned pal = ((pk & 0x1FFFu) - 18u) & 15u;
        const unsigned ctrb = bufBase + (unsigned)((k & 1) * OD32_BUF) + 18u * DESC_BOX_BLUR_W + 18u + pal - (unsigned)MAGIC_BITS * (DESC_BOX_BLUR_W + 1u);
        mbar_wait(barC + 8u * (k & 1), (k >> 1) & 1);
        unsigned val = 0;
#pragma unroll
        for (int bit = 0; bit < 8; bit++) {
            int r0, c0, r1, c1;
            if (fma) {     // contracted form g++ -O3 -march=native emits for x*b + y*a and x*a - y*b
                r0 = round_magic_bits(__fmaf_rn(pt[bit].x, b, __fmul_rn(pt[bit].y, a)));
                c0 = round_magic_bits(__fmaf_rn(pt[bit].x, a, -__fmul_rn(pt[bit].y, b)));
                r1 = round_magic_bits(__fmaf_rn(pt[bit].z, b, __fmul_rn(pt[bit].w, a)));
                c1 = round_magic_bits(__fmaf_rn(pt[bit].z, a, -__fmul_rn(pt[bit].w, b)));
            } else {
                r0 = round_magic_bits(__fadd_rn(__fmul_rn(pt[bit].x, b), __fmul_rn(pt[bit].y, a)));
                c0 = round_magic_bits(__fsub_rn(__fmul_rn(pt[bit].x, a), __fmul_rn(pt[bit].y, b)));
                r1 = round_magic_bits(__fadd_rn(__fmul_rn(pt[bit].z, b), __fmul_rn(pt[bit].w, a)));
                c1 = round_magic_bits(__fsub_rn(__fmul_rn(pt[bit].z, a), __fmul_rn(pt[bit].w, b)));
            }
            unsigned t0, t1;
            asm("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(ctrb + (unsigned)r0 * DESC_BOX_BLUR_W + (unsigned)c0));
            asm("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(ctrb + (unsigned)r1 * DESC_BOX_BLUR_W + (unsigned)c1));
            val |= (unsigned)(t0 < t1) << bit;
        }
        dout[(size_t)k * 32] = (uint8_t)val;
        __syncwarp();
    }
}

void launch_orient_describe(cudaStream_t s, const Plan& P, const FrameBufs& F, const uint2* icTab, const float4* patT, int B,
                            const LevelMaps* imgMaps, const LevelMaps* blurMaps)
{
    dim3 grd((P.totalNodes + DESC_WARPS - 1) / DESC_WARPS, B);
    OrientMaps M;
    if (imgMaps && blurMaps) {
        static const bool perWarp32 = !(getenv("ORBX_ORIENT32") && atoi(getenv("ORBX_ORIENT32")) == 0);      // A/B switch
        M.img = *imgMaps;
        M.blur = *blurMaps;
        // Measured (orientation + descriptors, ms): 16 keypoints per warp is the best or tied from ~40 VGA frames per launch up
        // (48 frames 0.042, 256 frames 0.165, 64 1080p frames 0.182 -- against 0.047 / 0.204 / 0.246 for one keypoint per warp and
        // 0.052 / 0.180 / 0.234 for 32); below that a warp's serial run of keypoints is the critical path of the launch and the
        // one-keypoint-per-warp form wins (one frame per call: 0.009 against 0.040 ms with 32 keypoints per warp)
        static const int kpwEnv = getenv("ORBX_ORIENT_KPW") ? atoi(getenv("ORBX_ORIENT_KPW")) : 0;      // tuning override: 1 / 8 / 16 / 32
        const int kpw = kpwEnv ? kpwEnv : ((long long)B * P.totalNodes >= 36000 ? 16 : 1);
        if (perWarp32 && kpw > 1) {
            dim3 g32((P.totalNodes + OD32_WARPS * kpw - 1) / (OD32_WARPS * kpw), B);
            if (kpw == 32) k_orient_describe32<32><<<g32, OD32_WARPS * 32, 0, s>>>(P, F, icTab, patT, M);
            else if (kpw == 16) k_orient_describe32<16><<<g32, OD32_WARPS * 32, 0, s>>>(P, F, icTab, patT, M);
            else k_orient_describe32<8><<<g32, OD32_WARPS * 32, 0, s>>>(P, F, icTab, patT, M);
        } else
            k_orient_describe<true><<<grd, DESC_WARPS * 32, 0, s>>>(P, F, icTab, patT, M);
    } else {
        memset(&M, 0, sizeof M);
        k_orient_describe<false><<<grd, DESC_WARPS * 32, 0, s>>>(P, F, icTab, patT, M);
    }
}

// ------------------------------------------------------------------------------------------------ padded copy-out
__global__ void k_reflect_copy(const uint8_t* __restrict__ src, int w, int h, int spitch, uint8_t* __restrict__ dst, int border, int dpitch)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w + 2 * border || y >= h + 2 * border) return;
    dst[(size_t)y * dpitch + x] = src[(size_t)reflect101(y - border, h) * spitch + reflect101(x - border, w)];
}

void launch_reflect_copy(cudaStream_t s, const uint8_t* src, int w, int h, int spitch, uint8_t* dst, int border, int dpitch)
{
    dim3 blk(32, 8), grd((w + 2 * border + 31) / 32, (h + 2 * border + 7) / 8);
    k_reflect_copy<<<grd, blk, 0, s>>>(src, w, h, spitch, dst, border, dpitch);
}

// ------------------------------------------------------------------------------------------------ grey conversion
// cvtColor(im, im, CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of Tracking::GrabImage* (src/Tracking.cc:280-305, :320-334,
// :350-364) written straight into the level-0 planes.  OpenCV's 8-bit path: (R*9798 + G*19235 + B*3735 + (1 << 14)) >> 15.
// A thread makes 4 output pixels (one word) from 3 (CN = 3) or 4 (CN = 4) input words.
__device__ __forceinline__ uint32_t gray1(uint32_t px, int c0, int c1, int c2)
{
    return ((px & 0xFF) * c0 + ((px >> 8) & 0xFF) * c1 + ((px >> 16) & 0xFF) * c2 + (1 << 14)) >> 15;
}

template <int CN>
__global__ void __launch_bounds__(256) k_gray(const uint8_t* __restrict__ src, size_t spitch, size_t sframe, uint8_t* __restrict__ dst,
                                               int w, int h, size_t dpitch, size_t dframe, int c0, int c1, int c2)
{
    const int w4 = (w + 3) >> 2;
    const uint8_t* sf = src + blockIdx.y * sframe;
    uint8_t* df = dst + blockIdx.y * dframe;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w4 * h; i += gridDim.x * blockDim.x) {
        const int y = i / w4, x4 = i - y * w4;
        const uint8_t* row = sf + (size_t)y * spitch + (size_t)x4 * 4 * CN;
        uint32_t p0, p1, p2, p3;
        if (4 * x4 + 4 <= w) {
            const uint32_t* r32 = reinterpret_cast<const uint32_t*>(row);
            if (CN == 4) {
                const uint4 v = *reinterpret_cast<const uint4*>(row);
                p0 = v.x; p1 = v.y; p2 = v.z; p3 = v.w;
            } else {
                const uint32_t a = r32[0], b = r32[1], c = r32[2];
                p0 = a;
                p1 = __byte_perm(a, b, 0x6543);
                p2 = __byte_perm(b, c, 0x5432);
                p3 = c >> 8;
            }
        } else {                                  // ragged last word of a row: byte loads, never past w * CN
            uint32_t px[4] = {0, 0, 0, 0};
            for (int k = 0; k < w - 4 * x4; k++) px[k] = row[CN * k] | (row[CN * k + 1] << 8) | (row[CN * k + 2] << 16);
            p0 = px[0]; p1 = px[1]; p2 = px[2]; p3 = px[3];
        }
        const uint32_t g = gray1(p0, c0, c1, c2) | (gray1(p1, c0, c1, c2) << 8) | (gray1(p2, c0, c1, c2) << 16) | (gray1(p3, c0, c1, c2) << 24);
        *reinterpret_cast<uint32_t*>(df + (size_t)y * dpitch + 4 * x4) = g;
    }
}

void launch_gray(cudaStream_t s, const uint8_t* src, size_t spitch, size_t sframe, uint8_t* dst, int w, int h, size_t dpitch, size_t dframe,
                 int channels, int rgb, int B)
{
    const int c0 = rgb ? 9798 : 3735, c1 = 19235, c2 = rgb ? 3735 : 9798;      // RY15, GY15, BY15 in memory order
    const int w4 = (w + 3) >> 2;
    dim3 grd(std::min((w4 * h + 255) / 256, 148 * 8), B);
    if (channels == 4) k_gray<4><<<grd, 256, 0, s>>>(src, spitch, sframe, dst, w, h, dpitch, dframe, c0, c1, c2);
    else k_gray<3><<<grd, 256, 0, s>>>(src, spitch, sframe, dst, w, h, dpitch, dframe, c0, c1, c2);
}

}  // namespace orbx
