// extract_kernels.cuh -- launch interface of the ORB extraction kernels (extract_kernels.cu).
#pragma once
#include "orbx_common.cuh"
#include "tma.cuh"

namespace orbx {

// Device buffers of one extractor handle for a batch of B frames (frame-major slabs).
struct FrameBufs {
    const uint8_t* lv0;            // level-0 planes (H2D staging or the caller's device images)
    unsigned long long lv0Pitch, lv0Frame;
    uint8_t* pyr;                  // B x Plan::pyrSlab      levels 1..L-1, unblurred (mvImagePyramid interiors)
    uint8_t* blur;                 // B x Plan::blurSlab     7x7 Gaussian of every level
    int* cellCount;                // B x totalCells         FAST candidates kept per cell
    uint32_t* slots;               // B x slotSlab           per-cell candidate slots (packed keys, row-major inside a cell)
    uint32_t* keys;                // B x keySlab            per-level compacted candidates in reference order
    uint16_t* knode;               // B x keySlab            node position (12 bits) | quadrant << 12 of every candidate
    uint32_t* levelKeys;           // B x totalNodes         quadtree output per level, list order
    int* levelCount;               // B x MAX_LEVELS
    orbx_kp* kps;                  // B x totalNodes
    uint8_t* desc;                 // B x totalNodes x 32
    int* counts;                   // B
    const uint4* cellTab;          // FAST_CELL_REC x totalCells (per shape, shared by all frames): FAST cell geometry, see fast_cell_record
    int frame0;                    // first frame of this launch: a large batch runs as two halves on two streams (orbx_extractor.cu)
};

struct UMaxTable { int v[16]; };
// IC_Angle weight table: [alignment 0..3][item] -> (u weights, v weights) as 4 signed bytes each; item = row * 9 + word
constexpr int IC_ITEMS = 31 * 9, IC_ITEMS_PAD = 288;

// cv::resize(INTER_LINEAR) of one level from the previous one for all frames (ComputePyramid, ORBextractor.cc:1120).
void launch_resize(cudaStream_t s, const uint8_t* src, int spitch, size_t sframe, uint8_t* dst, int dw, int dh, int dpitch,
                   size_t dframe, const void* xtab, const int4* ytab, int B, int sh);
// xtab: one 32-byte record per group of 4 output columns (see build_resize_tables)
// per-cell FAST-9/16 + NMS + ini/min retry, all levels and frames (ComputeKeyPointsOctTree cell loop, :789-829)
void launch_fast_cells(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, size_t smemBytes, const LevelMaps* maps);
// DistributeOctTree (:539-763), one CTA per (level, frame)
cudaError_t launch_octree(cudaStream_t s, const Plan& P, const FrameBufs& F, int B, int capS, int sortCap, int form = 0);   // form: orbx_set_octree_form
// GaussianBlur 7x7 sigma 2 of every level (:1085-1086)
void launch_blur(cudaStream_t s, const Plan& P, const FrameBufs& F, int B);
// IC_Angle (:77-104) + computeOrbDescriptor (:108-147) + operator()'s assembly (:1095-1103)
// imgMaps / blurMaps: TMA maps of the unblurred / blurred levels with boxes of 48 x 31 / 64 x 37 bytes (null: LDGSTS form)
void launch_orient_describe(cudaStream_t s, const Plan& P, const FrameBufs& F, const uint2* icTab, const float4* patT, int B,
                            const LevelMaps* imgMaps, const LevelMaps* blurMaps);
// padded (w+2b)x(h+2b) BORDER_REFLECT_101 copy of one plane into dst (row pitch dpitch)
void launch_reflect_copy(cudaStream_t s, const uint8_t* src, int w, int h, int spitch, uint8_t* dst, int border, int dpitch);

// cvtColor RGB(A)/BGR(A) -> GRAY (Tracking.cc:280-305) of B frames into level-0 planes; src rows 4-byte aligned (16 for 4 channels)
void launch_gray(cudaStream_t s, const uint8_t* src, size_t spitch, size_t sframe, uint8_t* dst, int w, int h, size_t dpitch, size_t dframe,
                 int channels, int rgb, int B);

size_t fast_smem_bytes(const Plan& P);
int fast_tile_pitch(const Plan& P);                                       // 64 or 128; 0 = a cell is too wide
constexpr int FAST_CELL_REC = 4;   // uint4 per cell of the FAST cell table
void fast_cell_record(const LevelGeom& G, int level, int i, int j, int tp, uint4* rec);   // the table entries of cell (i, j); tp = fast_tile_pitch
size_t octree_smem_bytes(int capS, int sortCap, int keysCap = 0);   // keysCap 0: the larger of the two kernel forms

}  // namespace orbx
