// orbx_common.cuh -- shared declarations of liborbx (sm_100a).  Internal; the public surface is include/orbx.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdarg.h>
#include <stdio.h>
#include <string>
#include "../../include/orbx.h"

namespace orbx {

void set_error(const char* fmt, ...);
bool cuda_ok(cudaError_t e, const char* what, const char* file, int line);

#define ORBX_CUDA(call)                                                        \
    do {                                                                       \
        if (!::orbx::cuda_ok((call), #call, __FILE__, __LINE__)) return ORBX_ERR_CUDA; \
    } while (0)

constexpr int MAX_LEVELS = 12;
constexpr int EDGE = 19;           // EDGE_THRESHOLD, reference src/ORBextractor.cc:72
constexpr int MIN_BORDER = 16;     // EDGE_THRESHOLD - 3, :773
constexpr int HALF_PATCH = 15;     // HALF_PATCH_SIZE, :71
constexpr int MAX_DIM = 4096;      // candidate coordinates are packed in 12 bits
constexpr int MAX_NODES = 4096;    // quadtree node positions are packed in 12 bits

// packed FAST candidate: x (12 bits, relative to MIN_BORDER) | y << 12 | response << 24
__host__ __device__ inline uint32_t pack_key(int x, int y, int r) { return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)r << 24); }
__host__ __device__ inline int key_x(uint32_t k) { return (int)(k & 0xFFFu); }
__host__ __device__ inline int key_y(uint32_t k) { return (int)((k >> 12) & 0xFFFu); }
__host__ __device__ inline int key_r(uint32_t k) { return (int)(k >> 24); }

struct LevelGeom {
    int w, h;             // level size (mvImagePyramid[level] interior)
    int pitch;            // bytes between rows of the level plane (levels >= 1) and of the blurred plane
    int maxBX, maxBY;     // w - 16, h - 16 (ComputeKeyPointsOctTree :775-776)
    int nCols, nRows, wCell, hCell;
    int cellBase;         // first cell of this level in the per-frame cell array
    int slotCap;          // candidate capacity of one cell
    int quota;            // mnFeaturesPerLevel[level]
    int nIni;             // root nodes of DistributeOctTree (:543)
    int nodeCap;          // capacity of the node list / per-level keypoint output
    int nodeBase;         // first output slot of this level in the per-frame level-key array
    int blurTileBase;     // first blur warp of this level
    int blurTilesX;       // words (4 columns) per row of the level: the blur's work items are (60-row strip, word) pairs, 32 per warp
    int scaledPatch;      // (int)(31 * mvScaleFactor[level]) (:845)
    float hX;             // root node width (:545)
    float scale;          // mvScaleFactor[level]
    unsigned long long planeOff;   // byte offset of the level plane in a frame's pyramid slab (level 0: unused)
    unsigned long long blurOff;    // byte offset of the blurred plane in a frame's blur slab
    unsigned long long slotOff;    // entry offset of the level's cell slots in a frame's slot array
    unsigned long long keyOff;     // entry offset of the level's compacted candidate array in a frame's key array
};

struct Plan {
    int nlevels;
    int iniTh, minTh;
    int fmaMode;
    int totalCells;         // per frame
    int totalNodes;         // per frame: sum of nodeCap == stride of the keypoint outputs
    int totalBlurTiles;     // per frame
    int fastPitch;          // row pitch (bytes) of the FAST cell tiles in shared memory: 64 or 128 (fast_tile_pitch)
    unsigned long long pyrSlab, blurSlab, slotSlab, keySlab;   // per-frame sizes (bytes / entries)
    LevelGeom lv[MAX_LEVELS];
};

}  // namespace orbx
