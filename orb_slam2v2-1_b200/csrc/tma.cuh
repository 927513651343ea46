// tma.cuh -- Tensor Memory Accelerator helpers for the extraction kernels (sm_100a).
//
// Every pyramid level (unblurred and blurred) is described once per handle by a 3-D tiled tensor map (x, y, frame) over
// its u8 planes; a kernel then fetches a whole box -- a FAST cell tile, the 31 x 31 orientation patch, the 37 x 37
// descriptor footprint -- with ONE cp.async.bulk.tensor instruction issued by one lane, instead of hundreds of per-lane
// LDGSTS with hand-computed addresses.  The plane base, the row pitch, the frame stride AND the x coordinate of a box must be
// multiples of 16 bytes (an unaligned x raises "illegal instruction"; tools/tma_probe.cu), y and the frame are free, and a box
// may leave the plane (zero fill): callers round x down to 16 and ask for a box wide enough to hold the slack.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "orbx_common.cuh"

namespace orbx {

// One map per pyramid level, passed to kernels as a __grid_constant__ parameter.  8 levels (the reference's default, config/Asus.yaml)
// keep a kernel's parameter block -- Plan + FrameBufs + two of these -- under 4 KB; extractors with more levels use the LDGSTS forms.
constexpr int TMA_LEVELS = 8;
struct LevelMaps { CUtensorMap m[TMA_LEVELS]; };

// Encodes a map over `frames` planes of w x h bytes (row pitch and frame stride in bytes) with a box of boxW x boxH x 1.
// Host only; resolves cuTensorMapEncodeTiled through the runtime (no link against libcuda).  Returns false when the driver
// entry point is missing or the geometry is not encodable (base / pitch / frame stride not multiples of 16, box too large).
bool tma_encode_planes(CUtensorMap* map, const void* base, int w, int h, int frames, size_t pitch, size_t frameStride, int boxW, int boxH);
inline bool tma_aligned(const void* base, size_t pitch, size_t frameStride) { return (((uintptr_t)base | pitch | frameStride) & 15) == 0; }

#ifdef __CUDACC__
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
// makes the initialised barrier visible to the async proxy (the TMA unit) before a copy may complete on it
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// box (x, y, frame) of the map -> shared memory at dst (128-byte aligned), completion counted in bytes on bar
__device__ __forceinline__ void tma_load_box(unsigned dst, const CUtensorMap* map, int x, int y, int frame, unsigned bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(frame), "r"(bar)
                 : "memory");
}
#endif

}  // namespace orbx
