// matcher_internal.cuh -- internal declarations shared by the matcher translation units.
#pragma once
#include "orbx_common.cuh"

namespace orbx {

constexpr int KEY_SHIFT = 22;                                   // key = distance << 22 | scan position
constexpr uint32_t KEY_POS_MASK = (1u << KEY_SHIFT) - 1;
constexpr uint32_t KEY_NONE = (256u << KEY_SHIFT) | KEY_POS_MASK;   // "no candidate": distance 256, position all ones

// growable device scratch of a matcher context
struct Arena {
    uint8_t* base = nullptr;
    size_t cap = 0;
    cudaStream_t stream = nullptr;
    int reserve(size_t bytes);
};
struct Carver {
    size_t off = 0;
    explicit Carver(const Arena&) {}
    size_t take(size_t bytes) { const size_t o = off; off += (bytes + 255) / 256 * 256; return o; }
    size_t total() const { return off; }
};

}  // namespace orbx

struct orbx_matcher {
    int device = 0;
    int smCount = 148;
    cudaStream_t stream = nullptr;
    orbx::Arena arena, arena2;     // arena2: buffers whose size is known only after a device count
    uint2* d_partial = nullptr;
    size_t partialCap = 0;
    int lastLaunches = 0;
};

namespace orbx {
int bruteforce_device(orbx_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int th_low, float nnratio,
                      int* d_bi, int* d_bd, int* d_sd, uint8_t* d_acc);
}
