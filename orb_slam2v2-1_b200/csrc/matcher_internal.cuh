// matcher_internal.cuh -- internal declarations shared by the matcher translation units.
#pragma once
#include "orbx_common.cuh"
#include <string.h>
#include <vector>

namespace orbx {

constexpr int KEY_SHIFT = 22;                                   // key = distance << 22 | scan position
constexpr uint32_t KEY_POS_MASK = (1u << KEY_SHIFT) - 1;
constexpr uint32_t KEY_NONE = (256u << KEY_SHIFT) | KEY_POS_MASK;   // "no candidate": distance 256, position all ones

// growable device scratch of a matcher context
struct Arena {
    uint8_t* base = nullptr;
    size_t cap = 0;
    cudaStream_t stream = nullptr;
    uint8_t* pin = nullptr;        // page-locked host mirror (same offsets), allocated on first use by mirror()
    size_t pinCap = 0;
    int reserve(size_t bytes);
    int mirror();
    void release();
};
// The host-array entry points move a dozen small arrays per call.  Issued one by one from pageable memory each costs a driver
// round trip; Xfer gathers them in the arena's pinned mirror instead and moves them with ONE copy per direction.
struct Xfer {
    Arena& A;
    size_t upEnd = 0, dnBeg = (size_t)-1, dnEnd = 0;
    struct Out { void* dst; size_t off, bytes; };
    Out outs[8];
    int nOut = 0;
    explicit Xfer(Arena& a) : A(a) {}
    struct Big { const void* src; size_t off, bytes; };
    Big bigs[4];
    int nBig = 0;
    void up(size_t off, const void* src, size_t bytes)
    {
        if (!bytes) return;
        if (bytes > (1u << 20) && nBig < 4) { bigs[nBig++] = Big{src, off, bytes}; return; }      // large arrays: no extra host copy
        memcpy(A.pin + off, src, bytes);
        if (off + bytes > upEnd) upEnd = off + bytes;
    }
    cudaError_t send()
    {
        cudaError_t e = upEnd ? cudaMemcpyAsync(A.base, A.pin, upEnd, cudaMemcpyHostToDevice, A.stream) : cudaSuccess;
        for (int i = 0; i < nBig && e == cudaSuccess; i++)      // after the gathered block, which may span their (stale) mirror range
            e = cudaMemcpyAsync(A.base + bigs[i].off, bigs[i].src, bigs[i].bytes, cudaMemcpyHostToDevice, A.stream);
        return e;
    }
    void down(void* dst, size_t off, size_t bytes)
    {
        if (!bytes) return;
        outs[nOut++] = Out{dst, off, bytes};
        if (off < dnBeg) dnBeg = off;
        if (off + bytes > dnEnd) dnEnd = off + bytes;
    }
    cudaError_t fetch()       // one device-to-host copy, waits for the stream, scatters to the caller's arrays
    {
        cudaError_t e = cudaSuccess;
        if (nOut) e = cudaMemcpyAsync(A.pin + dnBeg, A.base + dnBeg, dnEnd - dnBeg, cudaMemcpyDeviceToHost, A.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(A.stream);
        if (e == cudaSuccess)
            for (int i = 0; i < nOut; i++) memcpy(outs[i].dst, A.pin + outs[i].off, outs[i].bytes);
        return e;
    }
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
    int* d_tickets = nullptr;      // brute force: arrivals per query block (k_bf_scan), zero between calls
    int ticketCap = 0;
    int lastLaunches = 0;
    int windowSlots = 64;          // candidate slots per query of the device-frame searches: grows to what the last calls needed
};

namespace orbx {

constexpr int GRID_COLS = 64, GRID_ROWS = 48;      // FRAME_GRID_COLS / FRAME_GRID_ROWS, include/Frame.h:37-38
constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // src/ORBmatcher.cc:37-39

struct GridParams { float minX, minY, invW, invH; };

#ifdef __CUDACC__
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
#endif

// matcher_frame.cu
void build_grid_host(const orbx_kp* kps, int n, float minX, float maxX, float minY, float maxY, GridParams& gp,
                     std::vector<int>& start, std::vector<int>& entries);
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3);
int replay_initialization(const orbx_kp* k1, int n1, const orbx_kp* k2, int n2, const std::vector<int>& qmap, const std::vector<int>& off,
                          const std::vector<int>& idx, const std::vector<int>& dist, float nnratio, int checkOri, float* prev_matched,
                          int* matches12);
int replay_projection(const orbx_kp* kps, const float* u_right, const uint8_t* occupied_in, int n, const float* sf, const int* level,
                      const std::vector<float>& rr, const float* proj_xr, const uint8_t* mp_has_obs, int nmp, float nnratio,
                      const std::vector<int>& off, const std::vector<int>& idx, const std::vector<int>& dist, int* feature_match);
int replay_projection_frame(const orbx_kp* kps, const float* u_right, const uint8_t* occupied_in, int n, const std::vector<float>& qr,
                            const float* proj_ur, const float* last_angle, const uint8_t* mp_has_obs, int nmp, int checkOri,
                            const std::vector<int>& off, const std::vector<int>& idx, const std::vector<int>& dist, int* feature_match);
int windowed_distances(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, int n, float minX, float maxX, float minY, float maxY,
                       const float* qxy, const float* qr, const int* qlev2, const uint8_t* qdesc, int nq,
                       std::vector<int>& off, std::vector<int>& idx, std::vector<int>& dist);
int csr_distances(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, const std::vector<int>& off,
                  const std::vector<int>& idx, std::vector<int>& dist);

// frame_glue.cu
int undistort_keypoints_device(cudaStream_t s, orbx_kp* d_kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist);

int bruteforce_device(orbx_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int th_low, float nnratio,
                      int* d_bi, int* d_bd, int* d_sd, uint8_t* d_acc);
}
