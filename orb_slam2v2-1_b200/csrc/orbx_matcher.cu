// orbx_matcher.cu -- Hamming matcher half of the C ABI (include/orbx.h): ORBmatcher::DescriptorDistance and the
// best / second-best scans of the SearchBy* routines (reference src/ORBmatcher.cc).
//
// A sequential scan with "strict < for best, else-if < for second" (ORBmatcher.cc:205-226) keeps the two smallest
// elements under the total order (distance, scan position).  Every kernel therefore carries 32-bit keys
// distance << 22 | position and a running (min1, min2): min2 = min(min2, max(min1, key)); min1 = min(min1, key) -- three
// integer min/max per pair -- and any split of the scan (lanes, segments) merges exactly.  Distances come from
// __popc on the integer pipe (8 x 32-bit per pair); no tensor cores: this is not a floating-point contraction.
#include "orbx_common.cuh"
#include "matcher_internal.cuh"
#include <string.h>
#include <algorithm>
#include <mutex>
#include <vector>

namespace orbx {

// Brute-force variant: POPC issues at 16 lanes/clk/SM against 64 for LOP3 (tools/int_pipe_bench.cu), so eight POPCs per pair
// bind the scan.  Two carry-save adders (sum = a^b^c, carry = majority, one LOP3 each) fold six of the eight xor words into
// two "ones" and two "twos" words: 6 POPCs + 4 LOP3s give the same integer, and the POPC and ALU pipes end up balanced.
__device__ __forceinline__ unsigned lop3_xor3(unsigned a, unsigned b, unsigned c)
{
    unsigned d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ unsigned lop3_maj(unsigned a, unsigned b, unsigned c)
{
    unsigned d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int hamming256_csa(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    const unsigned s1 = lop3_xor3(x0, x1, x2), c1 = lop3_maj(x0, x1, x2);
    const unsigned s2 = lop3_xor3(x3, x4, x5), c2 = lop3_maj(x3, x4, x5);
    return (__popc(s1) + __popc(s2) + __popc(x6) + __popc(x7)) + 2 * (__popc(c1) + __popc(c2));
}

// a distance of 256 never passes "dist < 256" in the reference scan: map it to the empty key
__device__ __forceinline__ uint32_t make_key(int d, uint32_t pos) { return d > 255 ? KEY_NONE : ((uint32_t)d << KEY_SHIFT) | pos; }

__device__ __forceinline__ void top2_push(uint32_t& m1, uint32_t& m2, uint32_t key)
{
    m2 = min(m2, max(m1, key));
    m1 = min(m1, key);
}
__device__ __forceinline__ void top2_merge(uint32_t& m1, uint32_t& m2, uint32_t o1, uint32_t o2)
{
    m2 = min(max(m1, o1), min(m2, o2));
    m1 = min(m1, o1);
}

// ------------------------------------------------------------------------------------------------ pairwise distance
__global__ void k_pair_distance(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int* __restrict__ dist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dist[i] = hamming256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

// ------------------------------------------------------------------------------------------------ brute force
// One thread owns one query (its 256 bits live in 8 registers) and scans a contiguous segment of the train set in index
// order; train rows are staged through shared memory in tiles and read as warp-wide broadcasts.  Segments (blockIdx.y)
// give the grid enough CTAs to fill 148 SMs.  The per-segment (min1, min2) go to global memory and the LAST CTA of a query
// block to finish (ticket counter per query block, reset for the next call) folds them and writes the results: one launch
// per match call instead of scan + merge (2000 x 2000: 17 -> 13 us).  Keys carry the scan position, so the fold is exact
// in any order.
constexpr int BF_THREADS = 128, BF_TILE = 256;

__device__ __forceinline__ void bf_finish(uint32_t m1, uint32_t m2, int qi, int th_low, float nnratio, int* __restrict__ best_idx,
                                          int* __restrict__ best_dist, int* __restrict__ second_dist, uint8_t* __restrict__ accepted)
{
    if ((m1 >> KEY_SHIFT) >= 256u) m1 = KEY_NONE;        // "dist < 256" never held (ORBmatcher.cc:217)
    if ((m2 >> KEY_SHIFT) >= 256u) m2 = KEY_NONE;
    const int b = (int)(m1 >> KEY_SHIFT), sd = (int)(m2 >> KEY_SHIFT);
    best_idx[qi] = m1 == KEY_NONE ? -1 : (int)(m1 & KEY_POS_MASK);
    best_dist[qi] = b;
    second_dist[qi] = sd;
    // acceptance of ORBmatcher.cc:228-230: bestDist1 <= TH_LOW && (float)bestDist1 < mfNNratio * (float)bestDist2
    if (accepted) accepted[qi] = (m1 != KEY_NONE) && b <= th_low && (float)b < __fmul_rn(nnratio, (float)sd);
}

__global__ void __launch_bounds__(BF_THREADS) k_bf_scan(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t, int nt,
                                                        int segLen, uint2* partial, int* tickets, int th_low, float nnratio,
                                                        int* __restrict__ best_idx, int* __restrict__ best_dist,
                                                        int* __restrict__ second_dist, uint8_t* __restrict__ accepted)
{
    __shared__ uint4 st[BF_TILE * 2];
    __shared__ int s_last;
    const int tid = threadIdx.x;
    const int qi = blockIdx.x * BF_THREADS + tid;
    const int j0 = blockIdx.y * segLen, j1 = min(j0 + segLen, nt);
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (qi < nq) { qa = __ldg(&q[2 * qi]); qb = __ldg(&q[2 * qi + 1]); }
    uint32_t m1 = KEY_NONE, m2 = KEY_NONE;
    for (int base = j0; base < j1; base += BF_TILE) {
        const int cnt = min(BF_TILE, j1 - base);
        __syncthreads();
        for (int i = tid; i < cnt * 2; i += BF_THREADS) st[i] = __ldg(&t[2 * (size_t)base + i]);
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            // distance 256 (complementary rows) yields the key 256 << 22 | pos: above every real key, folded to "none" at the end
            const int d = hamming256_csa(qa, qb, st[2 * j], st[2 * j + 1]);
            top2_push(m1, m2, ((uint32_t)d << KEY_SHIFT) + (uint32_t)(base + j));
        }
    }
    const int nseg = gridDim.y;
    if (nseg > 1) {
        if (qi < nq) partial[(size_t)blockIdx.y * nq + qi] = make_uint2(m1, m2);
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const int ticket = atomicAdd(&tickets[blockIdx.x], 1);
            s_last = ticket == nseg - 1;
            if (s_last) tickets[blockIdx.x] = 0;          // every segment of this query block has arrived: ready for the next call
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        if (qi >= nq) return;
        m1 = m2 = KEY_NONE;
        for (int s = 0; s < nseg; s++) {
            const uint2 p = __ldcg(&partial[(size_t)s * nq + qi]);
            top2_merge(m1, m2, p.x, p.y);
        }
    }
    if (qi < nq) bf_finish(m1, m2, qi, th_low, nnratio, best_idx, best_dist, second_dist, accepted);
}

// ------------------------------------------------------------------------------------------------ candidate lists
// One warp per query; lanes stride over the query's candidate list, position = index inside the list.
__global__ void __launch_bounds__(256) k_candidates(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t,
                                                    const int* __restrict__ off, const int* __restrict__ idx,
                                                    int* __restrict__ best_idx, int* __restrict__ best_dist,
                                                    int* __restrict__ second_idx, int* __restrict__ second_dist)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const uint4 qa = __ldg(&q[2 * qi]), qb = __ldg(&q[2 * qi + 1]);
    const int o0 = off[qi], cnt = off[qi + 1] - o0;
    uint32_t m1 = KEY_NONE, m2 = KEY_NONE;
    for (int c = lane; c < cnt; c += 32) {
        const int j = __ldg(&idx[o0 + c]);
        const int d = hamming256(qa, qb, __ldg(&t[2 * (size_t)j]), __ldg(&t[2 * (size_t)j + 1]));
        top2_push(m1, m2, make_key(d, (uint32_t)c));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t o1 = __shfl_xor_sync(0xFFFFFFFFu, m1, o), o2 = __shfl_xor_sync(0xFFFFFFFFu, m2, o);
        top2_merge(m1, m2, o1, o2);
    }
    if (lane == 0) {
        best_idx[qi] = m1 == KEY_NONE ? -1 : idx[o0 + (int)(m1 & KEY_POS_MASK)];
        best_dist[qi] = (int)(m1 >> KEY_SHIFT);
        second_idx[qi] = m2 == KEY_NONE ? -1 : idx[o0 + (int)(m2 & KEY_POS_MASK)];
        second_dist[qi] = (int)(m2 >> KEY_SHIFT);
    }
}

__global__ void __launch_bounds__(256) k_candidate_distances(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t,
                                                             const int* __restrict__ off, const int* __restrict__ idx,
                                                             int* __restrict__ dist)
{
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= nq) return;
    const uint4 qa = __ldg(&q[2 * qi]), qb = __ldg(&q[2 * qi + 1]);
    const int o0 = off[qi], o1 = off[qi + 1];
    for (int c = o0 + lane; c < o1; c += 32) {
        const int j = __ldg(&idx[c]);
        dist[c] = hamming256(qa, qb, __ldg(&t[2 * (size_t)j]), __ldg(&t[2 * (size_t)j + 1]));
    }
}

// ------------------------------------------------------------------------------------------------ host helpers
int Arena::reserve(size_t bytes)
{
    if (bytes <= cap) return ORBX_OK;
    ORBX_CUDA(cudaStreamSynchronize(stream));
    cudaFree(base);
    base = nullptr;
    cap = 0;
    const size_t want = std::max(bytes + bytes / 4, (size_t)1 << 20);
    ORBX_CUDA(cudaMalloc(&base, want));
    cap = want;
    return ORBX_OK;
}

int Arena::mirror()
{
    if (pinCap >= cap) return ORBX_OK;
    if (pin) cudaFreeHost(pin);
    pin = nullptr;
    pinCap = 0;
    ORBX_CUDA(cudaMallocHost(&pin, cap));
    pinCap = cap;
    return ORBX_OK;
}

void Arena::release()
{
    cudaFree(base);
    if (pin) cudaFreeHost(pin);
    base = pin = nullptr;
    cap = pinCap = 0;
}

int bruteforce_device(orbx_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int th_low, float nnratio,
                      int* d_bi, int* d_bd, int* d_sd, uint8_t* d_acc)
{
    if (nq == 0) return ORBX_OK;
    const int qblocks = (nq + BF_THREADS - 1) / BF_THREADS;
    // enough CTAs for 4 per SM (a whole number of CTAs per SM when the query blocks divide it), but segments of at least 32 train rows
    int nseg = std::max(1, (4 * m->smCount + qblocks - 1) / qblocks);
    nseg = std::min(nseg, std::max(1, nt / 32));
    nseg = std::min(nseg, 65535);
    const int segLen = nt > 0 ? (nt + nseg - 1) / nseg : 1;
    nseg = nt > 0 ? (nt + segLen - 1) / segLen : 1;
    // per-segment partial results followed by one ticket counter per query block (zero between calls)
    const size_t ticketBytes = sizeof(int) * (size_t)qblocks;
    const size_t need = sizeof(uint2) * (size_t)nseg * nq + ticketBytes;
    if (need > m->partialCap || qblocks > m->ticketCap) {
        ORBX_CUDA(cudaStreamSynchronize(m->stream));
        cudaFree(m->d_partial);
        cudaFree(m->d_tickets);
        m->d_partial = nullptr;
        m->d_tickets = nullptr;
        m->partialCap = 0;
        m->ticketCap = 0;
        const int tcap = std::max(qblocks, 1024);
        ORBX_CUDA(cudaMalloc(&m->d_partial, need));
        ORBX_CUDA(cudaMalloc(&m->d_tickets, sizeof(int) * (size_t)tcap));
        ORBX_CUDA(cudaMemsetAsync(m->d_tickets, 0, sizeof(int) * (size_t)tcap, m->stream));
        m->partialCap = need;
        m->ticketCap = tcap;
    }
    dim3 grd(qblocks, nseg);
    k_bf_scan<<<grd, BF_THREADS, 0, m->stream>>>((const uint4*)d_q, nq, (const uint4*)d_t, nt, segLen, m->d_partial, m->d_tickets, th_low,
                                                 nnratio, d_bi, d_bd, d_sd, d_acc);
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

// The reference builds an ORBmatcher on the stack at every call site (Tracking.cc:742,907,1012,1329,1502,1541,
// LocalMapping.cc:223,491, LoopClosing.cc:249,598): creating a stream and growing the device / page-locked arenas, then
// cudaFree-ing them (which synchronises the device), would cost milliseconds per Tracking step.  Contexts are therefore
// recycled: orbm_destroy parks a context (stream + arenas alive) in a per-thread free list and orbm_create hands a parked
// context of the same device back out, so a stack-constructed matcher costs two vector operations.  Parked contexts are
// really destroyed when their thread exits or when the list is full.
namespace {
constexpr int PARK_MAX = 16;

void destroy_now(orbx_matcher* m)
{
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    m->arena.release();
    m->arena2.release();
    cudaFree(m->d_partial);
    cudaFree(m->d_tickets);
    if (m->stream) cudaStreamDestroy(m->stream);
    cudaGetLastError();
    delete m;
}

struct Parked {
    std::vector<orbx_matcher*> free;
    ~Parked() { for (orbx_matcher* m : free) destroy_now(m); }
};
thread_local Parked t_parked;
}  // namespace

int orbm_create(int device, orbx_matcher** out)
{
    if (!out) { set_error("orbm_create: out is null"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (device < 0) ORBX_CUDA(cudaGetDevice(&device));          // -1: the calling thread's current device
    for (size_t i = 0; i < t_parked.free.size(); i++)
        if (t_parked.free[i]->device == device) {
            *out = t_parked.free[i];
            t_parked.free.erase(t_parked.free.begin() + (long)i);
            return ORBX_OK;
        }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device >= ndev) { set_error("orbm_create: device %d not present (%d CUDA devices)", device, ndev); return ORBX_ERR_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    orbx_matcher* m = new orbx_matcher;
    m->device = device;
    int sms = 0;
    if (!cuda_ok(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute", __FILE__, __LINE__) ||
        !cuda_ok(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking), "cudaStreamCreate", __FILE__, __LINE__)) {
        delete m;
        return ORBX_ERR_CUDA;
    }
    m->smCount = sms;
    m->arena.stream = m->stream;
    m->arena2.stream = m->stream;
    *out = m;
    return ORBX_OK;
}

void orbm_destroy(orbx_matcher* m)
{
    if (!m) return;
    if ((int)t_parked.free.size() < PARK_MAX) { t_parked.free.push_back(m); return; }    // every entry point leaves the stream idle
    destroy_now(m);
}

// really frees the calling thread's parked contexts (tests, and hosts that want the memory back)
void orbm_trim(void)
{
    for (orbx_matcher* m : t_parked.free) destroy_now(m);
    t_parked.free.clear();
}

int orbm_device(const orbx_matcher* m) { return m ? m->device : -1; }

void* orbm_stream(orbx_matcher* m) { return m ? (void*)m->stream : nullptr; }
int orbm_last_launches(const orbx_matcher* m) { return m ? m->lastLaunches : 0; }

int orbm_descriptor_distance(orbx_matcher* m, const uint8_t* a, const uint8_t* b, int n, int* dist)
{
    if (!m || n < 0 || (n > 0 && (!a || !b || !dist))) { set_error("orbm_descriptor_distance: bad argument"); return ORBX_ERR_ARG; }
    if (n == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oa = cv.take((size_t)n * 32), ob = cv.take((size_t)n * 32), od = cv.take(sizeof(int) * (size_t)n);
    int rc = m->arena.reserve(cv.total());
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    ORBX_CUDA(cudaMemcpyAsync(B + oa, a, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
    ORBX_CUDA(cudaMemcpyAsync(B + ob, b, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
    k_pair_distance<<<(n + 255) / 256, 256, 0, m->stream>>>((const uint4*)(B + oa), (const uint4*)(B + ob), n, (int*)(B + od));
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(dist, B + od, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    m->lastLaunches = 1;
    return ORBX_OK;
}

int orbm_match_bruteforce_device(orbx_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                                 int* d_best_idx, int* d_best_dist, int* d_second_dist)
{
    if (!m || nq < 0 || nt < 0 || nt >= (int)KEY_POS_MASK || (nq > 0 && (!d_q || !d_best_idx || !d_best_dist || !d_second_dist)) ||
        (nt > 0 && !d_t) || ((uintptr_t)d_q & 15) || ((uintptr_t)d_t & 15)) {
        set_error("orbm_match_bruteforce_device: bad argument (descriptor pointers must be 16-byte aligned, nt < 2^22)");
        return ORBX_ERR_ARG;
    }
    ORBX_CUDA(cudaSetDevice(m->device));
    return bruteforce_device(m, d_q, nq, d_t, nt, 0, 0.f, d_best_idx, d_best_dist, d_second_dist, nullptr);
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// orbm_match_bruteforce_device_noalloc: the same scan (k_bf_scan) for callers that may not call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).  The kernel's scratch -- per-segment (min1, min2) partials
// and one ticket counter per query block -- lives in a file-scope device array of fixed size instead of the context's grown
// buffers.  Every GPU that loads the library holds the array (1 MiB + 4 KiB) until exit.  In host code `g_noalloc_scratch`
// is only a stand-in for the symbol: its device address is taken once per device with cudaGetSymbolAddress, with that
// device current, and that pointer is what the kernel receives.
constexpr int NOALLOC_MAX_QUERIES = 16384;                        // 128 query blocks
constexpr int NOALLOC_PARTIALS = 131072;                          // uint2 entries: segments x queries of one call
constexpr int NOALLOC_TICKETS = 1024;                             // >= NOALLOC_MAX_QUERIES / BF_THREADS
constexpr size_t NOALLOC_BYTES = sizeof(uint2) * (size_t)NOALLOC_PARTIALS + sizeof(int) * (size_t)NOALLOC_TICKETS;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

namespace {
struct NoAllocDevice { unsigned char* scratch = nullptr; int smCount = 0; };
NoAllocDevice g_noalloc_dev[64];
std::mutex g_noalloc_lock;
}  // namespace

int orbm_match_bruteforce_device_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int* d_best_idx, int* d_best_dist,
                                         int* d_second_dist, void* stream)
{
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_q || !d_best_idx || !d_best_dist || !d_second_dist)) || (nt > 0 && !d_t) ||
        ((uintptr_t)d_q & 15) || ((uintptr_t)d_t & 15)) {
        set_error("orbm_match_bruteforce_device_noalloc: bad argument (descriptor pointers must be 16-byte aligned)");
        return ORBX_ERR_ARG;
    }
    if (nq > NOALLOC_MAX_QUERIES || nt >= (int)KEY_POS_MASK) {
        set_error("orbm_match_bruteforce_device_noalloc: %d queries x %d train rows is beyond the fixed scratch of this entry point "
                  "(at most %d queries and %d train rows); use orbm_match_bruteforce_device", nq, nt, NOALLOC_MAX_QUERIES, (int)KEY_POS_MASK - 1);
        return ORBX_ERR_ARG;
    }
    if (nq == 0) return ORBX_OK;
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("orbm_match_bruteforce_device_noalloc: device ordinal %d out of range", dev); return ORBX_ERR_ARG; }
    NoAllocDevice nd;
    {   // set-up, once per device, with that device current
        std::lock_guard<std::mutex> hold(g_noalloc_lock);
        if (!g_noalloc_dev[dev].scratch) {
            void* p = nullptr;
            int sms = 0;
            ORBX_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            ORBX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            if (!p || sms <= 0) { set_error("orbm_match_bruteforce_device_noalloc: no device address for the scratch array"); return ORBX_ERR_CUDA; }
            g_noalloc_dev[dev].smCount = sms;
            g_noalloc_dev[dev].scratch = (unsigned char*)p;
        }
        nd = g_noalloc_dev[dev];
    }
    cudaStream_t s = (cudaStream_t)stream;
    uint2* partial = reinterpret_cast<uint2*>(nd.scratch);
    int* tickets = reinterpret_cast<int*>(nd.scratch + sizeof(uint2) * (size_t)NOALLOC_PARTIALS);
    // the same segmentation as bruteforce_device, with the segment count also bounded by the fixed partial array
    const int qblocks = (nq + BF_THREADS - 1) / BF_THREADS;
    int nseg = std::max(1, (4 * nd.smCount + qblocks - 1) / qblocks);
    nseg = std::min(nseg, std::max(1, nt / 32));
    nseg = std::min(nseg, std::max(1, NOALLOC_PARTIALS / nq));
    const int segLen = nt > 0 ? (nt + nseg - 1) / nseg : 1;
    nseg = nt > 0 ? (nt + segLen - 1) / segLen : 1;
    // the array keeps the last call's contents: partials are written before they are read, the ticket counters must start at
    // zero (an aborted call may have left them otherwise), so they are cleared on the caller's stream ahead of the kernel
    ORBX_CUDA(cudaMemsetAsync(tickets, 0, sizeof(int) * (size_t)qblocks, s));
    dim3 grd(qblocks, nseg);
    k_bf_scan<<<grd, BF_THREADS, 0, s>>>((const uint4*)d_q, nq, (const uint4*)d_t, nt, segLen, partial, tickets, 0, 0.f, d_best_idx,
                                         d_best_dist, d_second_dist, nullptr);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

static int bruteforce_host(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int th_low, float nnratio,
                           int* best_idx, int* best_dist, int* second_dist, uint8_t* accepted)
{
    if (!m || nq < 0 || nt < 0 || nt >= (int)KEY_POS_MASK || (nq > 0 && (!q || !best_idx || !best_dist || !second_dist)) || (nt > 0 && !t)) {
        set_error("orbm_match_bruteforce: bad argument");
        return ORBX_ERR_ARG;
    }
    if (nq == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oq = cv.take((size_t)nq * 32), ot = cv.take((size_t)std::max(nt, 1) * 32);
    const size_t obi = cv.take(sizeof(int) * (size_t)nq), obd = cv.take(sizeof(int) * (size_t)nq), osd = cv.take(sizeof(int) * (size_t)nq);
    const size_t oac = cv.take((size_t)nq);
    int rc = m->arena.reserve(cv.total());
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    ORBX_CUDA(cudaMemcpyAsync(B + oq, q, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    if (nt > 0) ORBX_CUDA(cudaMemcpyAsync(B + ot, t, (size_t)nt * 32, cudaMemcpyHostToDevice, m->stream));
    rc = bruteforce_device(m, B + oq, nq, B + ot, nt, th_low, nnratio, (int*)(B + obi), (int*)(B + obd), (int*)(B + osd),
                           accepted ? B + oac : nullptr);
    if (rc) return rc;
    ORBX_CUDA(cudaMemcpyAsync(best_idx, B + obi, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaMemcpyAsync(best_dist, B + obd, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaMemcpyAsync(second_dist, B + osd, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    if (accepted) ORBX_CUDA(cudaMemcpyAsync(accepted, B + oac, (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_match_bruteforce(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt,
                          int* best_idx, int* best_dist, int* second_dist)
{
    return bruteforce_host(m, q, nq, t, nt, 0, 0.f, best_idx, best_dist, second_dist, nullptr);
}

int orbm_match_ratio(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int th_low, float nnratio,
                     int* best_idx, int* best_dist, int* second_dist, uint8_t* accepted)
{
    if (!accepted) { set_error("orbm_match_ratio: accepted is null"); return ORBX_ERR_ARG; }
    return bruteforce_host(m, q, nq, t, nt, th_low, nnratio, best_idx, best_dist, second_dist, accepted);
}

static int candidates_host(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, const int* cand_off, const int* cand_idx,
                           int* best_idx, int* best_dist, int* second_idx, int* second_dist, int* dist)
{
    if (!m || nq < 0 || nt < 0 || (nq > 0 && (!q || !cand_off))) { set_error("orbm candidates: bad argument"); return ORBX_ERR_ARG; }
    if (nq == 0) return ORBX_OK;
    const int total = cand_off[nq];
    if (total < 0 || cand_off[0] != 0 || (total > 0 && (!cand_idx || !t))) { set_error("orbm candidates: bad CSR"); return ORBX_ERR_ARG; }
    for (int i = 0; i < nq; i++)
        if (cand_off[i + 1] < cand_off[i] || cand_off[i + 1] - cand_off[i] >= (int)KEY_POS_MASK) { set_error("orbm candidates: bad CSR offsets at %d", i); return ORBX_ERR_ARG; }
    for (int c = 0; c < total; c++)
        if (cand_idx[c] < 0 || cand_idx[c] >= nt) { set_error("orbm candidates: candidate %d out of range", c); return ORBX_ERR_ARG; }
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oq = cv.take((size_t)nq * 32), ot = cv.take((size_t)std::max(nt, 1) * 32);
    const size_t oo = cv.take(sizeof(int) * ((size_t)nq + 1)), oi = cv.take(sizeof(int) * (size_t)std::max(total, 1));
    const size_t o1 = cv.take(sizeof(int) * (size_t)nq), o2 = cv.take(sizeof(int) * (size_t)nq), o3 = cv.take(sizeof(int) * (size_t)nq),
                 o4 = cv.take(sizeof(int) * (size_t)nq);
    const size_t od = cv.take(sizeof(int) * (size_t)std::max(total, 1));
    int rc = m->arena.reserve(cv.total());
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    ORBX_CUDA(cudaMemcpyAsync(B + oq, q, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    if (nt > 0) ORBX_CUDA(cudaMemcpyAsync(B + ot, t, (size_t)nt * 32, cudaMemcpyHostToDevice, m->stream));
    ORBX_CUDA(cudaMemcpyAsync(B + oo, cand_off, sizeof(int) * ((size_t)nq + 1), cudaMemcpyHostToDevice, m->stream));
    if (total > 0) ORBX_CUDA(cudaMemcpyAsync(B + oi, cand_idx, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, m->stream));
    const int blocks = (nq + 7) / 8;
    if (dist) {
        k_candidate_distances<<<blocks, 256, 0, m->stream>>>((const uint4*)(B + oq), nq, (const uint4*)(B + ot), (const int*)(B + oo),
                                                             (const int*)(B + oi), (int*)(B + od));
        ORBX_CUDA(cudaGetLastError());
        if (total > 0) ORBX_CUDA(cudaMemcpyAsync(dist, B + od, sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost, m->stream));
    } else {
        k_candidates<<<blocks, 256, 0, m->stream>>>((const uint4*)(B + oq), nq, (const uint4*)(B + ot), (const int*)(B + oo),
                                                    (const int*)(B + oi), (int*)(B + o1), (int*)(B + o2), (int*)(B + o3), (int*)(B + o4));
        ORBX_CUDA(cudaGetLastError());
        ORBX_CUDA(cudaMemcpyAsync(best_idx, B + o1, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA(cudaMemcpyAsync(best_dist, B + o2, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA(cudaMemcpyAsync(second_idx, B + o3, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA(cudaMemcpyAsync(second_dist, B + o4, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    }
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    m->lastLaunches = 1;
    return ORBX_OK;
}

int orbm_match_candidates(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, const int* cand_off, const int* cand_idx,
                          int* best_idx, int* best_dist, int* second_idx, int* second_dist)
{
    if (nq > 0 && (!best_idx || !best_dist || !second_idx || !second_dist)) { set_error("orbm_match_candidates: null output"); return ORBX_ERR_ARG; }
    return candidates_host(m, q, nq, t, nt, cand_off, cand_idx, best_idx, best_dist, second_idx, second_dist, nullptr);
}

int orbm_candidate_distances(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, const int* cand_off, const int* cand_idx, int* dist)
{
    if (nq > 0 && !dist) { set_error("orbm_candidate_distances: null output"); return ORBX_ERR_ARG; }
    return candidates_host(m, q, nq, t, nt, cand_off, cand_idx, nullptr, nullptr, nullptr, nullptr, dist);
}

}  // extern "C"
