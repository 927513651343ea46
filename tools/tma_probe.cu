// tma_probe.cu -- stand-alone check of the TMA wrappers of csrc/tma.cuh (development tool): loads a box of a u8 plane stack into shared
// memory in a few variants and compares with the host.  usage: tma_probe <variant>   build: nvcc -gencode arch=compute_100a,code=sm_100a
//   0: one map as __grid_constant__ parameter   1: map inside a struct array, dynamic index   2: two boxes, two barriers, 8 warps
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../orb_slam2v2-1_b200/csrc/tma.cuh"
using namespace orbx;

namespace orbx {
void set_error(const char*, ...) {}
bool cuda_ok(cudaError_t e, const char*, const char*, int) { return e == cudaSuccess; }
bool tma_encode_planes(CUtensorMap* map, const void* base, int w, int h, int frames, size_t pitch, size_t frameStride, int boxW, int boxH)
{
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frameStride};
    const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1u}, estr[3] = {1u, 1u, 1u};
    CUresult r = ((EncodeFn)fn)(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) printf("encode failed: %d\n", (int)r);
    return r == CUDA_SUCCESS;
}
}

constexpr int BW = 32, BH = 31;
__global__ void k_one(const __grid_constant__ CUtensorMap map, int x, int y, int f, uint8_t* out)
{
    __shared__ __align__(128) uint8_t box[BW * BH];
    __shared__ uint64_t bar;
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar), d = (unsigned)__cvta_generic_to_shared(box);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        mbar_fence_init();
        mbar_expect_tx(b, BW * BH);
        tma_load_box(d, &map, x, y, f, b);
    }
    __syncwarp();
    mbar_wait(b, 0);
    for (int i = threadIdx.x; i < BW * BH; i += 32) out[i] = box[i];
}
// variant 3: plain bulk copy (no tensor map): mbarrier machinery alone.  variant 4: 2-D map.
__global__ void k_bulk(const uint8_t* src, uint8_t* out)
{
    __shared__ __align__(128) uint8_t box[1024];
    __shared__ uint64_t bar;
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar), d = (unsigned)__cvta_generic_to_shared(box);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        mbar_fence_init();
        mbar_expect_tx(b, 1024);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d), "l"(src), "r"(1024), "r"(b) : "memory");
    }
    __syncwarp();
    mbar_wait(b, 0);
    for (int i = threadIdx.x; i < 1024; i += 32) out[i] = box[i];
}
__global__ void k_2d(const __grid_constant__ CUtensorMap map, int x, int y, uint8_t* out, int bw, const CUtensorMap* gmap)
{
    __shared__ __align__(128) uint8_t box[256 * BH];
    __shared__ uint64_t bar;
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar), d = (unsigned)__cvta_generic_to_shared(box);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        mbar_fence_init();
        mbar_expect_tx(b, bw * BH);
        if (gmap) asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(reinterpret_cast<uint64_t>(gmap)) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(d),
                     "l"(gmap ? reinterpret_cast<uint64_t>(gmap) : reinterpret_cast<uint64_t>(&map)), "r"(x), "r"(y), "r"(b) : "memory");
    }
    __syncwarp();
    mbar_wait(b, 0);
    for (int i = threadIdx.x; i < bw * BH; i += 32) out[i] = box[i];
}
struct Maps { LevelMaps a, b; };
__global__ void k_arr(int pad0, int pad1, const __grid_constant__ Maps M, int l, int x, int y, int f, uint8_t* out)
{
    __shared__ __align__(128) uint8_t box[8][2 * 1024];
    __shared__ uint64_t bar[8][2];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar[w][0]), d = (unsigned)__cvta_generic_to_shared(&box[w][0]);
    if (lane == 0) {
        mbar_init(b, 1);
        mbar_init(b + 8, 1);
        mbar_fence_init();
        mbar_expect_tx(b, BW * BH);
        tma_load_box(d, &M.a.m[l], x + w, y, f, b);
        mbar_expect_tx(b + 8, BW * BH);
        tma_load_box(d + 1024, &M.b.m[l], x, y + w, f, b + 8);
    }
    __syncwarp();
    mbar_wait(b, 0);
    mbar_wait(b + 8, 0);
    for (int i = lane; i < BW * BH; i += 32) { out[(2 * w) * 1024 + i] = box[w][i]; out[(2 * w + 1) * 1024 + i] = box[w][1024 + i]; }
}

int main(int argc, char** argv)
{
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    const int W = 640, H = 480, F = 4, pitch = 640;
    std::vector<uint8_t> h((size_t)pitch * H * F);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *out;
    cudaMalloc(&d, h.size());
    cudaMalloc(&out, 16 * 1024);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    Maps M;
    memset(&M, 0, sizeof M);
    for (int l = 0; l < 3; l++)
        if (!tma_encode_planes(&M.a.m[l], d, W, H, F, pitch, (size_t)pitch * H, BW, BH) || !tma_encode_planes(&M.b.m[l], d, W, H, F, pitch, (size_t)pitch * H, BW, BH)) return 2;
    const int x = argc > 5 ? atoi(argv[5]) : 101, y = 37, f = 2, l = 1;
    std::vector<uint8_t> got(16 * 1024);
    int bad = 0;
    if (variant == 3) {
        k_bulk<<<1, 32>>>(d + 4096, out);
        cudaError_t e3 = cudaDeviceSynchronize();
        cudaMemcpy(got.data(), out, 1024, cudaMemcpyDeviceToHost);
        int b3 = 0;
        for (int i = 0; i < 1024; i++) b3 += got[i] != h[4096 + i];
        printf("variant 3 (bulk, no map): %s, %d mismatches\n", cudaGetErrorString(e3), b3);
        return 0;
    }
    if (variant == 4) {
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult st;
        cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &st);
        CUtensorMap m2;
        const cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)(H * F)}, strides[1] = {(cuuint64_t)pitch};
        const int bw = argc > 2 ? atoi(argv[2]) : BW;
        const int where = argc > 3 ? atoi(argv[3]) : 0;
        const int l2 = argc > 4 ? atoi(argv[4]) : 0;
        const cuuint32_t box[2] = {(cuuint32_t)bw, BH}, estr[2] = {1, 1};
        CUresult r = ((EncodeFn)fn)(&m2, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        const unsigned long long* q = (const unsigned long long*)&m2;
        printf("encode 2d rc=%d desc:", (int)r);
        for (int i = 0; i < 16; i++) printf(" %016llx", q[i]);
        printf("\n");
        CUtensorMap* gm = nullptr;
        if (where == 1) { cudaMalloc(&gm, 128); cudaMemcpy(gm, &m2, 128, cudaMemcpyHostToDevice); }
        k_2d<<<1, 32>>>(m2, x, y + f * H, out, bw, gm);
        cudaError_t e4 = cudaDeviceSynchronize();
        cudaMemcpy(got.data(), out, BW * BH, cudaMemcpyDeviceToHost);
        int b4 = 0;
        cudaMemcpy(got.data(), out, bw * BH, cudaMemcpyDeviceToHost);
        for (int r2 = 0; r2 < BH; r2++) for (int c = 0; c < bw; c++) b4 += got[r2 * bw + c] != h[(size_t)f * pitch * H + (size_t)(y + r2) * pitch + x + c];
        printf("variant 4 (2-D map, box %d, desc in %s, l2 %d): %s, %d mismatches\n", bw, where ? "global" : "param", l2, cudaGetErrorString(e4), b4);
        return 0;
    }
    if (variant == 0) {
        k_one<<<1, 32>>>(M.a.m[0], x, y, f, out);
    } else {
        k_arr<<<1, variant == 1 ? 32 : 256>>>(0, 0, M, l, x, y, f, out);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d: %s\n", variant, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(got.data(), out, got.size(), cudaMemcpyDeviceToHost);
    const int nw = variant == 0 ? 1 : (variant == 1 ? 1 : 8);
    for (int w = 0; w < nw; w++)
        for (int k = 0; k < (variant == 0 ? 1 : 2); k++)
            for (int r = 0; r < BH; r++)
                for (int c = 0; c < BW; c++) {
                    const int xx = x + (k == 0 ? w : 0) + c, yy = y + (k == 1 ? w : 0) + r;
                    const uint8_t want = h[(size_t)f * pitch * H + (size_t)yy * pitch + xx];
                    bad += got[(variant == 0 ? 0 : (2 * w + k) * 1024) + r * BW + c] != want;
                }
    printf("variant %d: %d mismatches\n", variant, bad);
    return bad != 0;
}
