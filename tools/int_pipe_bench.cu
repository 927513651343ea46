// int_pipe_bench.cu -- issue rates of the integer / SIMD-video / dot-product instructions the extractor and the matcher
// lean on, measured on the device (SURVEY.md section 8(d) asks for the sm_100 POPC and LOP3/IADD rates next to
// MEASURED_PEAKS.json).  Every kernel runs ILP independent dependency chains per thread, 1024 threads per CTA, one CTA
// per SM, and reports thread-operations per clock per SM from clock64() deltas (independent of the SM frequency).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -diag-suppress 177,550 -o int_pipe_bench int_pipe_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

constexpr int ILP = 8, ITER = 2048;

#define OPK(NAME, BODY)                                                                                              \
    __global__ void __launch_bounds__(1024) k_##NAME(unsigned* out, long long* cyc, unsigned seed)                   \
    {                                                                                                                \
        __shared__ unsigned sm[1024]; sm[threadIdx.x] = threadIdx.x * 2654435761u; unsigned x[ILP], z = seed * 3u + 1u;                                                 \
        for (int i = 0; i < ILP; i++) x[i] = seed + i * 7919u + threadIdx.x;                                         \
        __syncthreads();                                                                                             \
        const long long t0 = clock64();                                                                              \
        for (int it = 0; it < ITER; it++) {                                                                          \
            _Pragma("unroll") for (int i = 0; i < ILP; i++) { const unsigned b = x[(i + 1) % ILP]; BODY; }                                                \
        }                                                                                                            \
        const long long t1 = clock64();                                                                              \
        unsigned acc = 0;                                                                                            \
        for (int i = 0; i < ILP; i++) acc ^= x[i];                                                                   \
        out[blockIdx.x * blockDim.x + threadIdx.x] = acc;                                                            \
        __syncthreads();                                                                                             \
        if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;                                                             \
    }

OPK(iadd3, x[i] = x[i] + b + z)
OPK(lop3, x[i] = (x[i] & b) ^ z)
OPK(imad, x[i] = x[i] * b + z)
OPK(imad_hi, x[i] = __umulhi(x[i], b) + z)
OPK(shf, x[i] = __funnelshift_r(x[i], b, 5))
OPK(prmt, x[i] = __byte_perm(x[i], b, 0x5140))
OPK(popc, x[i] = __popc(x[i] ^ b))
OPK(popc_only, x[i] = __popc(x[i]) + b)
OPK(vabsdiff4, x[i] = __vabsdiffu4(x[i], b))
OPK(vmins2, x[i] = __vmins2(x[i], b) + 1u)
OPK(vmax_s32, x[i] = max((int)x[i], (int)b) + 1)
OPK(dp4a, x[i] = __dp4a(x[i], b, z))
OPK(dp2a, x[i] = __dp2a_lo(x[i], b, z))
OPK(shfl, x[i] = __shfl_xor_sync(0xFFFFFFFFu, x[i], 1) + b)
OPK(ffma, x[i] = __float_as_uint(__fmaf_rn(__uint_as_float(x[i]), 1.0001f, __uint_as_float(b))))
OPK(fadd, x[i] = __float_as_uint(__fadd_rn(__uint_as_float(x[i]), __uint_as_float(b))))
OPK(f2i, x[i] = (unsigned)__float2int_rn(__uint_as_float(x[i])) + b)
// mixes: does an ALU-pipe op dual-issue with an FMA-pipe op?
OPK(mix_lop3_imad, if (i & 1) x[i] = x[i] * b + z; else x[i] = (x[i] & b) ^ z)
OPK(mix_lop3_dp4a, if (i & 1) x[i] = __dp4a(x[i], b, z); else x[i] = (x[i] & b) ^ z)
OPK(mix_vabs_imad, if (i & 1) x[i] = x[i] * b + z; else x[i] = __vabsdiffu4(x[i], b))
OPK(mix_popc_lop3, if (i & 1) x[i] = __popc(x[i]) + b; else x[i] = (x[i] & b) ^ z)
OPK(mix_vmin_imad, if (i & 1) x[i] = x[i] * b + z; else x[i] = __vmins2(x[i], b) + 1u)

__device__ __forceinline__ unsigned hmin2u(unsigned a, unsigned b) { __half2 r = __hmin2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b)); return *reinterpret_cast<unsigned*>(&r); }
__device__ __forceinline__ unsigned hmax2u(unsigned a, unsigned b) { __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b)); return *reinterpret_cast<unsigned*>(&r); }
OPK(hmnmx2, x[i] = hmax2u(hmin2u(x[i], b), z))
OPK(vimnmx3, x[i] = __vmins2(__vmins2(x[i], b), z) ^ 0x00010001u)
OPK(mix_vmin_hmin, if (i & 1) x[i] = hmax2u(hmin2u(x[i], b), z); else x[i] = __vmaxs2(__vmins2(x[i], b), z))
OPK(vmnmx_pair, x[i] = __vmaxs2(__vmins2(x[i], b), z))
OPK(lds, x[i] = sm[(x[i] + b) & 1023])
OPK(mix_lop3_iadd, if (i & 1) x[i] = x[i] + b + z; else x[i] = (x[i] & b) ^ z)

template <typename K>
static void run(const char* name, K kern, int sms, unsigned* out, long long* cyc, bool last)
{
    kern<<<sms, 1024>>>(out, cyc, 12345u);
    kern<<<sms, 1024>>>(out, cyc, 12345u);
    cudaDeviceSynchronize();
    long long h[256];
    cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double s = 0;
    for (int i = 0; i < sms; i++) s += (double)h[i];
    const double per_sm = (double)ILP * ITER * 1024 / (s / sms);
    printf("  \"%s\": %.1f%s\n", name, per_sm, last ? "" : ",");
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    unsigned* out;
    long long* cyc;
    cudaMalloc(&out, sizeof(unsigned) * 1024 * sms);
    cudaMalloc(&cyc, sizeof(long long) * sms);
    printf("{\n  \"device\": \"%s\", \"sms\": %d, \"unit\": \"thread-ops per clock per SM (128 = one warp instruction per cycle on each of the 4 SMSPs)\",\n", p.name, sms);
#define R(N, L) run(#N, k_##N, sms, out, cyc, L)
    R(iadd3, 0); R(lop3, 0); R(imad, 0); R(imad_hi, 0); R(shf, 0); R(prmt, 0); R(popc, 0); R(popc_only, 0); R(vabsdiff4, 0); R(vmins2, 0);
    R(vmax_s32, 0); R(dp4a, 0); R(dp2a, 0); R(shfl, 0); R(ffma, 0); R(fadd, 0); R(f2i, 0);
    R(mix_lop3_imad, 0); R(mix_lop3_dp4a, 0); R(mix_vabs_imad, 0); R(mix_popc_lop3, 0); R(mix_vmin_imad, 0); R(hmnmx2, 0); R(vimnmx3, 0); R(vmnmx_pair, 0); R(mix_vmin_hmin, 0); R(lds, 0); R(mix_lop3_iadd, 1);
    printf("}\n");
    return cudaGetLastError() != cudaSuccess;
}
