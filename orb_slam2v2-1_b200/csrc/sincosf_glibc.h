// sincosf_glibc.h -- bit-exact restatement of glibc 2.39's sinf()/cosf() for arguments in [0, 2*pi].
//
// computeOrbDescriptor (reference src/ORBextractor.cc:112-113) evaluates cos/sin on a float inside
// `using namespace std`, i.e. glibc's cosf/sinf.  Those are not correctly rounded (<= 0.56 ulp), so a
// bit-exact descriptor needs the same algorithm, not CUDA's cosf/sinf: glibc's is the ARM optimized-routines
// design -- argument reduction and a degree-7/8 polynomial, all in double, rounded once to float.  The
// coefficient table below was read out of this image's libm.so.6 (__sincosf_table); every double operation is
// individually rounded (no FMA).  tests/native/sincosf_check.cpp (run by tests/test_abi.py::
// test_sincosf_restatement_matches_glibc) checks this header, compiled for the host, against the running glibc over all 1 086 953 882 floats in [0, 6.3]: 0 mismatches for both functions.
//
// ORBX_HD functions: usable from device code (with __dmul_rn/__dadd_rn so nvcc cannot contract) and from
// host code (compile with -ffp-contract=off).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define ORBX_HD __host__ __device__ __forceinline__
#else
#define ORBX_HD static inline
#endif

namespace orbx_sc {

#if defined(__CUDA_ARCH__)
ORBX_HD double dmul(double a, double b) { return __dmul_rn(a, b); }
ORBX_HD double dadd(double a, double b) { return __dadd_rn(a, b); }
ORBX_HD uint32_t fbits(float f) { return __float_as_uint(f); }
#else
ORBX_HD double dmul(double a, double b) { volatile double r = a * b; return r; }
ORBX_HD double dadd(double a, double b) { volatile double r = a + b; return r; }
ORBX_HD uint32_t fbits(float f) { union { float f; uint32_t u; } c; c.f = f; return c.u; }
#endif

// polynomial coefficients: cosine c0..c4 (even powers), sine s1..s3 (x^3, x^5, x^7)
#define ORBX_SC_C0 0x1p0
#define ORBX_SC_C1 (-0x1.ffffffd0c621cp-2)
#define ORBX_SC_C2 0x1.55553e1068f19p-5
#define ORBX_SC_C3 (-0x1.6c087e89a359dp-10)
#define ORBX_SC_C4 0x1.99343027bf8c3p-16
#define ORBX_SC_S1 (-0x1.555545995a603p-3)
#define ORBX_SC_S2 0x1.1107605230bc4p-7
#define ORBX_SC_S3 (-0x1.994eb3774cf24p-13)
#define ORBX_SC_HPI_INV 0x1.45F306DC9C883p+23   /* 2/pi * 2^24 */
#define ORBX_SC_HPI 0x1.921FB54442D18p0         /* pi/2 */

// n even: sine polynomial of x; n odd: cosine polynomial.  neg selects the negated cosine table.
ORBX_HD float poly(double x, double x2, int n, bool neg)
{
    if ((n & 1) == 0) {
        const double x3 = dmul(x, x2);
        const double s1 = dadd(ORBX_SC_S2, dmul(x2, ORBX_SC_S3));
        const double x7 = dmul(x3, x2);
        const double s = dadd(x, dmul(x3, ORBX_SC_S1));
        return (float)dadd(s, dmul(x7, s1));
    }
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = dmul(x2, x2);
    const double c2 = dadd(sg * ORBX_SC_C3, dmul(x2, sg * ORBX_SC_C4));
    const double c1 = dadd(sg * ORBX_SC_C0, dmul(x2, sg * ORBX_SC_C1));
    const double x6 = dmul(x4, x2);
    const double c = dadd(c1, dmul(x4, sg * ORBX_SC_C2));
    return (float)dadd(c, dmul(x6, c2));
}

ORBX_HD uint32_t abstop12(float f) { return (fbits(f) >> 20) & 0x7ff; }

// valid for 0 <= y < 120 (the descriptor's angles are in [0, 2*pi])
ORBX_HD void sincosf_glibc(float y, float* sn, float* cs)
{
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {          // |y| < pi/4
        const double x2 = dmul(x, x);
        if (abstop12(y) < abstop12(0x1p-12f)) { *sn = y; *cs = 1.0f; return; }
        *sn = poly(x, x2, 0, false);
        *cs = poly(x, x2, 1, false);
        return;
    }
    const double r = dmul(x, ORBX_SC_HPI_INV);
    const int n = ((int32_t)r + 0x800000) >> 24;
    x = dadd(x, -dmul((double)n, ORBX_SC_HPI));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;   // sign table {1,-1,-1,1}
    const bool neg = (n & 2) != 0;
    const double xs = dmul(x, sgn), x2 = dmul(x, x);
    *sn = poly(xs, x2, n, neg);
    *cs = poly(xs, x2, n ^ 1, neg);
}

}  // namespace orbx_sc
