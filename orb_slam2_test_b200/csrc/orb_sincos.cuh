/*
 * orb_sincos.cuh -- sin/cos of a float angle, bit-identical to what the reference binary gets.
 *
 * The reference rotates the BRIEF pattern with (float)cos(angle), (float)sin(angle)
 * (src/ORBextractor.cc:121-122); with `using namespace std` and the reference's -O3 build GCC
 * merges the two into ONE sincosf() call.  glibc's sincosf is not correctly rounded (about 0.09 %
 * of the angles in [0, 2*pi) differ from the rounded double result by one ulp), so "double
 * sincos, then cast" is not good enough for bit-exact descriptors.  This is the published glibc
 * >= 2.28 algorithm (sysdeps/ieee754/flt-32/s_sincosf.c: range reduction by pi/2 in double,
 * degree-7/8 minimax polynomials in double, single final rounding), written with the
 * multiply-adds fused exactly where the x86-64 FMA build of glibc 2.39 (the image's libm,
 * __sincosf_fma, selected by ifunc on every FMA-capable host) fuses them.  Checked against the
 * host libm on 1.6e8 angles: 0 differences (tests/test_sincos.py).
 *
 * Valid for |x| < 120 (the path only produces [0, 2*pi]).
 */
#ifndef ORB_SINCOS_CUH
#define ORB_SINCOS_CUH

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define ORB_HD __host__ __device__ __forceinline__
#else
#define ORB_HD static inline
#endif

ORB_HD double orb_fma_(double a, double b, double c)
{
#ifdef __CUDA_ARCH__
    return __fma_rn(a, b, c);
#else
    return fma(a, b, c);
#endif
}
ORB_HD double orb_mul_(double a, double b)
{
#ifdef __CUDA_ARCH__
    return __dmul_rn(a, b);
#else
    return a * b; /* host objects are built with -ffp-contract=off */
#endif
}

ORB_HD void orb_sincosf_exact(float y, float *sinp, float *cosp)
{
    /* polynomial tables for even / odd multiples of pi (second = negated cosine terms) */
    const double C0 = 1.0, C1 = -0x1.ffffffd0c621cp-2, S1 = -0x1.555545995a603p-3,
                 C2 = 0x1.55553e1068f19p-5, S2 = 0x1.1107605230bc4p-7,
                 C3 = -0x1.6c087e89a359dp-10, S3 = -0x1.994eb3774cf24p-13,
                 C4 = 0x1.99343027bf8c3p-16;
    uint32_t bits;
#ifdef __CUDA_ARCH__
    bits = __float_as_uint(y);
#else
    memcpy(&bits, &y, 4);
#endif
    const uint32_t top = (bits >> 20) & 0x7ff;
    const double x = (double)y;
    double xr = x, sgn = 1.0;
    int n = 0;
    if (top < 0x3f4) {          /* |y| < pi/4 */
        if (top < 0x398) {      /* |y| < 2^-12 */
            *sinp = y;
            *cosp = 1.0f;
            return;
        }
    } else {                    /* pi/4 <= |y| < 120: n = round(y / (pi/2)) */
        const double r = orb_mul_(x, 0x1.45f306dc9c883p+23);
        n = ((int32_t)r + 0x800000) >> 24;
        xr = orb_fma_(-(double)n, 0x1.921fb54442d18p+0, x);
        sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    }
    const double f = (n & 2) ? -1.0 : 1.0; /* odd multiple of pi: cosine terms change sign */
    const double c0 = f * C0, c1 = f * C1, c2 = f * C2, c3 = f * C3, c4 = f * C4;
    const double X = orb_mul_(xr, sgn), x2 = orb_mul_(xr, xr);
    const double s1v = orb_fma_(x2, S3, S2), c2v = orb_fma_(x2, c4, c3);
    const double x3 = orb_mul_(x2, X), x4 = orb_mul_(x2, x2);
    const double x5 = orb_mul_(x2, x3), x6 = orb_mul_(x2, x4);
    const double c1v = orb_fma_(x2, c1, c0);
    const double sv = orb_fma_(x3, S1, X), cv = orb_fma_(x4, c2, c1v);
    const float S = (float)orb_fma_(s1v, x5, sv), C = (float)orb_fma_(c2v, x6, cv);
    if (n & 1) { *sinp = C; *cosp = S; }
    else       { *sinp = S; *cosp = C; }
}

#endif
