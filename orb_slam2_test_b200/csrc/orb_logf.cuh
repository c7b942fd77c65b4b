/*
 * orb_logf.cuh -- logf of a float, bit-identical to what the reference binary gets.
 *
 * MapPoint::PredictScale (src/MapPoint.cc:557-590) computes ceil(log(ratio) / mfLogScaleFactor) on floats; with
 * `using namespace std` in effect (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:36, reached through ORBmatcher.h) the call
 * is the float overload, i.e. the host libm's logf.  glibc's logf is not correctly rounded (it is the ARM optimized-routines
 * algorithm: 16-entry table of 1/c and log c, degree-3 polynomial in double, one final rounding; sysdeps/ieee754/flt-32/
 * e_logf.c since glibc 2.27), so "double log, then cast" is not good enough when the quotient sits next to an integer.
 * This is that algorithm with the table of the image's glibc 2.39 (read out of libm.so.6's rodata).  Checked against the
 * host libm on ALL positive normal floats: 0 differences, with or without the multiply-adds fused (the x86-64 FMA ifunc
 * variant and the plain one agree everywhere), so the form below -- every operation rounded on its own -- stands for both.
 * tests/test_logf.py repeats the check on a sample through borb_debug_logf (host copy) and borb_debug_logf_device.
 */
#ifndef ORB_LOGF_CUH
#define ORB_LOGF_CUH

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define ORB_LOGF_HD __host__ __device__ __forceinline__
#else
#define ORB_LOGF_HD static inline
#endif

#ifdef __CUDA_ARCH__
#define ORB_LOGF_MUL(a, b) __dmul_rn((a), (b))
#define ORB_LOGF_ADD(a, b) __dadd_rn((a), (b))
#else
#define ORB_LOGF_MUL(a, b) ((a) * (b)) /* host objects are built without contraction */
#define ORB_LOGF_ADD(a, b) ((a) + (b))
#endif

ORB_LOGF_HD float orb_logf_exact(float x)
{
    /* {1/c, log c} for the 16 sub-intervals of [0x1.66p-1, 0x1.66p0) */
    const double T[16][2] = {
        {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
        {0x1.49539f0f010b0p+0, -0x1.01eae7f513a67p-2}, {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
        {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8ea0p+0, -0x1.1aa2bc79c8100p-3},
        {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
        {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
        {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aa0p-1, 0x1.c5e53aa362eb4p-4},
        {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d224770p-3},
        {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};
    const double Ln2 = 0x1.62e42fefa39efp-1, A0 = -0x1.00ea348b88334p-2, A1 = 0x1.5575b0be00b6ap-2, A2 = -0x1.ffffef20a4123p-2;
    uint32_t ix;
#ifdef __CUDA_ARCH__
    ix = __float_as_uint(x);
#else
    memcpy(&ix, &x, 4);
#endif
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) { /* x < 0x1p-126, infinite or NaN */
        if (ix * 2u == 0u) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return NAN;
        const float xs = x * 0x1p23f; /* subnormal: normalise (exact) */
#ifdef __CUDA_ARCH__
        ix = __float_as_uint(xs);
#else
        memcpy(&ix, &xs, 4);
#endif
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int32_t)tmp >> 23; /* arithmetic shift */
    const uint32_t iz = ix - (tmp & 0xff800000u);
    float zf;
#ifdef __CUDA_ARCH__
    zf = __uint_as_float(iz);
#else
    memcpy(&zf, &iz, 4);
#endif
    const double z = (double)zf, invc = T[i][0], logc = T[i][1];
    const double r = ORB_LOGF_ADD(ORB_LOGF_MUL(z, invc), -1.0);
    const double y0 = ORB_LOGF_ADD(logc, ORB_LOGF_MUL((double)k, Ln2));
    const double r2 = ORB_LOGF_MUL(r, r);
    double y = ORB_LOGF_ADD(ORB_LOGF_MUL(A1, r), A2);
    y = ORB_LOGF_ADD(ORB_LOGF_MUL(A0, r2), y);
    y = ORB_LOGF_ADD(ORB_LOGF_MUL(y, r2), ORB_LOGF_ADD(y0, r));
    return (float)y;
}

/* (int)f as x86-64's cvttss2si gives it (the reference's `int nScale = ceil(...)`): out-of-range and NaN become INT_MIN */
ORB_LOGF_HD int orb_cvtt_x86(float f)
{
    if (!(f > -2147483904.0f && f < 2147483648.0f)) return (int)0x80000000u;
    return (int)f;
}

#endif
