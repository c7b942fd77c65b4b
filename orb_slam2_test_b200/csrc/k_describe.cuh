/*
 * k_describe.cuh -- K4 + K6 + K7: intensity-centroid orientation, rotated-BRIEF descriptor and
 * output packing, one warp per surviving keypoint.
 *
 *   K4  IC_Angle (src/ORBextractor.cc:83-111): integer moments over the 31-px disc of the
 *       UNBLURRED level, lane r sums disc row v = r-15, warp-shuffle reduction, then
 *       cv::fastAtan2 (scalar float32 polynomial, every operation individually rounded).
 *   K6  computeOrbDescriptor (:117-157) on the BLURRED level: lane k builds descriptor byte k
 *       from 16 pattern points; rotation in the FMA-contracted form the reference's -O3
 *       -march=native object contains, sin/cos as glibc's sincosf (orb_sincos.cuh), cvRound =
 *       round-half-even.
 *   K7  operator() assembly (:1361-1395): levels concatenated in order, pt scaled by
 *       mvScaleFactor[level] for level > 0 (one float multiply), written as cv::KeyPoint records.
 */
#ifndef K_DESCRIBE_CUH
#define K_DESCRIBE_CUH

#include "borb_internal.h"
#include "orb_sincos.cuh"

struct BorbKpRec { /* == cv::KeyPoint == borb_keypoint */
    float x, y, size, angle, response;
    int octave, class_id;
};

__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float r2d = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = 0.9997878412794807f * r2d, p3 = -0.3258083974640975f * r2d,
                p5 = 0.1555786518463281f * r2d, p7 = -0.04432655554792128f * r2d;
    const float eps = (float)2.2204460492503131e-16; /* (float)DBL_EPSILON */
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__global__ void __launch_bounds__(256)
k_orient_describe(const BorbPlan *__restrict__ plan, const signed char *__restrict__ pattern,
                  const uint8_t *__restrict__ pyr, const uint8_t *__restrict__ blur,
                  const uint32_t *__restrict__ lkp_xy, const uint32_t *__restrict__ lkp_score,
                  const int *__restrict__ nlkp, BorbKpRec *__restrict__ out_kp,
                  uint8_t *__restrict__ out_desc, int *__restrict__ nout)
{
    __shared__ signed char pat[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) pat[i] = pattern[i];
    __syncthreads();

    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int nl = plan->nlevels;
    const int *cnt = nlkp + frame * BORB_MAXL;
    if (slot == 0 && lane == 0) {
        int t = 0;
        for (int l = 0; l < nl; ++l) t += cnt[l];
        nout[frame] = t;
    }
    if (slot >= plan->kp_total) return;
    int level = 0, base = 0;
    for (int l = 1; l < nl; ++l)
        if (slot >= plan->lv[l].kp_off) { level = l; }
    for (int l = 0; l < level; ++l) base += cnt[l];
    const BorbLevel &L = plan->lv[level];
    const int idx = slot - L.kp_off;
    if (idx >= cnt[level]) return;

    const uint32_t v = lkp_xy[(size_t)frame * plan->kp_total + slot];
    const int x = (int)(v & 0xffff), y = (int)(v >> 16);
    const size_t fo = (size_t)frame * (size_t)plan->pyr_bytes + L.off +
                      (size_t)(BORB_EDGE + y) * L.pitch + BORB_XPAD + x;

    /* ---- K4 ---- */
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int vrow = lane - 15;
        const int d = plan->umax[vrow < 0 ? -vrow : vrow];
        const uint8_t *row = pyr + fo + (ptrdiff_t)vrow * L.pitch;
        int rs = 0;
        for (int u = -d; u <= d; ++u) {
            const int p = row[u];
            rs += p;
            m10 += u * p;
        }
        m01 = vrow * rs;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    /* ---- K6 ---- */
    const float factor_pi = (float)(3.1415926535897932384626433832795 / 180.f);
    float a, b;
    orb_sincosf_exact(__fmul_rn(angle, factor_pi), &b, &a);
    const uint8_t *c = blur + fo;
    const signed char *p = pat + lane * 32;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        int t[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float px = (float)p[4 * k + 2 * e], py = (float)p[4 * k + 2 * e + 1];
            const int row = __float2int_rn(__fmaf_rn(px, b, __fmul_rn(py, a)));
            const int col = __float2int_rn(__fmaf_rn(px, a, -__fmul_rn(py, b)));
            t[e] = c[(ptrdiff_t)row * L.pitch + col];
        }
        val |= (t[0] < t[1]) << k;
    }
    const int o = base + idx;
    out_desc[((size_t)frame * plan->kp_total + o) * 32 + lane] = (uint8_t)val;

    /* ---- K7 ---- */
    if (lane == 0) {
        BorbKpRec r;
        r.x = (float)x;
        r.y = (float)y;
        if (level != 0) {
            r.x = __fmul_rn(r.x, L.scale);
            r.y = __fmul_rn(r.y, L.scale);
        }
        r.size = L.patch_size;
        r.angle = angle;
        r.response = (float)lkp_score[(size_t)frame * plan->kp_total + slot];
        r.octave = level;
        r.class_id = -1;
        out_kp[(size_t)frame * plan->kp_total + o] = r;
    }
}

#endif
