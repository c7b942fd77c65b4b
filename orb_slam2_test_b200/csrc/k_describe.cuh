/*
 * k_describe.cuh -- K4 + K6 + K7: intensity-centroid orientation, rotated-BRIEF descriptor and
 * output packing, one warp per surviving keypoint (a warp walks several keypoints; the 8 pattern
 * words of its lanes stay in registers).
 *
 *   K4  IC_Angle (src/ORBextractor.cc:83-111): integer moments over the 31-px disc of the
 *       UNBLURRED level.  Lane u owns disc COLUMN u-15 and walks the 31 rows, so every load is one
 *       coalesced row segment; m10 = sum_u u * colsum_u and m01 = sum_u sum_v v*p by warp-shuffle
 *       reduction, then cv::fastAtan2 (scalar float32 polynomial, every operation individually
 *       rounded).
 *   K6  computeOrbDescriptor (:117-157) on the BLURRED level: the 37x37 patch the rotated pattern
 *       can reach (|offset| <= 18) is staged into shared memory with aligned word loads, then lane
 *       k builds descriptor byte k from 16 pattern points gathered from the staged patch;
 *       rotation in the FMA-contracted form the reference's -O3 -march=native object contains,
 *       sin/cos as glibc's sincosf (orb_sincos.cuh), cvRound = round-half-even.
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

#ifndef DESC_NT
#define DESC_NT 128
#endif
#define DESC_NW (DESC_NT / 32)
#define DESC_PR 37  /* staged patch rows: -18 .. +18 */
#define DESC_PW 10  /* staged words per row: covers 37 columns at any alignment */
#ifndef DESC_KPW
#define DESC_KPW 16 /* keypoint slots per warp */
#endif
#ifndef DESC_MINB
#define DESC_MINB 8 /* 64 registers: the kernel is latency bound, occupancy beats the few spills (measured) */
#endif

__global__ void __launch_bounds__(DESC_NT, DESC_MINB)
k_orient_describe(const BorbPlan *__restrict__ plan, const signed char *__restrict__ pattern,
                  const uint8_t *__restrict__ pyr, const uint8_t *__restrict__ blur,
                  const uint32_t *__restrict__ lkp_xy, const uint32_t *__restrict__ lkp_score,
                  const int *__restrict__ nlkp, BorbKpRec *__restrict__ out_kp,
                  uint8_t *__restrict__ out_desc, int *__restrict__ nout)
{
    __shared__ uint32_t patch_all[DESC_NW][DESC_PR * DESC_PW];
    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *patch = patch_all[warp];
    const int nl = plan->nlevels;
    const int kp_total = plan->kp_total;

    /* the 8 pattern words of this lane: bit k of descriptor byte `lane` compares the blurred
     * image at pattern points (xa, ya) and (xb, yb), packed as 4 int8 */
    uint32_t pw[8];
    {
        const uint4 p0 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2);
        const uint4 p1 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2 + 1);
        pw[0] = p0.x; pw[1] = p0.y; pw[2] = p0.z; pw[3] = p0.w;
        pw[4] = p1.x; pw[5] = p1.y; pw[6] = p1.z; pw[7] = p1.w;
    }
    /* lane l < nl: survivors of level l, slot offset of level l, output offset of level l */
    const int my_cnt = lane < nl ? nlkp[frame * BORB_MAXL + lane] : 0;
    const int my_off = lane < nl ? plan->lv[lane].kp_off : 0x7fffffff;
    int my_base = my_cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, my_base, o);
        if (lane >= o) my_base += t;
    }
    if (blockIdx.x == 0 && warp == 0 && lane == 31) nout[frame] = my_base; /* total of the frame */
    my_base -= my_cnt; /* exclusive */

    /* per-lane constants: staging position (3 rows x 10 words per step), disc column */
    const int st_r = lane / DESC_PW, st_w = lane - st_r * DESC_PW;
    const int ic_u = lane - 15;
    const int ic_vmax = lane < 31 ? plan->umax[ic_u < 0 ? -ic_u : ic_u] : -1;

    const int warps_per_frame = gridDim.x * DESC_NW;
    for (int slot = blockIdx.x * DESC_NW + warp; slot < kp_total; slot += warps_per_frame) {
        const int level = __popc(__ballot_sync(0xffffffffu, slot >= my_off)) - 1;
        const int idx = slot - __shfl_sync(0xffffffffu, my_off, level);
        if (idx >= __shfl_sync(0xffffffffu, my_cnt, level)) continue;
        const int o = __shfl_sync(0xffffffffu, my_base, level) + idx;
        const BorbLevel &L = plan->lv[level];
        const int pitch = L.pitch;

        const uint32_t v = lkp_xy[(size_t)frame * kp_total + slot];
        const int x = (int)(v & 0xffff), y = (int)(v >> 16);
        const size_t lo = (size_t)frame * (size_t)plan->pyr_bytes + L.off + (size_t)BORB_EDGE * pitch + BORB_XPAD;

        /* ---- K6 staging first (independent of the angle): blurred rows y-18..y+18, columns from
         * the aligned word holding x-18.  Lanes 0..29 cover 3 rows x 10 words per step. ---- */
        const int xa = (x - 18) & ~3, xo = (x - 18) - xa;
        if (lane < 30) {
            const uint8_t *bp = blur + lo + (ptrdiff_t)(y - 18 + st_r) * pitch + xa + 4 * st_w;
            const ptrdiff_t adv = (ptrdiff_t)3 * pitch;
            uint32_t *pp = patch + st_r * DESC_PW + st_w;
#pragma unroll
            for (int it = 0; it < 12; ++it) {
                pp[it * 3 * DESC_PW] = __ldg(reinterpret_cast<const uint32_t *>(bp));
                bp += adv;
            }
            if (st_r == 0) pp[12 * 3 * DESC_PW] = __ldg(reinterpret_cast<const uint32_t *>(bp)); /* row 36 */
        }

        /* ---- K4: lane u-15 owns disc column u and the rows |v| <= umax[|u|] (the disc is
         * symmetric: column u reaches as far as row u does, src/ORBextractor.cc:504-520) ---- */
        int m10 = 0, m01 = 0;
        if (lane < 31) {
            const uint8_t *dn = pyr + lo + (ptrdiff_t)y * pitch + x + ic_u;
            const uint8_t *up = dn;
            int colsum = *dn;
#pragma unroll
            for (int vr = 1; vr <= 15; ++vr) {
                dn += pitch;
                up -= pitch;
                if (vr <= ic_vmax) {
                    const int pd = *dn, pu = *up;
                    colsum += pd + pu;
                    m01 += vr * (pd - pu);
                }
            }
            m10 = ic_u * colsum;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, s);
            m01 += __shfl_xor_sync(0xffffffffu, m01, s);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        /* ---- K6 ---- */
        const float factor_pi = (float)(3.1415926535897932384626433832795 / 180.f);
        float a, b;
        orb_sincosf_exact(__fmul_rn(angle, factor_pi), &b, &a);
        __syncwarp();
        const uint8_t *c = reinterpret_cast<const uint8_t *>(patch) + 18 * (DESC_PW * 4) + xo + 18;
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int t[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const float px = (float)(int)(signed char)(pw[k] >> (16 * e));
                const float py = (float)(int)(signed char)(pw[k] >> (16 * e + 8));
                const int row = __float2int_rn(__fmaf_rn(px, b, __fmul_rn(py, a)));
                const int col = __float2int_rn(__fmaf_rn(px, a, -__fmul_rn(py, b)));
                t[e] = c[row * (DESC_PW * 4) + col];
            }
            val |= (t[0] < t[1]) << k;
        }
        out_desc[((size_t)frame * kp_total + o) * 32 + lane] = (uint8_t)val;
        __syncwarp(); /* the patch is rewritten by the next keypoint */

        /* ---- K7: lanes 0..6 write one 4-byte field each of the cv::KeyPoint record ---- */
        if (lane < 7) {
            const float s = level != 0 ? L.scale : 1.f;
            uint32_t f;
            if (lane == 0) f = __float_as_uint(level != 0 ? __fmul_rn((float)x, s) : (float)x);
            else if (lane == 1) f = __float_as_uint(level != 0 ? __fmul_rn((float)y, s) : (float)y);
            else if (lane == 2) f = __float_as_uint(L.patch_size);
            else if (lane == 3) f = __float_as_uint(angle);
            else if (lane == 4) f = __float_as_uint((float)lkp_score[(size_t)frame * kp_total + slot]);
            else if (lane == 5) f = (uint32_t)level;
            else f = 0xffffffffu; /* class_id = -1 */
            reinterpret_cast<uint32_t *>(out_kp + (size_t)frame * kp_total + o)[lane] = f;
        }
    }
}

#endif
