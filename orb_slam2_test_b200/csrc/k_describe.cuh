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
 *       can reach (|offset| <= 18) is staged into shared memory by one TMA box load, then lane
 *       k builds descriptor byte k from 16 pattern points gathered from the staged patch;
 *       rotation in the FMA-contracted form the reference's -O3 -march=native object contains,
 *       sin/cos as glibc's sincosf (orb_sincos.cuh), cvRound = round-half-even.
 *   K7  operator() assembly (:1361-1395): levels concatenated in order, pt scaled by
 *       mvScaleFactor[level] for level > 0 (one float multiply), written as cv::KeyPoint records.
 */
#ifndef K_DESCRIBE_CUH
#define K_DESCRIBE_CUH

#include <cuda.h>

#include "borb_internal.h"
#include "borb_cta.cuh"
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

/* K6, one descriptor byte: bit k compares the blurred image at the two rotated pattern points of pattern word
 * pw[k] (xa, ya, xb, yb as int8); c = the keypoint's pixel, pitch = bytes per row (src/ORBextractor.cc:117-157).
 * The rotation has the form the reference's -O3 -march=native object contains -- x*b + y*a contracted to
 * fma(x, b, rn(y*a)), x*a - y*b to fma(x, a, -rn(y*b)) --, sin/cos are glibc's sincosf, cvRound is round-half-even:
 * tests/test_gpu_edge.py holds inputs within an ulp of a rounding tie where the unfused form reads another pixel. */
__device__ __forceinline__ int orb_desc_byte(const uint8_t *c, int pitch, float angle_deg, const uint32_t *pw)
{
    const float factor_pi = (float)(3.1415926535897932384626433832795 / 180.f);
    float a, b;
    orb_sincosf_exact(__fmul_rn(angle_deg, factor_pi), &b, &a);
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
            t[e] = c[row * pitch + col];
        }
        val |= (t[0] < t[1]) << k;
    }
    return val;
}

/* parity tap (borb_debug_describe): K6 on a caller-supplied patch and angle, one warp */
__global__ void k_debug_describe(const uint8_t *__restrict__ patch, int pitch, int cx, int cy, float angle_deg,
                                 const signed char *__restrict__ pattern, uint8_t *__restrict__ desc)
{
    const int lane = threadIdx.x;
    uint32_t pw[8];
    const uint4 p0 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2);
    const uint4 p1 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2 + 1);
    pw[0] = p0.x; pw[1] = p0.y; pw[2] = p0.z; pw[3] = p0.w;
    pw[4] = p1.x; pw[5] = p1.y; pw[6] = p1.z; pw[7] = p1.w;
    desc[lane] = (uint8_t)orb_desc_byte(patch + (size_t)cy * pitch + cx, pitch, angle_deg, pw);
}

#ifndef DESC_NT
#define DESC_NT 128
#endif
#define DESC_NW (DESC_NT / 32)
#define DESC_PR 37  /* staged patch rows: -18 .. +18 */
#define DESC_PB 64  /* staged bytes per patch row: 37 columns from a 16-byte boundary (<= 15 + 37) */
#define DESC_IR 31  /* staged rows of the unblurred level for IC_Angle: -15 .. +15 */
#define DESC_IB 48  /* staged bytes per disc row: 31 columns from a 16-byte boundary (<= 15 + 31) */
#define DESC_PBYTES (DESC_PR * DESC_PB)
#define DESC_IBYTES (DESC_IR * DESC_IB)
#define DESC_POFF 0
#define DESC_IOFF ((DESC_PBYTES + 127) & ~127)
#define DESC_BUF ((DESC_IOFF + DESC_IBYTES + 127) & ~127) /* bytes per buffer, 128-B aligned (TMA) */
#define DESC_SMEM_BYTES (128 + DESC_NW * 2 * DESC_BUF)
#ifndef DESC_KPW
#define DESC_KPW 16 /* keypoint slots per warp */
#endif
#ifndef DESC_MINB
#define DESC_MINB 7 /* shared memory allows 7 CTAs per SM */
#endif

struct DescMaps { /* kernel parameter (__grid_constant__): TMA descriptors of the unblurred and the blurred
                     planes of every level (boxes DESC_IB x DESC_IR and DESC_PB x DESC_PR x 1 frame) and the
                     few plan fields the kernel reads */
    CUtensorMap pyr[BORB_MAXL], blur[BORB_MAXL];
    int nlevels, kp_total;
    int umax[16];
    int kp_off[BORB_MAXL];
    float scale[BORB_MAXL], patch_size[BORB_MAXL];
};

template <bool PDL>
__global__ void __launch_bounds__(DESC_NT, DESC_MINB)
k_orient_describe(const __grid_constant__ DescMaps dm, const signed char *__restrict__ pattern, const int frame0,
                  const uint32_t *__restrict__ lkp_xy_p, const uint32_t *__restrict__ lkp_score_p,
                  const int *__restrict__ nlkp_p, BorbKpRec *__restrict__ out_kp,
                  uint8_t *__restrict__ out_desc, int *__restrict__ nout)
{
    /* [0,64): mbarriers [warp][buffer]; per warp two buffers, each the blurred 37-row patch and the
     * unblurred 31-row disc region of one keypoint */
    extern __shared__ __align__(128) unsigned char desc_smem[];
    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned char *patch = desc_smem + 128 + warp * (2 * DESC_BUF);
    const uint32_t patch_s = (uint32_t)__cvta_generic_to_shared(patch);
    const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(desc_smem + 16 * warp);
    const int nl = dm.nlevels;
    const int kp_total = dm.kp_total;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s + 8) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    /* the 8 pattern words of this lane: bit k of descriptor byte `lane` compares the blurred
     * image at pattern points (xa, ya) and (xb, yb), packed as 4 int8 */
    uint32_t pw[8];
    {
        const uint4 p0 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2);
        const uint4 p1 = __ldg(reinterpret_cast<const uint4 *>(pattern) + lane * 2 + 1);
        pw[0] = p0.x; pw[1] = p0.y; pw[2] = p0.z; pw[3] = p0.w;
        pw[4] = p1.x; pw[5] = p1.y; pw[6] = p1.z; pw[7] = p1.w;
    }
    /* PDL: launched programmatically behind K3 (the single-frame shape).  What K3 wrote is then read with ordinary
     * loads through pointers that exist only after the wait (borb_cta.cuh); the other instantiation keeps the read-only
     * loads the compiler derives from `const __restrict__` (they are what it would hoist above the wait). */
    const uint32_t *lkp_xy = lkp_xy_p, *lkp_score = lkp_score_p;
    const int *nlkp = nlkp_p;
    if (PDL) {
        borb_pdl_wait(); /* K3 is complete: the survivors and their counts (the blurred planes come over a full dependency) */
        lkp_xy = borb_pdl_after_wait(lkp_xy_p);
        lkp_score = borb_pdl_after_wait(lkp_score_p);
        nlkp = borb_pdl_after_wait(nlkp_p);
    }
    /* lane l < nl: survivors of level l, slot offset of level l, output offset of level l */
    const int my_cnt = lane < nl ? nlkp[frame * BORB_MAXL + lane] : 0;
    const int my_off = lane < nl ? dm.kp_off[lane] : 0x7fffffff;
    int my_base = my_cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, my_base, o);
        if (lane >= o) my_base += t;
    }
    if (blockIdx.x == 0 && warp == 0 && lane == 31) nout[frame] = my_base; /* total of the frame */
    my_base -= my_cnt; /* exclusive */

    const int ic_u = lane - 15; /* disc column of this lane */
    const int ic_vmax = lane < 31 ? dm.umax[ic_u < 0 ? -ic_u : ic_u] : -1;

    const int warps_per_frame = gridDim.x * DESC_NW;
    const size_t fkp = (size_t)frame * kp_total;

    /* A warp walks its slots as a three-deep software pipeline: in the iteration that orients and
     * describes keypoint i, the blurred patch and the unblurred disc region of keypoint i+1 are on
     * their way into the other buffer (two TMA box loads issued by lane 0, completion on the
     * buffer's mbarrier), and the coordinates of keypoint i+2 are being loaded -- no global-load
     * latency is left on the critical path and staging costs the warp a handful of instructions
     * (v3 issued 23 four-byte cp.async per lane: 18 % of the kernel's instructions). */
    int slot = blockIdx.x * DESC_NW + warp;
    struct Item { int level, o, slot; uint32_t v; bool ok; };
/* resolve the next occupied slot at or after `slot` and start the load of its coordinates */
#define DESC_NEXT(I)                                                                         \
    do {                                                                                     \
        I.ok = false; I.level = 0; I.o = 0; I.slot = 0; I.v = 0;                             \
        while (slot < kp_total) {                                                            \
            const int lv_ = __popc(__ballot_sync(0xffffffffu, slot >= my_off)) - 1;          \
            const int idx_ = slot - __shfl_sync(0xffffffffu, my_off, lv_);                   \
            const int s_ = slot;                                                             \
            slot += warps_per_frame;                                                         \
            if (idx_ < __shfl_sync(0xffffffffu, my_cnt, lv_)) {                              \
                I.ok = true; I.level = lv_; I.slot = s_;                                     \
                I.o = __shfl_sync(0xffffffffu, my_base, lv_) + idx_;                         \
                I.v = lkp_xy[fkp + s_];                                                             \
                break;                                                                       \
            }                                                                                \
        }                                                                                    \
    } while (0)
/* into buffer `bufi`: blurred rows y-18..y+18 from the 16-byte boundary at or before x-18, unblurred
 * rows y-15..y+15 from the 16-byte boundary at or before x-15 (a TMA box starts on a 16-byte
 * boundary; BORB_XPAD is a multiple of 16); box parts past the padded plane are zero-filled and
 * never read */
#define DESC_TMA(dst_, map_, x_, y_, bar_)                                                                       \
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" \
                 ::"r"(dst_), "l"(map_), "r"(x_), "r"(y_), "r"(frame0 + frame), "r"(bar_) : "memory")
#define DESC_STAGE(I, bufi)                                                                  \
    do {                                                                                     \
        if (I.ok && lane == 0) {                                                             \
            const int xs = (int)(I.v & 0xffff), ys = (int)(I.v >> 16);                       \
            const uint32_t b_ = bar_s + 8 * (bufi), d_ = patch_s + (bufi) * DESC_BUF;        \
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b_), "r"(DESC_PBYTES + DESC_IBYTES) : "memory"); \
            DESC_TMA(d_ + DESC_POFF, &dm.blur[I.level], BORB_XPAD + ((xs - 18) & ~15), BORB_EDGE + ys - 18, b_); \
            DESC_TMA(d_ + DESC_IOFF, &dm.pyr[I.level], BORB_XPAD + ((xs - 15) & ~15), BORB_EDGE + ys - 15, b_);  \
        }                                                                                    \
    } while (0)

    Item cur, nxt;
    DESC_NEXT(cur);
    DESC_STAGE(cur, 0);
    DESC_NEXT(nxt);
    int buf = 0;
    uint32_t it = 0;
    while (cur.ok) {
        DESC_STAGE(nxt, buf ^ 1);
        Item nn;
        DESC_NEXT(nn);

        const int level = cur.level, o = cur.o;
        const int x = (int)(cur.v & 0xffff), y = (int)(cur.v >> 16);
        const unsigned char *bufp = patch + buf * DESC_BUF;
        {   /* this keypoint's two regions have landed */
            const uint32_t b_ = bar_s + 8 * buf, par_ = (it >> 1) & 1u;
            asm volatile("{\n\t.reg .pred p;\n\tDESC_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                         "@p bra DESC_DONE;\n\tbra DESC_WAIT;\n\tDESC_DONE:\n\t}" ::"r"(b_), "r"(par_) : "memory");
        }
        ++it;

        /* ---- K4: lane u-15 owns disc column u and the rows |v| <= umax[|u|] (the disc is
         * symmetric: column u reaches as far as row u does, src/ORBextractor.cc:504-520) ---- */
        int m10 = 0, m01 = 0;
        if (lane < 31) {
            const uint8_t *dn = bufp + DESC_IOFF + 15 * DESC_IB + ((x - 15) & 15) + lane;
            const uint8_t *up = dn;
            int colsum = *dn;
#pragma unroll
            for (int vr = 1; vr <= 15; ++vr) {
                dn += DESC_IB;
                up -= DESC_IB;
                if (vr <= ic_vmax) {
                    const int pd = *dn, pu = *up;
                    colsum += pd + pu;
                    m01 += vr * (pd - pu);
                }
            }
            m10 = ic_u * colsum;
        }
        m10 = __reduce_add_sync(0xffffffffu, m10); /* integer sums: one REDUX each instead of a five-step butterfly */
        m01 = __reduce_add_sync(0xffffffffu, m01);
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        /* ---- K6 ---- */
        const int val = orb_desc_byte(bufp + DESC_POFF + 18 * DESC_PB + ((x - 18) & 15) + 18, DESC_PB, angle, pw);
        out_desc[((size_t)frame * kp_total + o) * 32 + lane] = (uint8_t)val;
        __syncwarp(); /* this buffer is refilled during the next iteration */

        /* ---- K7: lanes 0..6 write one 4-byte field each of the cv::KeyPoint record ---- */
        if (lane < 7) {
            const float s = level != 0 ? dm.scale[level] : 1.f;
            uint32_t f;
            if (lane == 0) f = __float_as_uint(level != 0 ? __fmul_rn((float)x, s) : (float)x);
            else if (lane == 1) f = __float_as_uint(level != 0 ? __fmul_rn((float)y, s) : (float)y);
            else if (lane == 2) f = __float_as_uint(dm.patch_size[level]);
            else if (lane == 3) f = __float_as_uint(angle);
            else if (lane == 4) f = __float_as_uint((float)lkp_score[fkp + cur.slot]);
            else if (lane == 5) f = (uint32_t)level;
            else f = 0xffffffffu; /* class_id = -1 */
            reinterpret_cast<uint32_t *>(out_kp + (size_t)frame * kp_total + o)[lane] = f;
        }
        cur = nxt;
        nxt = nn;
        buf ^= 1;
    }
#undef DESC_NEXT
#undef DESC_STAGE
#undef DESC_TMA
}

#endif
