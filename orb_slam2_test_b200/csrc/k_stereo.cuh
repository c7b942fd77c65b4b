/*
 * k_stereo.cuh -- S1/S2: Frame::ComputeStereoMatches on the device (src/Frame.cc:619-834), the
 * immediate consumer of BOTH extractors' keypoints, descriptors and pyramids.
 * Running it where those already live removes the two pyramid downloads from the stereo path.
 *
 *   S1  one warp per left keypoint.  (a) The reference's per-row candidate tables (:627-645) are
 *       replaced by a scan of all right keypoints: lane i tests right keypoints i, i+32, ... for
 *       "row band contains (size_t)yL", octave within +-1 and x inside [uL - maxD, uL]; 256-bit
 *       Hamming distance by popc (= ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1846-1862);
 *       the first smallest distance in right-index order wins, as the reference's strict `<` over
 *       its index-ordered candidate list does.  (b) If it is below (TH_HIGH+TH_LOW)/2, 11 shifted
 *       11x11 SADs on the keypoint's pyramid level (centre-subtracted, :697-740), parabola fit,
 *       disparity and depth with every float operation individually rounded (:742-781).
 *   S2  one CTA per frame pair: the median SAD (element size/2 of the sorted list, :786-789) by a
 *       counting bisection, then the 1.5*1.4*median outlier cut (:792-803).
 */
#ifndef K_STEREO_CUH
#define K_STEREO_CUH

#include "borb_internal.h"
#include "k_describe.cuh"

#define ST_NT 128

__global__ void __launch_bounds__(ST_NT)
k_stereo_match(const BorbPlan *__restrict__ plan, const BorbKpRec *__restrict__ kp_l,
               const uint8_t *__restrict__ desc_l, const int *__restrict__ n_l,
               const BorbKpRec *__restrict__ kp_r, const uint8_t *__restrict__ desc_r,
               const int *__restrict__ n_r, const uint8_t *__restrict__ pyr_l,
               const uint8_t *__restrict__ pyr_r, float mbf, float mb, float *__restrict__ u_right,
               float *__restrict__ depth, int *__restrict__ sad_out)
{
    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int o = blockIdx.x * (ST_NT / 32) + (threadIdx.x >> 5);
    const int kp_total = plan->kp_total;
    if (o >= n_l[frame]) return;
    const size_t fo = (size_t)frame * kp_total;
    const BorbKpRec kl = kp_l[fo + o];
    float res_u = -1.f, res_z = -1.f;
    int res_sad = -1;

    const int lv = kl.octave;
    const float ul = kl.x, vl = kl.y;
    const int row = (int)vl; /* (size_t)vL of a non-negative float */
    const float max_d = __fdiv_rn(mbf, mb);
    const float min_u = __fsub_rn(ul, max_d), max_u = ul;
    uint32_t best = (100u << 16) | 0xffffu; /* TH_HIGH, no index */
    if (row < plan->lv[0].h && !(max_u < 0.f)) {
        const int nr = n_r[frame];
        uint32_t dl[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) dl[k] = reinterpret_cast<const uint32_t *>(desc_l + (fo + o) * 32)[k];
        for (int ir = lane; ir < nr; ir += 32) {
            const BorbKpRec &q = kp_r[fo + ir];
            const int oc = q.octave;
            const float y = q.y, r = __fmul_rn(2.0f, plan->lv[oc].scale);
            const int maxr = (int)ceilf(__fadd_rn(y, r)), minr = (int)floorf(__fsub_rn(y, r));
            if (row < minr || row > maxr || oc < lv - 1 || oc > lv + 1) continue;
            const float ur = q.x;
            if (!(ur >= min_u && ur <= max_u)) continue;
            const uint32_t *dr = reinterpret_cast<const uint32_t *>(desc_r + (fo + ir) * 32);
            int d = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) d += __popc(dl[k] ^ dr[k]);
            const uint32_t key = ((uint32_t)d << 16) | (uint32_t)ir;
            if (d < (int)(best >> 16)) best = key; /* strict: the earliest index of this lane stays */
        }
    }
    best = __reduce_min_sync(0xffffffffu, best); /* (distance, index): first smallest in index order */
    const int best_dist = (int)(best >> 16);
    if (best_dist < (100 + 50) / 2) {
        const int best_r = (int)(best & 0xffffu);
        const BorbLevel &L = plan->lv[lv];
        const float ur0 = kp_r[fo + best_r].x, sf = L.inv_scale;
        const float su = roundf(__fmul_rn(ul, sf)), sv = roundf(__fmul_rn(vl, sf)), sr = roundf(__fmul_rn(ur0, sf));
        const int w = 5, LL = 5;
        const float iniu = sr + (float)(LL - w), endu = sr + (float)(LL + w + 1);
        if (!(iniu < 0.f || endu >= (float)L.w)) {
            const int cu = (int)su, cv = (int)sv, cr = (int)sr;
            const size_t lo = (size_t)frame * (size_t)plan->pyr_bytes + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
            const uint8_t *pl = pyr_l + lo + (ptrdiff_t)cv * L.pitch + cu;
            const uint8_t *pr = pyr_r + lo + (ptrdiff_t)cv * L.pitch + cr;
            const int cl = pl[0];
            int sad[11];
#pragma unroll
            for (int i = 0; i < 11; ++i) sad[i] = 0;
            for (int p = lane; p < 121; p += 32) {
                const int dy = p / 11 - w, dx = p - (dy + w) * 11 - w;
                const int a = (int)pl[dy * L.pitch + dx] - cl;
                const uint8_t *q = pr + dy * L.pitch + dx;
#pragma unroll
                for (int i = 0; i < 11; ++i) {
                    const int b = (int)q[i - LL] - (int)pr[i - LL];
                    sad[i] += abs(a - b);
                }
            }
#pragma unroll
            for (int i = 0; i < 11; ++i) sad[i] = __reduce_add_sync(0xffffffffu, sad[i]);
            int best_sad = 0x7fffffff, best_inc = 0;
#pragma unroll
            for (int i = 0; i < 11; ++i)
                if (sad[i] < best_sad) { best_sad = sad[i]; best_inc = i - LL; }
            if (best_inc != -LL && best_inc != LL) {
                float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                for (int i = 1; i < 10; ++i)
                    if (i - LL == best_inc) { d1 = (float)sad[i - 1]; d2 = (float)sad[i]; d3 = (float)sad[i + 1]; }
                const float delta = __fdiv_rn(__fsub_rn(d1, d3),
                                              __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                if (!(delta < -1.f || delta > 1.f)) {
                    float best_ur = __fmul_rn(L.scale, __fadd_rn(__fadd_rn(sr, (float)best_inc), delta));
                    float disparity = __fsub_rn(ul, best_ur);
                    if (disparity >= 0.f && disparity < max_d) {
                        if (disparity <= 0.f) {
                            disparity = 0.01f;
                            best_ur = (float)((double)ul - 0.01);
                        }
                        res_z = __fdiv_rn(mbf, disparity);
                        res_u = best_ur;
                        res_sad = best_sad;
                    }
                }
            }
        }
    }
    if (lane == 0) {
        u_right[fo + o] = res_u;
        depth[fo + o] = res_z;
        sad_out[fo + o] = res_sad;
    }
}

__global__ void __launch_bounds__(256)
k_stereo_cut(const BorbPlan *__restrict__ plan, const int *__restrict__ n_l, float *__restrict__ u_right,
             float *__restrict__ depth, const int *__restrict__ sad, int *__restrict__ n_kept)
{
    __shared__ int s_cnt;
    const int frame = blockIdx.x, tid = threadIdx.x;
    const size_t fo = (size_t)frame * plan->kp_total;
    const int n = n_l[frame];
    int m = 0;
    for (int i = tid; i < n; i += 256) m += sad[fo + i] >= 0;
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    atomicAdd(&s_cnt, m);
    __syncthreads();
    m = s_cnt;
    __syncthreads();
    if (m == 0) { if (tid == 0) n_kept[frame] = 0; return; }
    /* median = smallest v with #{sad <= v} >= m/2 + 1; SADs are < 121 * 510 < 2^16 */
    int lo = 0, hi = 65535;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        int c = 0;
        for (int i = tid; i < n; i += 256) { const int s = sad[fo + i]; c += s >= 0 && s <= mid; }
        if (tid == 0) s_cnt = 0;
        __syncthreads();
        atomicAdd(&s_cnt, c);
        __syncthreads();
        const int tot = s_cnt;
        __syncthreads();
        if (tot >= m / 2 + 1) hi = mid; else lo = mid + 1;
    }
    const float th = __fmul_rn(1.5f * 1.4f, (float)lo);
    int kept = 0;
    for (int i = tid; i < n; i += 256) {
        const int s = sad[fo + i];
        if (s < 0) continue;
        if ((float)s < th) ++kept;
        else { u_right[fo + i] = -1.f; depth[fo + i] = -1.f; }
    }
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    atomicAdd(&s_cnt, kept);
    __syncthreads();
    if (tid == 0) n_kept[frame] = s_cnt;
}

#endif
