/*
 * k_stereo.cuh -- S1/S2: Frame::ComputeStereoMatches on the device (src/Frame.cc:619-834), the
 * immediate consumer of BOTH extractors' keypoints, descriptors and pyramids.
 * Running it where those already live removes the two pyramid downloads from the stereo path.
 *
 *   S0  one CTA per pair: the reference's per-row candidate tables vRowIndices (:627-645) in CSR
 *       form -- right keypoint iR is listed under rows floor(yR - 2s) .. ceil(yR + 2s), s its level's
 *       scale factor (shared-memory histogram, scan, scatter; order inside a row is irrelevant, see
 *       S1).
 *   S1  one warp per left keypoint.  (a) The candidates of row (size_t)yL, 32 at a time: octave
 *       within +-1 and x inside [uL - maxD, uL]; 256-bit Hamming distance by popc
 *       (= ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1846-1862); the minimum of
 *       (distance << 16 | index) over the row = the first smallest distance in right-index order,
 *       which is what the reference's strict `<` over its index-ordered list keeps.  (b) If it is below (TH_HIGH+TH_LOW)/2, 11 shifted
 *       11x11 SADs on the keypoint's pyramid level (centre-subtracted, :697-740), parabola fit,
 *       disparity and depth with every float operation individually rounded (:742-781).
 *   S2  one CTA per frame pair: the median SAD (element size/2 of the sorted list, :786-789) by a
 *       counting bisection, then the 1.5*1.4*median outlier cut (:792-803).
 */
#ifndef K_STEREO_CUH
#define K_STEREO_CUH

#include "borb_internal.h"
#include "k_describe.cuh"
#include "borb_cta.cuh"

#define ST_NT 128

#define ST_ROWS_NT 1024
#define ST_CUT_NT 1024

/* S0: row_start[rows + 1], row_ent[kp_total * span] per frame; span = the most rows one right
 * keypoint can be listed under = floor(4 * largest scale factor) + 3 (computed on the host) */
__global__ void __launch_bounds__(ST_ROWS_NT)
k_stereo_rows(const BorbPlan *__restrict__ plan, const BorbKpRec *__restrict__ kp_r, const int *__restrict__ n_r,
              int span, int *__restrict__ row_start, uint16_t *__restrict__ row_ent)
{
    extern __shared__ int st_sm[]; /* cnt[rows], cur[rows], part[ST_ROWS_NT] */
    const int rows = plan->lv[0].h, kp_total = plan->kp_total;
    int *cnt = st_sm, *cur = st_sm + rows, *part = st_sm + 2 * rows;
    const int frame = blockIdx.x, tid = threadIdx.x;
    const size_t fo = (size_t)frame * kp_total;
    const int n = n_r[frame];
    for (int y = tid; y < rows; y += ST_ROWS_NT) cnt[y] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += ST_ROWS_NT) {
        const BorbKpRec &q = kp_r[fo + i];
        const float r = __fmul_rn(2.0f, plan->lv[q.octave].scale);
        const int maxr = min((int)ceilf(__fadd_rn(q.y, r)), rows - 1), minr = max((int)floorf(__fsub_rn(q.y, r)), 0);
        for (int y = minr; y <= maxr; ++y) atomicAdd(&cnt[y], 1);
    }
    __syncthreads();
    const int per = (rows + ST_ROWS_NT - 1) / ST_ROWS_NT;
    int sum = 0;
    for (int c = 0; c < per; ++c) { const int y = tid * per + c; if (y < rows) sum += cnt[y]; }
    int total;
    int base = cta_exscan<ST_ROWS_NT>(sum, part, &total);
    int *start = row_start + (size_t)frame * (rows + 1);
    for (int c = 0; c < per; ++c) {
        const int y = tid * per + c;
        if (y < rows) { start[y] = base; cur[y] = base; base += cnt[y]; }
    }
    if (tid == 0) start[rows] = total;
    __syncthreads();
    uint16_t *ent = row_ent + fo * span; /* total <= n * span by construction */
    for (int i = tid; i < n; i += ST_ROWS_NT) {
        const BorbKpRec &q = kp_r[fo + i];
        const float r = __fmul_rn(2.0f, plan->lv[q.octave].scale);
        const int maxr = min((int)ceilf(__fadd_rn(q.y, r)), rows - 1), minr = max((int)floorf(__fsub_rn(q.y, r)), 0);
        for (int y = minr; y <= maxr; ++y) ent[atomicAdd(&cur[y], 1)] = (uint16_t)i;
    }
}

__global__ void __launch_bounds__(ST_NT)
k_stereo_match(const BorbPlan *__restrict__ plan, const BorbKpRec *__restrict__ kp_l,
               const uint8_t *__restrict__ desc_l, const int *__restrict__ n_l,
               const BorbKpRec *__restrict__ kp_r, const uint8_t *__restrict__ desc_r,
               const int *__restrict__ n_r, int span, const int *__restrict__ row_start,
               const uint16_t *__restrict__ row_ent, const uint8_t *__restrict__ pyr_l,
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
        const int rows = plan->lv[0].h;
        const int *start = row_start + (size_t)frame * (rows + 1);
        const uint16_t *ent = row_ent + fo * span;
        const int s0 = start[row], s1 = start[row + 1];
        uint32_t dl[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) dl[k] = reinterpret_cast<const uint32_t *>(desc_l + (fo + o) * 32)[k];
        for (int i = s0 + lane; i < s1; i += 32) {
            const int ir = ent[i];
            const BorbKpRec &q = kp_r[fo + ir];
            const int oc = q.octave;
            if (oc < lv - 1 || oc > lv + 1) continue;
            const float ur = q.x;
            if (!(ur >= min_u && ur <= max_u)) continue;
            const uint4 *dr = reinterpret_cast<const uint4 *>(desc_r + (fo + ir) * 32);
            const uint4 a = __ldg(dr), b = __ldg(dr + 1);
            const int d = __popc(dl[0] ^ a.x) + __popc(dl[1] ^ a.y) + __popc(dl[2] ^ a.z) + __popc(dl[3] ^ a.w) +
                          __popc(dl[4] ^ b.x) + __popc(dl[5] ^ b.y) + __popc(dl[6] ^ b.z) + __popc(dl[7] ^ b.w);
            const uint32_t key = ((uint32_t)d << 16) | (uint32_t)ir;
            if (key < best) best = key; /* (distance, index) lexicographic; best starts at (TH_HIGH, 0xffff) */
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
            int sad[11], cr_v[11];
#pragma unroll
            for (int i = 0; i < 11; ++i) { sad[i] = 0; cr_v[i] = (int)pr[i - LL]; }
            for (int p = lane; p < 121; p += 32) {
                const int dy = p / 11 - w, dx = p - (dy + w) * 11 - w;
                const int a = (int)pl[dy * L.pitch + dx] - cl;
                const uint8_t *q = pr + dy * L.pitch + dx;
#pragma unroll
                for (int i = 0; i < 11; ++i) sad[i] += abs(a - ((int)q[i - LL] - cr_v[i]));
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

__global__ void __launch_bounds__(ST_CUT_NT)
k_stereo_cut(const BorbPlan *__restrict__ plan, const int *__restrict__ n_l, float *__restrict__ u_right,
             float *__restrict__ depth, const int *__restrict__ sad, int *__restrict__ n_kept)
{
    /* median = element m/2 (0-based) of the sorted matched SADs = smallest v with #{sad <= v} >= m/2+1.
     * SADs are < 121 * 510 < 2^16: a histogram of the high byte finds its bin, a histogram of the
     * low byte inside that bin finds the value. */
    __shared__ int hist[256];
    __shared__ int sel[2]; /* chosen bin, count below it */
    __shared__ int s_cnt;
    const int frame = blockIdx.x, tid = threadIdx.x;
    const size_t fo = (size_t)frame * plan->kp_total;
    const int n = n_l[frame];
    if (tid < 256) hist[tid] = 0;
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    for (int i = tid; i < n; i += ST_CUT_NT) {
        const int s = sad[fo + i];
        if (s >= 0) atomicAdd(&hist[s >> 8], 1);
    }
    __syncthreads();
    if (tid < 32) { /* one warp walks the 256 bins */
        int c[8], sum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { c[k] = hist[tid * 8 + k]; sum += c[k]; }
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (tid >= o) inc += t;
        }
        const int m = __shfl_sync(0xffffffffu, inc, 31);
        if (tid == 0) s_cnt = m;
        const int need = m / 2 + 1;
        int run = inc - sum;
        if (m > 0 && run < need && inc >= need) { /* exactly one lane */
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (run < need && run + c[k] >= need) { sel[0] = tid * 8 + k; sel[1] = run; }
                run += c[k];
            }
        }
    }
    __syncthreads();
    const int m = s_cnt;
    if (m == 0) { if (tid == 0) n_kept[frame] = 0; return; }
    const int bin = sel[0], below = sel[1];
    __syncthreads();
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += ST_CUT_NT) {
        const int s = sad[fo + i];
        if (s >= 0 && (s >> 8) == bin) atomicAdd(&hist[s & 255], 1);
    }
    __syncthreads();
    if (tid < 32) {
        int c[8], sum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { c[k] = hist[tid * 8 + k]; sum += c[k]; }
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (tid >= o) inc += t;
        }
        const int need = m / 2 + 1 - below;
        int run = inc - sum;
        if (run < need && inc >= need) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (run < need && run + c[k] >= need) sel[0] = (bin << 8) | (tid * 8 + k);
                run += c[k];
            }
        }
    }
    __syncthreads();
    const float th = __fmul_rn(1.5f * 1.4f, (float)sel[0]);
    int kept = 0;
    for (int i = tid; i < n; i += ST_CUT_NT) {
        const int s = sad[fo + i];
        if (s < 0) continue;
        if ((float)s < th) ++kept;
        else { u_right[fo + i] = -1.f; depth[fo + i] = -1.f; }
    }
    kept = __reduce_add_sync(0xffffffffu, kept);
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    if ((tid & 31) == 0 && kept) atomicAdd(&s_cnt, kept);
    __syncthreads();
    if (tid == 0) n_kept[frame] = s_cnt;
}

#endif
