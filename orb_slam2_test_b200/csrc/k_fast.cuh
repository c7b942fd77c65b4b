/*
 * k_fast.cuh -- K2: per-cell FAST-9/16 with the iniThFAST -> minThFAST fallback
 * (src/ORBextractor.cc:1000-1094; cv::FAST semantics per SURVEY.md A.3 / oracle/cvprim.c).
 *
 * One CTA per ~30x30 cell.  The reference runs cv::FAST twice on a cell that yields nothing at
 * the high threshold; here ONE score plane is computed at t_lo = min(ini, min):
 *   S(p) = (largest t for which p is still a FAST-9 corner), 0 for non-corners
 *   K(t) = { p : S(p) >= t and S(p) > S(q) for the 8 neighbours q inside the cell's domain }
 * and K(ini) = K(t_lo) restricted to S >= ini (a pixel that beats every neighbour's score beats
 * in particular those of the neighbours that are corners at the higher threshold), so the cell
 * emits K(ini) if that is non-empty, else K(min).  NMS never looks across the cell's domain
 * (3 px inside the window), exactly as a per-cell cv::FAST call does.
 *
 * Candidates are appended UNORDERED (warp-aggregated atomics).  The reference order
 * (cell row, cell col, y, x) only ever matters as the tie-break between equal responses inside
 * one quadtree node, and that key is recomputed from the coordinates in K3.
 */
#ifndef K_FAST_CUH
#define K_FAST_CUH

#include "borb_internal.h"

#define FAST_WPITCH 68 /* >= BORB_MAXCELL + 6, multiple of 4 */
#define FAST_SPITCH 64 /* >= BORB_MAXCELL + 2 */


/* 9 contiguous set bits in a 16-bit circular mask */
__host__ __device__ __forceinline__ bool has_arc9(uint32_t m)
{
    uint32_t x = m | (m << 16);
    x &= x >> 1;  /* runs of 2 */
    x &= x >> 2;  /* runs of 4 */
    x &= x >> 4;  /* runs of 8 */
    x &= x >> 1;  /* runs of 9 */
    return (x & 0xffffu) != 0;
}

/* FAST score of the pixel at p (row pitch `pitch`): 0 if it is not a FAST-9 corner at threshold
 * t_lo, else the largest threshold at which it still is one (>= t_lo). */
__host__ __device__ __forceinline__ int fast_score_px(const uint8_t *p, int pitch, int t_lo)
{
    const int v = p[0];
    int d[16];
    d[0] = v - p[3 * pitch];       d[8] = v - p[-3 * pitch];
    d[4] = v - p[3];               d[12] = v - p[-3];
    /* a 9-arc contains at least two of the four compass points, on the same side */
    {
        const int nb = (d[0] > t_lo) + (d[4] > t_lo) + (d[8] > t_lo) + (d[12] > t_lo);
        const int nd = (d[0] < -t_lo) + (d[4] < -t_lo) + (d[8] < -t_lo) + (d[12] < -t_lo);
        if (nb < 2 && nd < 2) return 0;
    }
    d[1] = v - p[3 * pitch + 1];   d[2] = v - p[2 * pitch + 2];
    d[3] = v - p[pitch + 3];       d[5] = v - p[-pitch + 3];
    d[6] = v - p[-2 * pitch + 2];  d[7] = v - p[-3 * pitch + 1];
    d[9] = v - p[-3 * pitch - 1];  d[10] = v - p[-2 * pitch - 2];
    d[11] = v - p[-pitch - 3];     d[13] = v - p[pitch - 3];
    d[14] = v - p[2 * pitch - 2];  d[15] = v - p[3 * pitch - 1];
    uint32_t mb = 0, md = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        mb |= (uint32_t)(d[k] > t_lo) << k;
        md |= (uint32_t)(d[k] < -t_lo) << k;
    }
    const bool bright = has_arc9(mb);
    if (!bright && !has_arc9(md)) return 0;
    /* OpenCV's cornerScore<16> equals max over the 16 arcs of min|d| minus 1, i.e. the largest
     * threshold at which the pixel still passes the segment test.  A bright 9-arc and a dark
     * 9-arc cannot coexist (18 > 16 samples), so only the side found above is searched.
     * Bisection over the threshold with the same bit-mask test; NOTE: a min/max-chain
     * formulation of this score is miscompiled by ptxas 12.9 -O1..-O3 for sm_100a (VIMNMX3
     * fusion; tools/dbg_fast.cu reproduces it), do not reintroduce it without that check. */
    int e[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) e[k] = bright ? d[k] : -d[k];
    int lo = t_lo, hi = 254;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        uint32_t m = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) m |= (uint32_t)(e[k] > mid) << k;
        if (has_arc9(m)) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(128)
k_fast_cells(const BorbPlan *__restrict__ plan, const BorbCell *__restrict__ cells,
             const uint8_t *__restrict__ pyr, uint32_t *__restrict__ cand_xy,
             uint32_t *__restrict__ cand_score, int *__restrict__ ncand)
{
    __shared__ uint8_t win[(BORB_MAXCELL + 6) * FAST_WPITCH];
    __shared__ uint8_t sc[(BORB_MAXCELL + 2) * FAST_SPITCH];

    const BorbCell cell = cells[blockIdx.x];
    const int frame = blockIdx.y;
    const BorbLevel &L = plan->lv[cell.level];
    const uint8_t *img = pyr + (size_t)frame * (size_t)plan->pyr_bytes + L.off +
                         (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    const int ww = cell.ww, wh = cell.wh;
    const int dw = ww - 6, dh = wh - 6; /* detection domain */
    const int tid = threadIdx.x;

    for (int i = tid; i < wh * ww; i += blockDim.x) {
        const int y = i / ww, x = i - y * ww;
        win[y * FAST_WPITCH + x] = img[(size_t)(cell.y0 + y) * L.pitch + cell.x0 + x];
    }
    for (int i = tid; i < (dh + 2) * FAST_SPITCH; i += blockDim.x) sc[i] = 0;
    __syncthreads();

    const int t_lo = min(plan->ini_th, plan->min_th);
    for (int i = tid; i < dw * dh; i += blockDim.x) {
        const int y = i / dw, x = i - y * dw;
        const int s = fast_score_px(win + (y + 3) * FAST_WPITCH + x + 3, FAST_WPITCH, t_lo);
        if (s == 0) continue;
        sc[(y + 1) * FAST_SPITCH + x + 1] = (uint8_t)s;
    }
    __syncthreads();

    /* strict 3x3 NMS inside the domain; survivors at the high threshold counted block-wide */
    const int ini = plan->ini_th, mn = plan->min_th;
    int n_hi = 0;
    for (int i0 = 0; i0 < dw * dh; i0 += blockDim.x) {
        const int i = i0 + tid;
        bool hi = false;
        if (i < dw * dh) {
            const int y = i / dw, x = i - y * dw;
            const uint8_t *q = sc + (y + 1) * FAST_SPITCH + x + 1;
            const int c = q[0];
            hi = c >= ini && c > q[-1] && c > q[1] && c > q[-FAST_SPITCH - 1] &&
                 c > q[-FAST_SPITCH] && c > q[-FAST_SPITCH + 1] && c > q[FAST_SPITCH - 1] &&
                 c > q[FAST_SPITCH] && c > q[FAST_SPITCH + 1];
        }
        n_hi += __syncthreads_count(hi);
    }
    const int t_emit = n_hi ? ini : mn;
    /* with ini < min the second cv::FAST pass can only return a subset of an empty set */
    if (!n_hi && mn > ini) return;

    uint32_t *oxy = cand_xy + (size_t)frame * (size_t)plan->cand_total + L.cand_off;
    uint32_t *osc = cand_score + (size_t)frame * (size_t)plan->cand_total + L.cand_off;
    int *cnt = ncand + frame * BORB_MAXL + cell.level;
    const int lane = tid & 31;
    for (int i0 = 0; i0 < dw * dh; i0 += blockDim.x) {
        const int i = i0 + tid;
        bool keep = false;
        int x = 0, y = 0, c = 0;
        if (i < dw * dh) {
            y = i / dw; x = i - y * dw;
            const uint8_t *q = sc + (y + 1) * FAST_SPITCH + x + 1;
            c = q[0];
            keep = c >= t_emit && c > q[-1] && c > q[1] && c > q[-FAST_SPITCH - 1] &&
                   c > q[-FAST_SPITCH] && c > q[-FAST_SPITCH + 1] && c > q[FAST_SPITCH - 1] &&
                   c > q[FAST_SPITCH] && c > q[FAST_SPITCH + 1];
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (bal) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cnt, __popc(bal));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) {
                const int slot = base + __popc(bal & ((1u << lane) - 1));
                if (slot < L.cand_cap) {
                    /* window-relative (x+3, y+3) shifted by the cell origin: coords relative
                     * to (minBorderX, minBorderY), src/ORBextractor.cc:1087-1088 */
                    oxy[slot] = (uint32_t)(x + 3 + cell.sx) | ((uint32_t)(y + 3 + cell.sy) << 16);
                    osc[slot] = (uint32_t)c;
                }
            }
        }
    }
}

#endif
