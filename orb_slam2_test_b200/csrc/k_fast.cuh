/*
 * k_fast.cuh -- K2: per-cell FAST-9/16 with the iniThFAST -> minThFAST fallback
 * (src/ORBextractor.cc:1000-1094; cv::FAST semantics per SURVEY.md A.3 / oracle/cvprim.c).
 *
 * The reference runs cv::FAST twice on a cell that yields nothing at the high threshold; here ONE
 * score plane is computed per cell at t_lo = min(ini, min):
 *   S(p) = (largest t for which p is still a FAST-9 corner), 0 for non-corners
 *   K(t) = { p : S(p) >= t and S(p) > S(q) for the 8 neighbours q inside the cell's domain }
 * and K(ini) = K(t_lo) restricted to S >= ini (a pixel that beats every neighbour's score beats
 * in particular those of the neighbours that are corners at the higher threshold), so the cell
 * emits K(ini) if that is non-empty, else K(min).  NMS never looks across the cell's domain
 * (3 px inside the window), exactly as a per-cell cv::FAST call does.
 *
 * Execution shape (v2).  Corner-ness is sparse and its cost is wildly uneven per pixel (a 5-load
 * compass test rejects ~3/4 of the pixels, scoring a corner costs ~400 instructions), so a
 * pixel-per-thread kernel spends its time in divergent warps (v1: 17 of 32 lanes active, ALU pipe
 * 87 % busy -- profiles/r1_fast_v1_raw.csv).  v2 is a persistent kernel, one CTA per cell at a
 * time, that runs the test as four DENSE stages with a shared-memory compaction between them:
 *   stage 0  window -> shared memory, aligned 32-bit loads
 *   stage 1  every pixel: compass pre-test (two ADJACENT compass points beyond the threshold on
 *            the same side -- necessary for a 9-arc), survivors appended to list1 with their side
 *   stage 2  list1: 16-point segment test on that side, corners appended to list2
 *   stage 3  list2: score -> score plane
 *   stage 4  list2: strict 3x3 NMS on the score plane, survivors to list3, count of those >= ini
 *   emit     one global atomicAdd per cell reserves the output slots
 * The score plane is kept zero between cells by clearing only the entries of list2.
 *
 * Candidates are appended UNORDERED.  The reference order (cell row, cell col, y, x) only ever
 * matters as the tie-break between equal responses inside one quadtree node, and that key is
 * recomputed from the coordinates in K3.
 */
#ifndef K_FAST_CUH
#define K_FAST_CUH

#include "borb_internal.h"

#define FAST_NT 128
#define FAST_WPITCH 72 /* >= BORB_MAXCELL + 6 + 3 (alignment slack), multiple of 4 */
#define FAST_SPITCH 64 /* >= BORB_MAXCELL + 2 */
#define FAST_MAXPIX (BORB_MAXCELL * BORB_MAXCELL)

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

/* ring offsets of the 16-point Bresenham circle (OpenCV order) */
#define FAST_RING(P, pitch, k)                                                                     \
    ((k) == 0 ? (P)[3 * (pitch)] : (k) == 1 ? (P)[3 * (pitch) + 1] : (k) == 2 ? (P)[2 * (pitch) + 2] \
     : (k) == 3 ? (P)[(pitch) + 3] : (k) == 4 ? (P)[3] : (k) == 5 ? (P)[-(pitch) + 3]               \
     : (k) == 6 ? (P)[-2 * (pitch) + 2] : (k) == 7 ? (P)[-3 * (pitch) + 1] : (k) == 8 ? (P)[-3 * (pitch)] \
     : (k) == 9 ? (P)[-3 * (pitch) - 1] : (k) == 10 ? (P)[-2 * (pitch) - 2] : (k) == 11 ? (P)[-(pitch) - 3] \
     : (k) == 12 ? (P)[-3] : (k) == 13 ? (P)[(pitch) - 3] : (k) == 14 ? (P)[2 * (pitch) - 2]         \
     : (P)[3 * (pitch) - 1])

/* Largest threshold t >= t_lo at which the 16 signed contrasts e[] (positive = beyond the centre
 * on the corner's side) still contain a 9-arc; the caller guarantees there is one at t_lo.
 * = OpenCV's cornerScore<16> for a corner (max over arcs of min contrast, minus 1).
 * Bisection over the threshold with the bit-mask arc test; NOTE: the min/max-chain formulation of
 * this score is miscompiled by ptxas 12.9 -O1..-O3 for sm_100a (3-input min/max fusion;
 * tools/dbg_fast.cu reproduces it), do not reintroduce it without that check. */
__host__ __device__ __forceinline__ int fast_score_from_contrasts(const int *e, int t_lo)
{
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

/* FAST score of the pixel at p (row pitch `pitch`): 0 if it is not a FAST-9 corner at threshold
 * t_lo, else the largest threshold at which it still is one (>= t_lo).  Reference form of the
 * staged kernel below (same arithmetic, one pixel at a time); used by the host tap
 * borb_debug_fast_score and tools/dbg_fast.cu. */
__host__ __device__ __forceinline__ int fast_score_px(const uint8_t *p, int pitch, int t_lo)
{
    const int v = p[0];
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = v - (int)FAST_RING(p, pitch, k);
    uint32_t mb = 0, md = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        mb |= (uint32_t)(d[k] > t_lo) << k;
        md |= (uint32_t)(d[k] < -t_lo) << k;
    }
    const bool bright = has_arc9(mb);
    /* a bright 9-arc and a dark 9-arc cannot coexist (18 > 16 samples) */
    if (!bright && !has_arc9(md)) return 0;
    int e[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) e[k] = bright ? d[k] : -d[k];
    return fast_score_from_contrasts(e, t_lo);
}

/* i / d and i % d for 0 <= i < 4096, 1 <= d <= 64, with m = fast_div_magic(d) */
__host__ __device__ __forceinline__ uint32_t fast_div_magic(int d) { return (1u << 20) / (uint32_t)d + 1u; }
__host__ __device__ __forceinline__ int fast_div(int i, uint32_t m) { return (int)(((uint32_t)i * m) >> 20); }

/* Shared memory of one CTA, sized on the host for the largest cell of the plan (K2k: 33x40 px ->
 * 15 KB, so 14 CTAs fit on an SM; the 60x60 worst case of the cell formula needs 36 KB). */
struct FastSmemLayout {
    int win_off, sc_off, l1_off, l2_off, l3_off, bytes;
};
__host__ __device__ __forceinline__ FastSmemLayout fast_smem_layout(int max_wh, int max_pix)
{
    FastSmemLayout o;
    int b = 64; /* counters */
    o.win_off = b; b += max_wh * FAST_WPITCH;
    o.sc_off = b;  b += ((max_wh - 4) * FAST_SPITCH + 15) & ~15;
    o.l1_off = b;  b += (2 * max_pix * 2 + 15) & ~15;  /* u16 pixel | side<<12; one slot per side */
    o.l2_off = b;  b += (max_pix * 2 + 15) & ~15;      /* u16 corners */
    o.l3_off = b;  b += (max_pix / 4 + 64) * 4;        /* u32 NMS survivors: pixel | score<<12 */
    o.bytes = b;
    return o;
}
struct FastCounters {
    int n1, n2, n3, n_hi, out_base, out_fill;
};

__global__ void __launch_bounds__(FAST_NT)
k_fast_cells(const BorbPlan *__restrict__ plan, const BorbCell *__restrict__ cells, int n_frames,
             const uint8_t *__restrict__ pyr, uint32_t *__restrict__ cand_xy,
             uint32_t *__restrict__ cand_score, int *__restrict__ ncand)
{
    extern __shared__ __align__(16) unsigned char fast_raw[];
    const FastSmemLayout lay = fast_smem_layout(plan->fast_max_wh, plan->fast_max_pix);
    struct {
        FastCounters &c;
        uint8_t *win, *sc;
        uint16_t *list1, *list2;
        uint32_t *list3;
    } s = {*reinterpret_cast<FastCounters *>(fast_raw), fast_raw + lay.win_off, fast_raw + lay.sc_off,
           reinterpret_cast<uint16_t *>(fast_raw + lay.l1_off), reinterpret_cast<uint16_t *>(fast_raw + lay.l2_off),
           reinterpret_cast<uint32_t *>(fast_raw + lay.l3_off)};
    const int tid = threadIdx.x, lane = tid & 31;
    const int n_cells = plan->n_cells;
    const int total = n_cells * n_frames;
    const int ini = plan->ini_th, mn = plan->min_th;
    const int t_lo = min(ini, mn);

    for (int i = tid; i < (lay.l1_off - lay.sc_off) / 4; i += FAST_NT) reinterpret_cast<uint32_t *>(s.sc)[i] = 0u;
    if (tid == 0) { s.c.n1 = 0; s.c.n2 = 0; s.c.n3 = 0; s.c.n_hi = 0; s.c.out_fill = 0; }
    __syncthreads();

    for (int work = blockIdx.x; work < total; work += gridDim.x) {
        const int frame = work / n_cells;
        const BorbCell cell = cells[work - frame * n_cells];
        const BorbLevel &L = plan->lv[cell.level];
        const uint8_t *img = pyr + (size_t)frame * (size_t)plan->pyr_bytes + L.off +
                             (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
        const int ww = cell.ww, wh = cell.wh;
        const int dw = ww - 6, dh = wh - 6; /* detection domain */
        const int npix = dw * dh;
        const uint32_t dmagic = fast_div_magic(dw);

        /* ---- stage 0: window -> smem, whole aligned words (level rows are 32-B aligned) ---- */
        const int ax0 = cell.x0 & ~3, xoff = cell.x0 - ax0;
        {
            const int wpr = (xoff + ww + 3) >> 2; /* words per row, <= 18 */
            const uint32_t wmagic = fast_div_magic(wpr);
            const int nw = wpr * wh;
            for (int i = tid; i < nw; i += FAST_NT) {
                const int y = fast_div(i, wmagic), w = i - y * wpr;
                const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(img + (size_t)(cell.y0 + y) * L.pitch + ax0) + w);
                reinterpret_cast<uint32_t *>(s.win + y * FAST_WPITCH)[w] = v;
            }
        }
        __syncthreads();

        /* ---- stage 1: compass pre-test on every pixel of the domain ---- */
        for (int i0 = 0; i0 < npix; i0 += FAST_NT) {
            const int i = i0 + tid;
            uint32_t side = 0;
            int y = 0, x = 0;
            if (i < npix) {
                y = fast_div(i, dmagic);
                x = i - y * dw;
                const uint8_t *p = s.win + (y + 3) * FAST_WPITCH + xoff + x + 3;
                const int v = p[0];
                const int lo = v - t_lo, hi = v + t_lo;
                const int r0 = p[3 * FAST_WPITCH], r4 = p[3], r8 = p[-3 * FAST_WPITCH], r12 = p[-3];
                const bool b0 = r0 < lo, b4 = r4 < lo, b8 = r8 < lo, b12 = r12 < lo;
                const bool k0 = r0 > hi, k4 = r4 > hi, k8 = r8 > hi, k12 = r12 > hi;
                /* a 9-arc covers two adjacent compass points */
                if ((b0 && (b4 || b12)) || (b8 && (b4 || b12))) side |= 1u;
                if ((k0 && (k4 || k12)) || (k8 && (k4 || k12))) side |= 2u;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, side != 0);
            if (bal) {
                /* both sides flagged is possible but rare: such a pixel takes two slots */
                const uint32_t both = __ballot_sync(0xffffffffu, side == 3u);
                int base = 0;
                if (lane == 0) base = atomicAdd(&s.c.n1, __popc(bal) + __popc(both));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (side) {
                    const uint32_t below = (1u << lane) - 1u;
                    int slot = base + __popc(bal & below) + __popc(both & below);
                    const uint32_t pix = (uint32_t)(y * 64 + x);
                    if (side & 1u) s.list1[slot++] = (uint16_t)(pix | (1u << 12));
                    if (side & 2u) s.list1[slot] = (uint16_t)(pix | (2u << 12));
                }
            }
        }
        __syncthreads();

        /* ---- stage 2: full segment test of the listed (pixel, side) pairs ---- */
        const int n1 = s.c.n1;
        for (int i0 = 0; i0 < n1; i0 += FAST_NT) {
            const int i = i0 + tid;
            bool corner = false;
            uint32_t ent = 0;
            if (i < n1) {
                ent = s.list1[i];
                const int x = ent & 63, y = (ent >> 6) & 63;
                const uint8_t *p = s.win + (y + 3) * FAST_WPITCH + xoff + x + 3;
                const int v = p[0];
                /* contrast beyond the threshold on this side: g = sgn*r + off > 0 */
                const int sgn = (ent & (1u << 12)) ? -1 : 1;
                const int off = (ent & (1u << 12)) ? v - t_lo : -(v + t_lo);
                uint32_t m = 0;
#pragma unroll
                for (int k = 0; k < 16; ++k) m |= (uint32_t)(sgn * (int)FAST_RING(p, FAST_WPITCH, k) + off > 0) << k;
                corner = has_arc9(m);
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, corner);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&s.c.n2, __popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (corner) s.list2[base + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)ent;
            }
        }
        __syncthreads();

        /* ---- stage 3: score of every corner -> score plane ---- */
        const int n2 = s.c.n2;
        for (int i = tid; i < n2; i += FAST_NT) {
            const uint32_t ent = s.list2[i];
            const int x = ent & 63, y = (ent >> 6) & 63;
            const uint8_t *p = s.win + (y + 3) * FAST_WPITCH + xoff + x + 3;
            const int v = p[0];
            const int sgn = (ent & (1u << 12)) ? -1 : 1;
            const int off = (ent & (1u << 12)) ? v : -v;
            int e[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) e[k] = sgn * (int)FAST_RING(p, FAST_WPITCH, k) + off;
            s.sc[(y + 1) * FAST_SPITCH + x + 1] = (uint8_t)fast_score_from_contrasts(e, t_lo);
        }
        __syncthreads();

        /* ---- stage 4: strict 3x3 NMS inside the domain ---- */
        for (int i0 = 0; i0 < n2; i0 += FAST_NT) {
            const int i = i0 + tid;
            bool keep = false;
            uint32_t rec = 0;
            if (i < n2) {
                const uint32_t ent = s.list2[i];
                const int x = ent & 63, y = (ent >> 6) & 63;
                const uint8_t *q = s.sc + (y + 1) * FAST_SPITCH + x + 1;
                const int c = q[0];
                keep = c > q[-1] && c > q[1] && c > q[-FAST_SPITCH - 1] && c > q[-FAST_SPITCH] &&
                       c > q[-FAST_SPITCH + 1] && c > q[FAST_SPITCH - 1] && c > q[FAST_SPITCH] &&
                       c > q[FAST_SPITCH + 1];
                rec = (ent & 0xfffu) | ((uint32_t)c << 12);
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (bal) {
                const uint32_t bhi = __ballot_sync(0xffffffffu, keep && (int)(rec >> 12) >= ini);
                int base = 0;
                if (lane == 0) {
                    base = atomicAdd(&s.c.n3, __popc(bal));
                    if (bhi) atomicAdd(&s.c.n_hi, __popc(bhi));
                }
                base = __shfl_sync(0xffffffffu, base, 0);
                if (keep) s.list3[base + __popc(bal & ((1u << lane) - 1u))] = rec;
            }
        }
        __syncthreads();

        /* ---- emit: K(ini) if non-empty else K(min).  (With ini < min every survivor already is
         * >= ini, so n_hi == n3: a cell empty at ini stays empty, its second pass being a subset.) ---- */
        const int n3 = s.c.n3, n_hi = s.c.n_hi;
        const int t_emit = n_hi ? ini : mn;
        const int n_emit = n_hi ? n_hi : n3;
        if (n_emit > 0) {
            int *cnt = ncand + frame * BORB_MAXL + cell.level;
            if (tid == 0) s.c.out_base = atomicAdd(cnt, n_emit);
            __syncthreads();
            const int base = s.c.out_base;
            uint32_t *oxy = cand_xy + (size_t)frame * (size_t)plan->cand_total + L.cand_off;
            uint32_t *osc = cand_score + (size_t)frame * (size_t)plan->cand_total + L.cand_off;
            for (int i = tid; i < n3; i += FAST_NT) {
                const uint32_t rec = s.list3[i];
                const int c = (int)(rec >> 12);
                if (c >= t_emit) {
                    const int slot = base + atomicAdd(&s.c.out_fill, 1);
                    if (slot < L.cand_cap) {
                        const int x = rec & 63, y = (rec >> 6) & 63;
                        /* window-relative (x+3, y+3) shifted by the cell origin: coords relative
                         * to (minBorderX, minBorderY), src/ORBextractor.cc:1087-1088 */
                        oxy[slot] = (uint32_t)(x + 3 + cell.sx) | ((uint32_t)(y + 3 + cell.sy) << 16);
                        osc[slot] = (uint32_t)c;
                    }
                }
            }
        }
        /* ---- restore the all-zero score plane and the counters for the next cell ---- */
        for (int i = tid; i < n2; i += FAST_NT) {
            const uint32_t ent = s.list2[i];
            s.sc[(((ent >> 6) & 63) + 1) * FAST_SPITCH + (ent & 63) + 1] = 0;
        }
        __syncthreads();
        if (tid == 0) { s.c.n1 = 0; s.c.n2 = 0; s.c.n3 = 0; s.c.n_hi = 0; s.c.out_fill = 0; }
        /* the next iteration's first write to these is after its stage-0 barrier */
    }
}

#endif
