/*
 * k_fast.cuh -- K2: per-cell FAST-9/16 with the iniThFAST -> minThFAST fallback
 * (src/ORBextractor.cc:1000-1094; cv::FAST semantics per SURVEY.md A.3 / oracle/cvprim.c).
 *
 * With S(p) = the largest t for which p is still a FAST-9 corner (OpenCV's cornerScore, 0 for
 * non-corners), cv::FAST(t, nms) on a cell returns
 *   K(t) = { p : S(p) >= t and S(p) > S_t(q) for the 8 neighbours q inside the cell's domain },
 * S_t(q) = S(q) if S(q) >= t else 0.  The reference keeps K(ini) if it is non-empty, else K(min).
 * NMS never looks across the cell's domain (3 px inside the window), exactly as a per-cell
 * cv::FAST call does.
 *
 * Execution shape (v3).  Corner-ness is sparse and its cost is wildly uneven per pixel (a 5-load
 * compass test rejects over half of the pixels, the full test + score costs ~150 instructions), so
 * a pixel-per-thread kernel spends its time in divergent warps (v1: 17 of 32 lanes active, ALU pipe
 * 87 % busy -- profiles/r1_fast_v1_raw.csv).  This is a persistent kernel, one CTA (2 warps) per
 * cell at a time; 32-pixel groups of the cell are dealt round-robin to the warps, each of which runs DENSE stages
 * over warp-private shared-memory lists (ballot/popc compaction, no atomics):
 *   stage 0  window -> shared memory, aligned 32-bit loads                       | barrier
 *   stage 1  own pixels: compass pre-test (two ADJACENT compass points beyond the threshold on
 *            the same side -- necessary for a 9-arc); (pixel, side) appended to list1
 *   stage 2  list1: the 16 contrasts on that side, M = max over the 16 arcs of the minimum of
 *            their 9 contrasts (sliding-window minimum: 4 min stages + 1 max tree); the pixel is
 *            a corner at t iff M > t and its score is M - 1 (= OpenCV cornerScore<16>);
 *            score -> score plane, pixel -> list2                               | barrier
 *   stage 3  list2: strict 3x3 NMS on the score plane, survivors to list3, per-warp counts
 *                                                                                | barrier
 *   stages 1-3 run at t = ini; only a cell that ends with no survivor repeats them at t = min
 *   emit     one global atomicAdd per warp
 * The score plane is kept zero between cells by clearing only the entries of list2.
 *
 * Candidates are appended UNORDERED.  The reference order (cell row, cell col, y, x) only ever
 * matters as the tie-break between equal responses inside one quadtree node, and that key is
 * recomputed from the coordinates in K3.
 */
#ifndef K_FAST_CUH
#define K_FAST_CUH

#include <cuda.h> /* CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint) */

#include "borb_internal.h"
#include "borb_cta.cuh"


/* staged window row pitch WP (template parameter of the kernel) = width of the TMA box.  A box must
 * start on a 16-byte boundary of the plane row (measured: any other start coordinate faults), so the
 * detection domain begins at byte 4..19 of a staged row; WP is the smallest of 64 / 80 / 96 that
 * holds that offset, the widest domain of the plan and the word stage 1 reads past it */
#define FAST_WP_FOR(max_ww) ((max_ww) + 20 <= 64 ? 64 : (max_ww) + 20 <= 80 ? 80 : 96)
/* byte of a staged row at which the detection domain of a window with origin x0 begins */
#define FAST_DOMAIN_BYTE(x0) (4 + (((x0) - 1) & 15))
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
 * Bisection over the threshold with the bit-mask arc test: exact by definition of the score.  This
 * is the REFERENCE form (host tap, tools/dbg_*.cu); the kernel uses fast_arc_minmax below.
 * NOTE: OpenCV's interleaved min/max chain for this score is miscompiled by ptxas 12.9 -O1..-O3 for
 * sm_100a (tools/dbg_fast.cu reproduces it); the sliding-window form below is not
 * (tools/dbg_score.cu: 0 mismatches vs this bisection on 0.9 M corners, int32 / float / s16x2). */
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

/* max over the 16 circular 9-arcs of the minimum contrast in the arc */
__device__ __forceinline__ int fast_arc_minmax(const int *e)
{
    int m2[16], m4[16], m8[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m8[k] = min(m4[k], m4[(k + 4) & 15]);
    int M = -65536;
#pragma unroll
    for (int k = 0; k < 16; ++k) M = max(M, min(m8[k], e[(k + 8) & 15]));
    return M;
}

/* The same on 16-bit pairs with three-input packed min/max (VIMNMX3.S16x2): P[j] holds the (biased, positive)
 * contrasts of ring positions j and j + 8 in its two halves, so one packed operation advances two arc
 * positions; the wrap-around of the circular windows is a half-word swap.  A 9-window is three 3-windows:
 *   A[j] = min3(X[j], X[j+1], X[j+2]),  B[j] = min3(A[j], A[j+3], A[j+6]),  M = max over B
 * on the 16-long circular sequence X[j] = P[j] (j < 8), X[j+8] = swap(X[j]) -- 16 min3 + 5 max + 9 swaps.
 * (A sliding-window form on two-input VIMNMX.S16x2, 40 min/max + 15 swaps, was 8 % slower: 2.24 -> 2.07 ms
 * per 512 K2k frames; the kernel is bound by the ALU pipe these run on.)  Inputs in [0, 32767]; returns the
 * max over the 16 circular 9-arcs of the arc minimum (same bias as the inputs) in BOTH halves of the word. */
__device__ __forceinline__ uint32_t fast_swap16(uint32_t v) { return __byte_perm(v, 0u, 0x1032); }
__device__ __forceinline__ uint32_t fast_arc_minmax3_x2(const uint32_t *P)
{
    uint32_t X[10], A[14], B[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) X[j] = P[j];
    X[8] = fast_swap16(P[0]);
    X[9] = fast_swap16(P[1]);
#pragma unroll
    for (int j = 0; j < 8; ++j) A[j] = __vimin3_s16x2(X[j], X[j + 1], X[j + 2]);
#pragma unroll
    for (int j = 8; j < 14; ++j) A[j] = fast_swap16(A[j - 8]);
#pragma unroll
    for (int j = 0; j < 8; ++j) B[j] = __vimin3_s16x2(A[j], A[j + 3], A[j + 6]);
    const uint32_t r0 = __vimax3_s16x2(B[0], B[1], B[2]), r1 = __vimax3_s16x2(B[3], B[4], B[5]);
    const uint32_t M = __vimax3_s16x2(B[6], B[7], r0);
    return __vimax3_s16x2(M, r1, fast_swap16(__vmaxs2(M, r1)));
}

/* Shared memory of one CTA, sized on the host for the largest cell of the plan (K2k: 33x40 px ->
 * 18 KB; the 60x60 worst case of the cell formula needs 44 KB).  chunk = pixels per warp. */
struct FastSmemLayout {
    int chunk, win_off, win_bytes, sc_off, l1_off, l2_off, l3_off, bytes;
};
__host__ __device__ __forceinline__ int fast_chunk(int npix, int nw) { return ((npix + 32 * nw - 1) / (32 * nw)) * 32; }
__host__ __device__ __forceinline__ FastSmemLayout fast_smem_layout(int max_wh, int max_pix, int wp, int nw)
{
    FastSmemLayout o;
    o.chunk = fast_chunk(max_pix, nw) + 64; /* a warp owns whole rows: up to one row (< 64 px) above its share */
    int b = 128; /* per-warp counts [0,64), the two window mbarriers [64,80) */
    o.win_off = b; o.win_bytes = (max_wh * wp + 127) & ~127; b += 2 * o.win_bytes; /* double buffered, 128-B aligned (TMA) */
    o.sc_off = b;  b += ((max_wh - 4) * FAST_SPITCH + 15) & ~15;
    /* list1 (u16 pixel | side<<12) is dead once stage 2 has consumed it; list3 (u32 NMS survivors:
     * pixel | score<<12, at most half a warp's pixels + one per row) reuses its storage */
    o.l1_off = b;  o.l3_off = b;  b += nw * (o.chunk + 64) * 2;
    o.l2_off = b;  b += nw * o.chunk * 2;            /* u16 corners */
    o.bytes = b;
    return o;
}

/* One cell's window into shared memory by TMA: a single thread arms the buffer's mbarrier with the
 * box size and issues one cp.async.bulk.tensor (box WP bytes x max_wh rows x 1 frame of the level's
 * padded plane; rows / columns of the box past the plane are zero-filled and never read).  The box
 * starts at the 16-byte boundary at or before the column left of the window. */
__device__ __forceinline__ uint32_t fast_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct FastMaps { /* kernel parameter (__grid_constant__): one TMA descriptor per pyramid level, and the
                     plan constants every CTA needs at once (a parameter read, not a global load) */
    CUtensorMap m[BORB_MAXL];
    int max_wh, max_pix, n_cells, ini_th, min_th;
};

/* box origin of a cell's window in its level's padded plane, packed x | y << 16 | level << 28 (x, y < 4096 + pad) */
__device__ __forceinline__ uint32_t fast_tma_origin(const BorbCell &cell)
{
    const int x = BORB_XPAD + ((cell.x0 - 1) & ~15), y = BORB_EDGE + cell.y0; /* BORB_XPAD is a multiple of 16 */
    return (uint32_t)x | ((uint32_t)y << 16);
}
__device__ __forceinline__ void fast_tma_window(const CUtensorMap *maps, uint32_t origin, int level, int frame, uint32_t dst,
                                                uint32_t bar, uint32_t bytes)
{
    const int x = (int)(origin & 0xffffu), y = (int)(origin >> 16);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(maps + level), "r"(x), "r"(y), "r"(frame), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void fast_mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tFAST_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra FAST_DONE;\n\tbra FAST_WAIT;\n\tFAST_DONE:\n\t}" ::"r"(bar), "r"(parity)
                 : "memory");
}

#ifndef FAST_MIN_CTAS
#define FAST_MIN_CTAS 16
#endif
/* NW = warps per cell: 2 for batches (measured best of 2 / 4 / 8: load balance between uneven cells), 4 for a frame or
 * two, where the grid is one CTA per cell and a cell's dependent chain (load -> pre-test -> score -> NMS) is the time */
/* -DBORB_FAST_PROBE (tools/probe_fast.py, never in the shipped library): %globaltimer of thread 0 at eight points of
 * a CTA's first item, for the phase split of the single-frame shape */
#ifdef BORB_FAST_PROBE
__device__ unsigned long long g_fast_probe[2048 * 8];
#define FAST_PROBE(k)                                                                      \
    do {                                                                                   \
        if (threadIdx.x == 0 && blockIdx.y == 0 && blockIdx.x < 2048) {                    \
            unsigned long long t_;                                                         \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                         \
            g_fast_probe[blockIdx.x * 8 + (k)] = t_;                                       \
        }                                                                                  \
    } while (0)
#else
#define FAST_PROBE(k) do { } while (0)
#endif

template <int WP, int NW>
__global__ void __launch_bounds__(32 * NW, FAST_MIN_CTAS * 2 / NW)
k_fast_cells(const __grid_constant__ BorbPlan plan_c, const BorbCell *__restrict__ cells, int n_frames,
             const __grid_constant__ FastMaps tmaps, int frame0, uint32_t *__restrict__ cand_xy,
             uint32_t *__restrict__ cand_score, int *__restrict__ ncand)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    extern __shared__ __align__(128) unsigned char fast_raw[];
    constexpr int FAST_NT = 32 * NW, FAST_NW = NW;
    const FastSmemLayout lay = fast_smem_layout(tmaps.max_wh, tmaps.max_pix, WP, NW);
    const uint32_t bar0 = fast_smem_u32(fast_raw + 64), win0 = fast_smem_u32(fast_raw + lay.win_off);
    const uint32_t box_bytes = (uint32_t)(WP * tmaps.max_wh);
    const CUtensorMap *maps = tmaps.m;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t below = (1u << lane) - 1u;
    int *wcnt = reinterpret_cast<int *>(fast_raw); /* [pass][warp]: NMS survivors */
    uint8_t *sc = fast_raw + lay.sc_off;
    uint16_t *list1 = reinterpret_cast<uint16_t *>(fast_raw + lay.l1_off) + warp * (lay.chunk + 64);
    uint16_t *list2 = reinterpret_cast<uint16_t *>(fast_raw + lay.l2_off) + warp * lay.chunk;
    uint32_t *list3 = reinterpret_cast<uint32_t *>(list1);
    const int n_cells = tmaps.n_cells;
    const int ini = tmaps.ini_th, mn = tmaps.min_th;
    FAST_PROBE(0);

    for (int i = tid; i < (lay.l1_off - lay.sc_off) / 4; i += FAST_NT) reinterpret_cast<uint32_t *>(sc)[i] = 0u;

    /* persistent walk over (cell, frame) items, FRAME-MAJOR per CTA: a CTA keeps its cell and walks the frames
     * blockIdx.y, blockIdx.y + gridDim.y, ... before it moves to cell ci + gridDim.x (the host sizes gridDim.x =
     * n_cells whenever it can, so in the batched shape a CTA never changes cell).  Everything that depends only on
     * the cell -- the stage-1 lane geometry, the level's candidate slab -- is computed when the cell changes, not
     * per item (it was 12.5 % of the kernel's instructions when the walk was cell-major).  The window of the NEXT
     * item is copied into the other buffer while this one is processed. */
    int frame = blockIdx.y, ci = blockIdx.x, buf = 0;
    bool have = frame < n_frames && ci < n_cells;
    uint32_t origin = 0; /* of the current cell's window (fast_tma_origin) */
    int level = 0;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (have) {
        const BorbCell c0 = cells[ci];
        origin = fast_tma_origin(c0);
        level = c0.level;
    }
    borb_pdl_wait(); /* the pyramid is complete from here on (nothing above reads it) */
    borb_pdl_launch_dependents();
    if (have && tid == 0) fast_tma_window(maps, origin, level, frame0 + frame, win0, bar0, box_bytes);
    FAST_PROBE(1);
    /* per-cell state (see the geometry block below) */
    int dw = 0, dh = 0, s1_R = 1, s1_rig = 0, s1_col = 0, org_off = 0, s1_qoff = 0, cand_cap = 0, sxy3 = 0;
    bool s1_on = false;
    uint32_t s1_vlo = 0, s1_vhi = 0;
    long long cand_off = 0;
    bool new_cell = true;
    uint32_t it = 0;
    while (have) {
        int nci = ci, nframe = frame + gridDim.y;
        if (nframe >= n_frames) { nci = ci + gridDim.x; nframe = blockIdx.y; }
        const bool nhave = nci < n_cells;
        uint8_t *win = fast_raw + lay.win_off + buf * lay.win_bytes;
        fast_mbar_wait(bar0 + 8 * buf, (it >> 1) & 1u); /* this item's window has landed */
        __syncthreads(); /* and everybody is done with the other buffer and the lists of the previous item */
        FAST_PROBE(2);
        uint32_t norigin = origin;
        int nlevel = level;
        if (nhave) {
            if (nci != ci) { /* never in the batched shape (gridDim.x == n_cells) */
                const BorbCell nc = cells[nci];
                norigin = fast_tma_origin(nc);
                nlevel = nc.level;
            }
            if (tid == 0)
                fast_tma_window(maps, norigin, nlevel, frame0 + nframe, win0 + (buf ^ 1) * lay.win_bytes, bar0 + 8 * (buf ^ 1), box_bytes);
        }
        ++it;
        if (new_cell) {
            const BorbCell cell = cells[ci];
            const BorbLevel &L = plan->lv[cell.level];
            dw = cell.ww - 6; dh = cell.wh - 6; /* detection domain */
            /* stage-1 lane geometry: Q words cover a domain row (bytes b0 .. b0+dw-1 of the staged row),
             * R rows per warp step; this lane: row s1_rig of the step, word s1_ql of the row */
            const int s1_b0 = FAST_DOMAIN_BYTE(cell.x0), s1_w0 = s1_b0 >> 2;
            org_off = 3 * WP + s1_b0; /* domain pixel (0,0): staged row 3 */
            const int s1_Q = ((s1_b0 + dw - 1) >> 2) - s1_w0 + 1;
            s1_R = (int)(cell.dmagic >> 8);
            s1_rig = 0;
#pragma unroll
            for (int k = 1; k < 8; ++k) s1_rig += lane >= k * s1_Q;
            s1_on = s1_rig < s1_R;
            const int s1_ql = lane - s1_rig * s1_Q;
            s1_col = 4 * (s1_w0 + s1_ql) - s1_b0; /* domain x of this lane's first pixel (may be < 0) */
            s1_vlo = ((unsigned)s1_col < (unsigned)dw ? 0x00000600u : 0u) |
                     ((unsigned)(s1_col + 1) < (unsigned)dw ? 0x06000000u : 0u);
            s1_vhi = ((unsigned)(s1_col + 2) < (unsigned)dw ? 0x00000600u : 0u) |
                     ((unsigned)(s1_col + 3) < (unsigned)dw ? 0x06000000u : 0u);
            s1_qoff = 3 * (WP / 4) + s1_w0 + s1_ql + s1_rig * (WP / 4); /* this lane's word of staged row 3 + rig */
            cand_off = L.cand_off; cand_cap = L.cand_cap;
            /* window-relative (x+3, y+3) shifted by the cell origin: coords relative to (minBorderX, minBorderY),
             * src/ORBextractor.cc:1087-1088 */
            sxy3 = (3 + cell.sx) | ((3 + cell.sy) << 16);
        }
        FAST_PROBE(3);
        const uint8_t *org = win + org_off;
        const uint32_t *s1_qp = reinterpret_cast<const uint32_t *>(win) + s1_qoff;
        int n2 = 0, n3 = 0, cell_total = 0;
        /* The reference's two cv::FAST calls (:1069-1075) as two passes, the second only for cells
         * where the first found nothing; K(ini) needs nothing below ini: a corner's score is at
         * least the threshold it was detected at, so neighbours that are corners only at the lower
         * threshold can never beat it.  With ini <= min a second pass could only return a subset
         * of the (empty) first. */
        const int n_pass = mn < ini ? 2 : 1;
        for (int pass = 0; pass < n_pass; ++pass) {
            const int t = pass == 0 ? ini : mn;

            /* ---- stage 1: compass pre-test, four pixels per lane ----
             * A lane owns one aligned word (4 adjacent pixels) of a staged row; a warp covers R = 32/Q
             * rows at a time (Q = words covering a domain row), row groups dealt round-robin to the
             * warps.  The pixels are widened to two 16-bit fields per register so that one integer
             * add/subtract tests two pixels against a ring pixel without borrows between fields:
             *   dark   r < v - t  <=>  bit 9 of (v + 511 - t) - r     (field value in [1, 766])
             *   bright r > v + t  <=>  bit 9 of (511 - t - v) + r
             * a 9-arc covers two adjacent compass points, so the pre-test is (N|S) & (E|W). */
            int n1 = 0;
            const uint32_t Kt = (uint32_t)(511 - min(t, 255)) * 0x00010001u;
            for (int y0 = warp * s1_R; y0 < dh; y0 += FAST_NW * s1_R) {
                const int y = y0 + s1_rig;
                uint32_t s_lo = 0, s_hi = 0;
                if (s1_on && y < dh) {
                    const uint32_t *rw = s1_qp + y0 * (WP / 4);
                    const uint32_t C = rw[0], N = rw[-3 * (WP / 4)], S = rw[3 * (WP / 4)];
                    const uint32_t E = __byte_perm(C, rw[1], 0x6543), W = __byte_perm(rw[-1], C, 0x4321);
#define F16LO(X) __byte_perm(X, 0u, 0x4140)
#define F16HI(X) __byte_perm(X, 0u, 0x4342)
                    {
                        const uint32_t v = F16LO(C), A = v + Kt, B = Kt - v;
                        const uint32_t n = F16LO(N), s_ = F16LO(S), e = F16LO(E), w = F16LO(W);
                        const uint32_t dk = ((A - n) | (A - s_)) & ((A - e) | (A - w));
                        const uint32_t br = ((B + n) | (B + s_)) & ((B + e) | (B + w));
                        s_lo = ((dk & 0x02000200u) | ((br & 0x02000200u) << 1)) & s1_vlo;
                    }
                    {
                        const uint32_t v = F16HI(C), A = v + Kt, B = Kt - v;
                        const uint32_t n = F16HI(N), s_ = F16HI(S), e = F16HI(E), w = F16HI(W);
                        const uint32_t dk = ((A - n) | (A - s_)) & ((A - e) | (A - w));
                        const uint32_t br = ((B + n) | (B + s_)) & ((B + e) | (B + w));
                        s_hi = ((dk & 0x02000200u) | ((br & 0x02000200u) << 1)) & s1_vhi;
                    }
#undef F16LO
#undef F16HI
                }
                /* flagged pixels of this lane (bit 9 = dark, bit 10 = bright, per field) */
                const uint32_t any_lo = (s_lo | (s_lo >> 1)) & 0x02000200u, any_hi = (s_hi | (s_hi >> 1)) & 0x02000200u;
                const int c = __popc(any_lo | (any_hi >> 1));
                if (__ballot_sync(0xffffffffu, c != 0) == 0u) continue; /* (folding this vote into the three below: slower) */
                /* exclusive prefix of the per-lane counts (0 .. 4) from three ballots, one per bit of the count:
                 * independent votes instead of a chain of five dependent shuffles */
                const uint32_t v0 = __ballot_sync(0xffffffffu, c & 1), v1 = __ballot_sync(0xffffffffu, c & 2),
                               v2 = __ballot_sync(0xffffffffu, c & 4);
                const int before = __popc(v0 & below) + 2 * __popc(v1 & below) + 4 * __popc(v2 & below);
                uint16_t *dst = list1 + n1 + before;
                n1 += __popc(v0) + 2 * __popc(v1) + 4 * __popc(v2);
                const uint32_t pixb = (uint32_t)((y << 6) + s1_col);
                const uint32_t sd0 = (s_lo >> 9) & 3u, sd1 = (s_lo >> 25) & 3u, sd2 = (s_hi >> 9) & 3u, sd3 = (s_hi >> 25) & 3u;
                if (sd0) *dst++ = (uint16_t)(pixb | (sd0 << 12));
                if (sd1) *dst++ = (uint16_t)((pixb + 1) | (sd1 << 12));
                if (sd2) *dst++ = (uint16_t)((pixb + 2) | (sd2 << 12));
                if (sd3) *dst++ = (uint16_t)((pixb + 3) | (sd3 << 12));
            }
            __syncwarp();
            FAST_PROBE(4);

            /* ---- stage 2: full test + score of the listed pixels ---- */
            n2 = 0;
            const uint32_t tcmp = (uint32_t)(t + 256) * 0x00010001u;
            for (int i = lane; i - lane < n1; i += 32) {
                bool corner = false;
                uint32_t ent = 0;
                if (i < n1) {
                    ent = list1[i];
                    const int x = ent & 63, y = (ent >> 6) & 63;
                    const uint8_t *p = org + y * WP + x;
                    const int v = p[0];
                    /* ring positions j and j + 8 packed as 16-bit halves */
                    uint32_t rp[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        rp[k] = (uint32_t)FAST_RING(p, WP, k) | ((uint32_t)FAST_RING(p, WP, k + 8) << 16);
                    /* contrast on the flagged side, biased by 256 so both halves stay positive: ring darker
                     * (bit 12): 256 + v - r, else 256 + r - v.  One multiply-add per pair (written as PTX: the
                     * compiler otherwise negates and selects, on the ALU pipe this kernel is bound by).  Mx holds the
                     * biased score in BOTH halves, so "score > t" is one unsigned compare against (t + 256) * 0x10001. */
                    uint32_t Mx;
                    {
                        const bool darker = (ent & (1u << 12)) != 0;
                        const uint32_t sg = darker ? 0xffffffffu : 1u;
                        const uint32_t off = (uint32_t)(256 + (darker ? v : -v)) * 0x00010001u;
                        uint32_t e[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(e[k]) : "r"(rp[k]), "r"(sg), "r"(off));
                        Mx = fast_arc_minmax3_x2(e);
                    }
                    if (Mx <= tcmp && (ent >> 12) == 3u) { /* flagged on both sides (rare): try the other */
                        const uint32_t off = (uint32_t)(256 - v) * 0x00010001u;
                        uint32_t e[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) e[k] = rp[k] + off;
                        Mx = fast_arc_minmax3_x2(e);
                    }
                    corner = Mx > tcmp;
                    if (corner) sc[(y + 1) * FAST_SPITCH + x + 1] = (uint8_t)(Mx - 1u); /* (M - 256) - 1 in the low byte */
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, corner);
                if (corner) list2[n2 + __popc(bal & below)] = (uint16_t)ent;
                n2 += __popc(bal);
            }
            __syncthreads();
            FAST_PROBE(5);

            /* ---- stage 3: strict 3x3 NMS inside the domain ---- */
            n3 = 0;
            for (int i = lane; i - lane < n2; i += 32) {
                bool keep = false;
                uint32_t rec = 0;
                if (i < n2) {
                    const uint32_t ent = list2[i];
                    const int x = ent & 63, y = (ent >> 6) & 63;
                    const uint8_t *q = sc + (y + 1) * FAST_SPITCH + x + 1;
                    const int c = q[0];
                    keep = c > q[-1] && c > q[1] && c > q[-FAST_SPITCH - 1] && c > q[-FAST_SPITCH] &&
                           c > q[-FAST_SPITCH + 1] && c > q[FAST_SPITCH - 1] && c > q[FAST_SPITCH] &&
                           c > q[FAST_SPITCH + 1];
                    rec = (ent & 0xfffu) | ((uint32_t)c << 12);
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                if (keep) list3[n3 + __popc(bal & below)] = rec;
                n3 += __popc(bal);
            }
            if (lane == 0) wcnt[pass * FAST_NW + warp] = n3;
            __syncthreads();
            FAST_PROBE(6);
            cell_total = 0;
#pragma unroll
            for (int w = 0; w < FAST_NW; ++w) cell_total += wcnt[pass * FAST_NW + w];
            if (cell_total > 0) break;
        }

        /* ---- emit this warp's survivors; one global atomicAdd per warp ---- */
        if (cell_total > 0 && n3 > 0) {
            int base = 0;
            if (lane == 0) base = atomicAdd(ncand + frame * BORB_MAXL + level, n3);
            base = __shfl_sync(0xffffffffu, base, 0);
            uint32_t *oxy = cand_xy + (size_t)frame * (size_t)plan->cand_total + cand_off;
            uint32_t *osc = cand_score + (size_t)frame * (size_t)plan->cand_total + cand_off;
            for (int i = lane; i < n3; i += 32) {
                const uint32_t rec = list3[i];
                const int slot = base + i;
                if (slot < cand_cap) {
                    oxy[slot] = (uint32_t)sxy3 + ((rec & 63u) | ((rec << 10) & 0x3f0000u)); /* x | y << 16 */
                    osc[slot] = rec >> 12;
                }
            }
        }
        /* ---- restore the all-zero score plane for the next cell (own corners only) ---- */
        for (int i = lane; i < n2; i += 32) {
            const uint32_t ent = list2[i];
            sc[(((ent >> 6) & 63) + 1) * FAST_SPITCH + (ent & 63) + 1] = 0;
        }
        /* the next cell's first write to the score plane / wcnt comes after its top barrier */
        FAST_PROBE(7);
        have = nhave;
        new_cell = nci != ci;
        if (nhave) { origin = norigin; level = nlevel; ci = nci; frame = nframe; buf ^= 1; }
    }
}

#endif
