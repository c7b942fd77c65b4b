/*
 * k_pyramid.cuh -- K1: scale pyramid level = bilinear resize of the previous level + reflect-101
 * frame, one kernel per level (the level chain is serial: src/ORBextractor.cc:1400-1443 resizes
 * level l from level l-1, not from the input).
 *
 * Arithmetic = OpenCV 4.x cv::resize(INTER_LINEAR) for CV_8UC1 (oracle/cvprim.c):
 *   horizontal  r  = s[x0]*a0 + s[x0+1]*a1            (11-bit weights, int32)
 *   vertical    d  = (((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2
 * (a0+a1 = b0+b1 = 2048, so d <= 255 and OpenCV's saturate_cast never clips.)
 * The tap tables depend only on (cols, rows) and are built on the host once per geometry.
 *
 * Frame pixels are computed at their reflected interior coordinate (no second pass, no
 * intra-level synchronisation).  v3: a CTA owns a 128 x 16 tile of the padded level and stages the
 * source rows / columns its taps reach in shared memory (16-byte coalesced loads, 32-bit shared
 * addressing for the byte gathers); a lane owns 4 adjacent columns (one aligned 32-bit word) with
 * its horizontal taps in registers, a warp 8 consecutive rows.
 */
#ifndef K_PYRAMID_CUH
#define K_PYRAMID_CUH

#include "borb_internal.h"
#include "borb_cta.cuh"

#define PYR_NT 128
#ifndef PYR_TH
#define PYR_TH 32      /* padded output rows per CTA (8 per warp); measured: 8 -> 1.56, 16 -> 1.19, 32 -> 1.05, 64 -> 1.14 ms */
#endif
#define PYR_TW 128     /* padded output columns per CTA (4 per lane) */
#define PYR_LPAD 20    /* columns computed left of the image: 19-px frame rounded up to a word */
#define PYR_SP_MAX 320 /* staged source row pitch: bytes; scale factors up to ~2.3 */

__device__ __forceinline__ int reflect101(int i, int n)
{
    /* |i| excursions are at most BORB_XPAD < n here */
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

__host__ __device__ __forceinline__ int pyr_words_per_row(int w) { return (w + BORB_EDGE + PYR_LPAD + 3) >> 2; }

/* level 0: copy of the input with the frame (src/ORBextractor.cc:1436-1439).  A thread writes 16
 * bytes (one aligned uint4) of a padded row.  Two launch shapes, so that no warp mixes the two paths (one launch for
 * everything ran with 11 of 32 lanes active: the frame units' per-byte path diverged two of the three warps of a row):
 *   INTERIOR  units that lie inside the image: the 16 contiguous source bytes, at whatever alignment the input's row
 *             step gives them, as 5 aligned words + funnel shifts.  Grid (interior units / 32, padded rows / 4, frames),
 *             block (32, 4).
 *   FRAME     the units that hold frame columns (two on the left, the rest on the right), byte by byte at the
 *             reflected coordinate; a warp = 32 consecutive padded ROWS of one such unit column.
 *             Grid (padded rows / 128, frame units, frames), block (32, 4). */
template <bool INTERIOR>
__global__ void __launch_bounds__(PYR_NT)
k_pyr_level0(const __grid_constant__ BorbPlan plan_c, const uint8_t *__restrict__ img, size_t step,
             size_t frame_stride, uint8_t *__restrict__ pyr)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    const BorbLevel &L = plan->lv[0];
    const int n_int = L.w >> 4; /* units with 0 <= c and c + 15 < w */
    int r, c;
    if (INTERIOR) {
        r = blockIdx.y * 4 + threadIdx.y;                  /* padded row */
        const int u = blockIdx.x * 32 + threadIdx.x;
        if (r >= L.prows || u >= n_int) return;
        c = 16 * u;                                        /* first of 16 level columns */
    } else {
        r = (blockIdx.x * 4 + threadIdx.y) * 32 + threadIdx.x;
        if (r >= L.prows) return;
        const int e = blockIdx.y;                          /* frame unit: 0, 1 left of the image, then right of it */
        c = e < 2 ? 16 * e - BORB_XPAD : 16 * (n_int + e - 2);
    }
    const uint8_t *src = img + (size_t)blockIdx.z * frame_stride;
    uint8_t *dst = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes + L.off + (size_t)r * L.pitch + BORB_XPAD + c;
    const int y = reflect101(r - BORB_EDGE, L.h);
    const uint8_t *srow = src + (size_t)y * step;
    uint32_t v[4];
    /* the fifth word is read only when it holds bytes of this unit, so nothing is read past the word that holds the row's
     * last needed byte; the image's very last row still takes the byte path (a checker would flag its last word) */
    if (INTERIOR && (y + 1 < L.h || blockIdx.z + 1 < gridDim.z)) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(srow + c);
        const uint32_t *base = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        uint32_t w[5];
#pragma unroll
        for (int k = 0; k < 4; ++k) w[k] = __ldg(base + k);
        w[4] = sh ? __ldg(base + 4) : 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __funnelshift_r(w[k], w[k + 1], sh);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                int x = min(max(c + 4 * k + b, -BORB_EDGE), L.w - 1 + BORB_EDGE); /* beyond the frame: don't care */
                x = reflect101(x, L.w);
                v[k] |= (uint32_t)__ldg(srow + x) << (8 * b);
            }
        }
    }
    *reinterpret_cast<uint4 *>(dst) = make_uint4(v[0], v[1], v[2], v[3]);
}

/* level l > 0 from level l-1.  CTA = PYR_TW x PYR_TH tile of the padded level; the source rows and
 * columns the tile's taps reach are staged in shared memory with 16-byte loads, then a warp
 * produces 4 consecutive rows, a lane 4 adjacent columns of each (one aligned 32-bit store).
 * Grid: (ceil(4*wpr / PYR_TW), ceil(prows / PYR_TH), frames). */
__global__ void __launch_bounds__(PYR_NT)
k_pyr_resize(const __grid_constant__ BorbPlan plan_c, int l, const int *__restrict__ tab_i,
             const short *__restrict__ tab_s, uint8_t *__restrict__ pyr, int sp /* staged pitch */, int srows)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    extern __shared__ __align__(16) unsigned char pyr_tile[];
    const BorbLevel &L = plan->lv[l];
    const BorbLevel &S = plan->lv[l - 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * PYR_TW + 4 * lane - PYR_LPAD;
    const int r0 = blockIdx.y * PYR_TH;
    uint8_t *frame = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes;
    const uint8_t *sbase = frame + S.off + (size_t)BORB_EDGE * S.pitch + BORB_XPAD;

    /* horizontal taps of this lane's 4 columns: source column and the weight pair (a0 | a1 << 16) */
    int sx[4];
    uint32_t aw[4];
    int sx_lo = 0x7fffffff, sx_hi = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = min(c + k, L.w - 1 + BORB_EDGE);
        x = reflect101(x, L.w);
        sx[k] = tab_i[L.xofs_off + x];
        aw[k] = *reinterpret_cast<const uint32_t *>(tab_s + L.xa_off + 2 * x); /* a0 | a1 << 16: the tables' offsets are even */
        sx_lo = min(sx_lo, sx[k]);
        sx_hi = max(sx_hi, min(sx[k] + 1, S.w - 1));
    }
    /* source window of the tile: columns (all 4 warps hold the same columns) and rows */
    sx_lo = __reduce_min_sync(0xffffffffu, sx_lo) & ~15;
    sx_hi = __reduce_max_sync(0xffffffffu, sx_hi);
    /* Row geometry once per CTA: lane rr of every warp holds output row r0 + rr of the tile (PYR_TH <= 32):
     * its two clamped source rows and the vertical weight pair; the row loop fetches them by shuffle. */
    static_assert(PYR_TH <= 32, "one tile row per lane");
    int sy_lo = 0x7fffffff, sy_hi = 0, my_sy0 = 0, my_sy1 = 0;
    uint32_t my_b = 0;
    if (lane < PYR_TH && r0 + lane < L.prows) {
        const int y = reflect101(r0 + lane - BORB_EDGE, L.h);
        const int sy = tab_i[L.yofs_off + y];
        my_sy0 = min(max(sy, 0), S.h - 1);
        my_sy1 = min(max(sy + 1, 0), S.h - 1);
        my_b = *reinterpret_cast<const uint32_t *>(tab_s + L.yb_off + 2 * y); /* b0 | b1 << 16 */
        sy_lo = my_sy0;
        sy_hi = my_sy1;
    }
    sy_lo = __reduce_min_sync(0xffffffffu, sy_lo);
    sy_hi = __reduce_max_sync(0xffffffffu, sy_hi);
    const uint32_t my_rows = (uint32_t)(my_sy0 - sy_lo) | ((uint32_t)(my_sy1 - sy_lo) << 16);
    /* Stage the window as byte PAIRS: element i of a staged row is (s[i], s[i+1]) as one aligned
     * 16-bit value, so a horizontal tap pair is one LDS.U16 and the horizontal sum
     * s[x0]*a0 + s[x0+1]*a1 one dp2a.  (Where cv::resize clamps x0+1 to the last column its weight
     * a1 is 0, so the pair's second byte is immaterial there.) */
    const int sp2 = 2 * sp;
    {
        const int vpr = (sx_hi - sx_lo) / 16 + 1;             /* uint4 per source row */
        const int nrow = sy_hi - sy_lo + 1;
        const int nv = vpr * nrow;                            /* host guarantees vpr*16 <= sp, nrow <= srows */
        const uint8_t *g = sbase + (size_t)sy_lo * S.pitch + sx_lo; /* 16-B aligned: pitch, XPAD, sx_lo are */
        const uint32_t vmagic = 65536u / (uint32_t)vpr + 1u;  /* i / vpr for i < 4096, vpr <= 32 */
        for (int i = threadIdx.x; i < nv; i += PYR_NT) {
            const int rr = (int)(((uint32_t)i * vmagic) >> 16), cc = i - rr * vpr;
            const uint32_t *gw = reinterpret_cast<const uint32_t *>(g + (size_t)rr * S.pitch) + 4 * cc;
            const uint4 q = __ldg(reinterpret_cast<const uint4 *>(gw));
            const uint32_t w[5] = {q.x, q.y, q.z, q.w, __ldg(gw + 4)};
            uint32_t o[8];
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                o[2 * m] = __byte_perm(w[m], 0u, 0x2110);            /* (b0,b1) (b1,b2) */
                o[2 * m + 1] = __byte_perm(w[m], w[m + 1], 0x4332);  /* (b2,b3) (b3,b4) */
            }
            uint4 *dstv = reinterpret_cast<uint4 *>(pyr_tile + rr * sp2 + cc * 32);
            dstv[0] = make_uint4(o[0], o[1], o[2], o[3]);
            dstv[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) sx[k] = 2 * (sx[k] - sx_lo); /* byte offset of the pair in a staged row */
    __syncthreads();
    const bool col_on = c < L.w + BORB_EDGE; /* lanes past the frame keep running: they serve the row shuffles */

    /* Both source rows' horizontal sums (already >> 4, as the vertical pass uses them) are computed for every output row:
     * keeping the previous row's sums to reuse the shared one (5 rows of 6 at 1.2) cost more instructions -- the
     * compares, branches and 8 register moves per row -- than the 4 dot products it saved (0.952 -> 0.944 ms). */
    /* a running store pointer: the row's address as (size_t)r * pitch cost 14 instructions per row in SASS */
    const int rw0 = r0 + warp * (PYR_TH / 4);
    uint8_t *drow = frame + L.off + BORB_XPAD + c + (size_t)rw0 * L.pitch;
    const int nrow = min(PYR_TH / 4, L.prows - rw0);
#pragma unroll
    for (int i = 0; i < PYR_TH / 4; ++i, drow += L.pitch) {
        if (i >= nrow) break;
        /* (b * h) >> 16 as the high word of (b << 16) * h: operands are non-negative */
        const uint32_t rw_ = __shfl_sync(0xffffffffu, my_rows, warp * (PYR_TH / 4) + i);
        const uint32_t bw_ = __shfl_sync(0xffffffffu, my_b, warp * (PYR_TH / 4) + i);
        const uint32_t b0 = bw_ << 16, b1 = bw_ & 0xffff0000u;
        const int sy0 = (int)(rw_ & 0xffffu), sy1 = (int)(rw_ >> 16);
        uint32_t n0[4], n1[4];
        {
            const uint8_t *row0 = pyr_tile + sy0 * sp2, *row1 = pyr_tile + sy1 * sp2;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                n0[k] = __dp2a_lo(aw[k], (uint32_t) * reinterpret_cast<const uint16_t *>(row0 + sx[k]), 0u) >> 4;
                n1[k] = __dp2a_lo(aw[k], (uint32_t) * reinterpret_cast<const uint16_t *>(row1 + sx[k]), 0u) >> 4;
            }
        }
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t d = (__umulhi(b0, n0[k]) + __umulhi(b1, n1[k]) + 2u) >> 2;
            v |= d << (8 * k);
        }
        if (col_on) *reinterpret_cast<uint32_t *>(drow) = v;
    }
}

/* level l > 0 from level l-1, SOURCE-row driven (the batch shape).  k_pyr_resize above computes, for every output row, the
 * horizontal sums of BOTH source rows it blends: 2 h-sums per output row, although consecutive output rows share a source
 * row or skip one (the level scale is >= 1: an output row d blends source rows yofs[d], yofs[d] + 1, and yofs is strictly
 * increasing).  Here a warp walks SOURCE rows: one h-sum per source row (1.2 per output row at scale 1.2, + 1 to seed the
 * stream), kept in registers under static names (the walk is fully unrolled, rows alternate between two register sets), and
 * the output row that ends at this source row -- at most one, yinv[s] -- is blended and stored.  The staged pitch is a
 * compile-time constant, so a source row's offset is an LDS immediate and the row loop carries no address arithmetic.
 * The reflect-101 frame ROWS are copies of interior rows: the warp that computes interior row d stores it a second time
 * at its mirror (frame columns are still computed through mirrored taps, as in k_pyr_resize).
 * The host falls back to k_pyr_resize for a level where the preconditions fail (borb_api.cu: pyr_stream_ok).
 * Grid: (ceil(4 * wpr / PYR_TW), ceil((S.h - 1) / (4 * PST_NSW)), frames), PYR_NT threads. */
#ifndef PST_NSW
#define PST_NSW 12                  /* source rows a warp walks (besides the one that seeds its stream) */
#endif
#define PST_ROWS (4 * PST_NSW + 1)
/* PST_SP = staged row pitch in bytes, a template parameter: PYR_SP_MAX holds any window the host admits, 192 the window of
 * the usual scale factors (<= 192 source columns for 128 output columns: 1.5).
 * Taps (second version): the staged rows are plain copies of the source rows; the four taps of a lane's pixels 0, 1 (x0,
 * x0 + 1 of each) lie within two aligned words of a staged row and so do those of pixels 2, 3 (the host checks it per
 * level: scale factors up to ~3), so a row's horizontal sums are 4 LDS.32 + 2 PRMT (a selector gathers a pair's taps:
 * first pixel in bytes 0, 1, second in bytes 2, 3) + 4 dp2a.  The first version staged every row as byte PAIRS (element i =
 * (s[i], s[i+1]), one LDS.U16 per tap pair): twice the shared-memory stores, 8 PRMT per staged vector, and lanes 9.6 bytes
 * apart on the tap loads (half of the kernel's shared-memory wavefronts were their bank conflicts). */
#ifndef PST_MINB
#define PST_MINB 9
#endif
template <int PST_SP>
__global__ void __launch_bounds__(PYR_NT, PST_MINB)
k_pyr_stream(const __grid_constant__ BorbPlan plan_c, int l, const int *__restrict__ tab_i, const short *__restrict__ tab_s,
             const uint2 *__restrict__ yinv /* per source row: output row or ~0u, b0 | b1 << 16 */, uint8_t *__restrict__ pyr)
{
    const BorbPlan *plan = &plan_c;
    extern __shared__ __align__(16) unsigned char pyr_tile[];
    const BorbLevel &L = plan->lv[l];
    const BorbLevel &S = plan->lv[l - 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    /* a lane owns four adjacent columns = one aligned word of the padded row.  (Two column pairs 64 apart per lane halve the
     * shared-memory bank conflicts of the tap loads -- 2.4 wavefronts per load here -- but cost a second, 16-bit store per
     * row: measured 129.0 against 129.6 k frames/s, not kept.) */
    const int c = blockIdx.x * PYR_TW + 4 * lane - PYR_LPAD;
    uint8_t *frame = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes;
    const uint8_t *sbase = frame + S.off + (size_t)BORB_EDGE * S.pitch + BORB_XPAD;

    int sx[4];
    uint32_t aw[4];
    int sx_lo = 0x7fffffff, sx_hi = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = min(c + k, L.w - 1 + BORB_EDGE);
        x = reflect101(x, L.w);
        sx[k] = tab_i[L.xofs_off + x];
        aw[k] = *reinterpret_cast<const uint32_t *>(tab_s + L.xa_off + 2 * x); /* a0 | a1 << 16: the tables' offsets are even */
        sx_lo = min(sx_lo, sx[k]);
        sx_hi = max(sx_hi, min(sx[k] + 1, S.w - 1));
    }
    sx_lo = __reduce_min_sync(0xffffffffu, sx_lo) & ~15;
    sx_hi = __reduce_max_sync(0xffffffffu, sx_hi);
    const int s_tile = blockIdx.y * (4 * PST_NSW);
    const int s_last = min(s_tile + 4 * PST_NSW, S.h - 1);
    {   /* stage source rows s_tile .. s_last of the window; the loads of four vectors are issued before the first of
         * them is stored (a thread stages 6-7 vectors: one at a time, the CTA waited out a DRAM latency per vector) */
        const int vpr = (sx_hi - sx_lo) / 16 + 1; /* host guarantees vpr * 16 <= PST_SP */
        const int nv = vpr * (s_last - s_tile + 1);
        const uint8_t *g = sbase + (size_t)s_tile * S.pitch + sx_lo;
        const uint32_t vmagic = 65536u / (uint32_t)vpr + 1u;
        for (int i0 = threadIdx.x; i0 < nv; i0 += 4 * PYR_NT) {
            uint4 q[4];
            int off[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * PYR_NT;
                const int rr = (int)(((uint32_t)min(i, nv - 1) * vmagic) >> 16), cc = min(i, nv - 1) - rr * vpr;
                q[u] = __ldg(reinterpret_cast<const uint4 *>(g + (size_t)rr * S.pitch) + cc);
                off[u] = i < nv ? rr * PST_SP + cc * 16 : -1;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (off[u] >= 0) *reinterpret_cast<uint4 *>(pyr_tile + off[u]) = q[u];
        }
    }
    const int s0 = s_tile + warp * PST_NSW; /* the row that seeds this warp's stream */
    /* lane j in 1 .. PST_NSW holds the output row that ends at source row s0 + j */
    uint2 g = make_uint2(0xffffffffu, 0u);
    if (lane >= 1 && lane <= PST_NSW && s0 + lane <= S.h - 1) g = __ldg(yinv + s0 + lane);
    __syncthreads();
    if (s0 >= S.h - 1) return;
    const bool col_on = c < L.w + BORB_EDGE;
    /* per pixel pair: the aligned word its taps start in (a pointer into the seed row) and the PRMT selector */
    const uint32_t *A[2];
    uint32_t sel[2];
#pragma unroll
    for (int pr = 0; pr < 2; ++pr) {
        const int xa = sx[2 * pr] - sx_lo, xb = sx[2 * pr + 1] - sx_lo, base = min(xa, xb) & ~3;
        A[pr] = reinterpret_cast<const uint32_t *>(pyr_tile + (warp * PST_NSW) * PST_SP + base);
        sel[pr] = (uint32_t)(xa - base) | ((uint32_t)(xa + 1 - base) << 4) | ((uint32_t)(xb - base) << 8) | ((uint32_t)(xb + 1 - base) << 12);
    }
    /* Output addressing.  The output rows a warp produces are CONSECUTIVE (yofs is strictly increasing), so the store
     * pointer runs: one 64-bit add per row.  Only the first and the last tiles of a level hold rows with a mirror in the
     * reflect-101 frame; they take the variant of the walk that carries the two mirror pointers. */
    const uint32_t ends = __ballot_sync(0xffffffffu, g.x != 0xffffffffu); /* bit j: an output row ends at source row s0 + j */
    if (!ends) return;
    const int nrow = min(PST_NSW, S.h - 1 - s0);
    const uint32_t d_first = __shfl_sync(0xffffffffu, g.x, __ffs(ends) - 1);
    const uint32_t d_last = d_first + (uint32_t)__popc(ends) - 1u;
    const size_t pitch = (size_t)L.pitch;
    uint8_t *drow = frame + L.off + BORB_XPAD + c + (size_t)(d_first + BORB_EDGE) * pitch;
    const bool mirrors = d_first <= (uint32_t)BORB_EDGE || (int)d_last >= L.h - 1 - BORB_EDGE;
    uint32_t H[2][4];
#define PST_HSUM(j)                                                                                                     \
    do {                                                                                                                \
        const uint32_t ta_ = __byte_perm(A[0][(j) * (PST_SP / 4)], A[0][(j) * (PST_SP / 4) + 1], sel[0]);               \
        const uint32_t tb_ = __byte_perm(A[1][(j) * (PST_SP / 4)], A[1][(j) * (PST_SP / 4) + 1], sel[1]);               \
        H[(j) & 1][0] = __dp2a_lo(aw[0], ta_, 0u) >> 4;                                                                 \
        H[(j) & 1][1] = __dp2a_hi(aw[1], ta_, 0u) >> 4;                                                                 \
        H[(j) & 1][2] = __dp2a_lo(aw[2], tb_, 0u) >> 4;                                                                 \
        H[(j) & 1][3] = __dp2a_hi(aw[3], tb_, 0u) >> 4;                                                                 \
    } while (0)
    PST_HSUM(0);
#define PST_BLEND(j)                                                                                                    \
    const uint32_t bw_ = __shfl_sync(0xffffffffu, g.y, (j));                                                            \
    const uint32_t b0 = bw_ << 16, b1 = bw_ & 0xffff0000u;                                                              \
    uint32_t v = 0;                                                                                                     \
    _Pragma("unroll") for (int k = 0; k < 4; ++k) {                                                                     \
        const uint32_t t = (__umulhi(b0, H[((j) - 1) & 1][k]) + __umulhi(b1, H[(j) & 1][k]) + 2u) >> 2;                 \
        v |= t << (8 * k);                                                                                              \
    }
    if (!mirrors && nrow == PST_NSW) { /* the common walk: a full one, no frame rows */
#pragma unroll
        for (int j = 1; j <= PST_NSW; ++j) {
            PST_HSUM(j);
            if (ends & (1u << j)) { /* warp-uniform */
                PST_BLEND(j)
                if (col_on) *reinterpret_cast<uint32_t *>(drow) = v;
                drow += pitch;
            }
        }
    } else {
        /* frame rows are copies of interior rows 1 .. 19 (padded row 19 - d) and h-20 .. h-2 (padded row 2h + 17 - d) */
        uint32_t d = d_first;
        const uint32_t bot0 = (uint32_t)(L.h - 1 - BORB_EDGE);
        uint8_t *const row19 = frame + L.off + BORB_XPAD + c + (size_t)BORB_EDGE * pitch;
        const long long mir2 = 2LL * (L.h - 1);
#pragma unroll
        for (int j = 1; j <= PST_NSW; ++j) {
            if (j > nrow) break;
            PST_HSUM(j);
            if (ends & (1u << j)) {
                PST_BLEND(j)
                if (col_on) {
                    *reinterpret_cast<uint32_t *>(drow) = v;
                    if (d - 1u < (uint32_t)BORB_EDGE) *reinterpret_cast<uint32_t *>(row19 - (size_t)d * pitch) = v;
                    if (d - bot0 < (uint32_t)BORB_EDGE) *reinterpret_cast<uint32_t *>(row19 + (size_t)(mir2 - (long long)d) * pitch) = v;
                }
                drow += pitch;
                ++d;
            }
        }
    }
#undef PST_HSUM
#undef PST_BLEND
}

/* ---- the whole level chain in ONE launch, for a frame or two (the latency shape) ----
 * Eight dependent launches of a few microseconds each are mostly launch gaps when the grid is a single frame.
 * Here a CTA owns one tile of every level (the levels are partitioned by the same gx x gy grid) and walks the
 * levels itself: level l of its tile is computed in shared memory from level l-1 of its tile, written to the
 * padded plane (with the reflect-101 frame where the tile touches the image edge) and kept as the source of
 * level l+1.  A pixel is a pure function of the previous level's pixels, so the few halo pixels that neighbouring
 * CTAs compute twice are identical: the planes are bit-for-bit those of k_pyr_level0 + k_pyr_resize.  The host
 * closes the dependencies per tile: need(l) = own(l) + frame sources + taps of need(l+1)  (PyrChainRect). */
#ifndef PCH_NT
#define PCH_NT 512
#endif
#define PCH_TW 20 /* tile size aimed at on the LAST level (measured, single KITTI frame: 24x16 -> 62.0, 20x16 -> 61.3, 20x12 -> 60.0 us per call) */
#define PCH_TH 12
#define PCH_EDGE_WX 0.5 /* relative width / height of the tiles on the image edge (borb_api.cu build_chain; measured, single KITTI frame, kernels p50: 1.0 x 1.0 51.3 us, 0.8 x 0.8 49.8, 0.6 x 0.5 49.2, 0.5 x 0.4 49.1, 0.4 x 0.3 48.8) */
#define PCH_EDGE_WY 0.4

struct PyrChainRect {
    short nx0, ny0, nw, nh;   /* rectangle of the level this CTA computes (level pixels) */
    short ox0, oy0, ox1, oy1; /* rectangle it writes out; x bounds are multiples of 4 or the level width */
};

struct PyrChainTile { /* one per tile, built on the host (borb_api.cu: build_chain) */
    PyrChainRect r[BORB_MAXL];
    int toff[BORB_MAXL + 1];      /* first tap-table entry of level l; [nlevels] = entries in all */
    uint32_t magic[BORB_MAXL][2]; /* divisors of the two per-level loops: words per computed row, per written row */
    int pad[3];                   /* sizeof is a multiple of 16 */
};
static_assert(sizeof(PyrChainTile) % 16 == 0, "copied as uint4");

__host__ __device__ __forceinline__ int pch_pitch(int nw) { return (nw + 3) & ~3; }
/* i / w for i, w < 65536 by one multiply (the divisor is the same for the whole CTA) */
__device__ __forceinline__ uint32_t pch_magic(int w) { return w > 1 ? 0xffffffffu / (uint32_t)w + 1u : 0u; }
__device__ __forceinline__ int pch_div(int i, uint32_t m, int w) { return w > 1 ? (int)__umulhi((uint32_t)i, m) : i; }

/* -DBORB_PCH_PROBE (tools/probe_chain.py, never in the shipped library): %globaltimer of thread 0 per tile: entry, level 0
 * staged, then per level: computed, written */
#ifdef BORB_PCH_PROBE
__device__ unsigned long long g_pch_probe[256 * 20];
#define PCH_PROBE(k)                                                                       \
    do {                                                                                   \
        if (threadIdx.x == 0 && blockIdx.y == 0 && blockIdx.x < 256) {                     \
            unsigned long long t_;                                                         \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                         \
            g_pch_probe[blockIdx.x * 20 + (k)] = t_;                                       \
        }                                                                                  \
    } while (0)
#else
#define PCH_PROBE(k) do { } while (0)
#endif

__global__ void __launch_bounds__(PCH_NT)
k_pyr_chain(const __grid_constant__ BorbPlan plan_c, const PyrChainTile *__restrict__ tiles,
            const uint2 *__restrict__ gtabs, int tstride, const uint8_t *__restrict__ img, size_t step,
            size_t frame_stride, uint8_t *__restrict__ pyr, int off_b, int off_t)
{
    const BorbPlan *plan = &plan_c;
    extern __shared__ __align__(16) unsigned char pch_smem[];
    borb_pdl_launch_dependents(); /* K2's CTAs may take their places and set up while the chain runs */
    PCH_PROBE(0);
    const int tid = threadIdx.x, nl = plan->nlevels;
    const PyrChainTile *const gT = tiles + blockIdx.x;
    uint2 *const tabs = reinterpret_cast<uint2 *>(pch_smem + off_t);
    uint8_t *const frame = pyr + (size_t)blockIdx.y * (size_t)plan->pyr_bytes;

    /* prologue, all plain copies in flight together: the tile's header and tap tables (built on the host) and
     * level 0 of the tile from the input */
    __shared__ __align__(16) PyrChainTile sT_;
    if (tid < (int)(sizeof(PyrChainTile) / 16))
        reinterpret_cast<uint4 *>(&sT_)[tid] = __ldg(reinterpret_cast<const uint4 *>(gT) + tid);
    {
        const int n4 = (__ldg(&gT->toff[nl]) + 1) >> 1; /* uint4 = two entries */
        const uint4 *g4 = reinterpret_cast<const uint4 *>(gtabs + (size_t)blockIdx.x * tstride);
        for (int i = tid; i < n4; i += PCH_NT) reinterpret_cast<uint4 *>(tabs)[i] = __ldg(g4 + i);
    }
    {   /* aligned words of the input (whatever its row step), stored byte by byte */
        const PyrChainRect r = gT->r[0];
        const int pitch = pch_pitch(r.nw), K = (r.nw + 6) >> 2;
        const uint8_t *src = img + (size_t)blockIdx.y * frame_stride + (size_t)r.ny0 * step + r.nx0;
        const uint32_t km = pch_magic(K);
        for (int i0 = tid; i0 < K * r.nh; i0 += 4 * PCH_NT) {
            uint32_t w[4];
            int p0[4], rr[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { /* the loads of four words first, then their stores */
                const int i = i0 + u * PCH_NT;
                rr[u] = pch_div(i, km, K);
                const int k = i - rr[u] * K;
                const uintptr_t a = reinterpret_cast<uintptr_t>(src + (size_t)rr[u] * step);
                const int ph = (int)(a & 3);
                p0[u] = 4 * k - ph;
                const bool on = i < K * r.nh && p0[u] < r.nw; /* every word read holds at least one byte of the row */
                w[u] = on ? __ldg(reinterpret_cast<const uint32_t *>(a - ph) + k) : 0u;
                if (!on) p0[u] = r.nw;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint8_t *d = pch_smem + rr[u] * pitch;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (p0[u] + j >= 0 && p0[u] + j < r.nw) d[p0[u] + j] = (uint8_t)(w[u] >> (8 * j));
            }
        }
    }
    __syncthreads();
    PCH_PROBE(1);
    const PyrChainRect *const sR = sT_.r;
    const int *const sToff = sT_.toff;

    for (int l = 0; l < nl; ++l) {
        const BorbLevel &L = plan->lv[l];
        const PyrChainRect r = sR[l];
        uint8_t *const buf = pch_smem + ((l & 1) ? off_b : 0);
        const int pitch = pch_pitch(r.nw);
        if (l > 0) {
            const uint8_t *const sbuf = pch_smem + ((l & 1) ? 0 : off_b);
            const int spitch = pch_pitch(sR[l - 1].nw);
            const int wpr = pitch >> 2;
            const uint2 *const xt = tabs + sToff[l], *const yt = xt + 3 * wpr;
            const uint32_t wm = sT_.magic[l][0];
            for (int i = tid; i < wpr * r.nh; i += PCH_NT) {
                const int y = pch_div(i, wm, wpr), xw = i - y * wpr;
                const uint2 ey = yt[y];
                const uint32_t *row0 = reinterpret_cast<const uint32_t *>(sbuf + (ey.x & 0xffffu) * spitch);
                const uint32_t *row1 = reinterpret_cast<const uint32_t *>(sbuf + (ey.x >> 16) * spitch);
                const uint32_t b0 = ey.y << 16, b1 = ey.y & 0xffff0000u;
                /* four pixels: two aligned words of each source row per pixel pair, one PRMT gathers the pair's four taps
                 * (x0, x1 of the first pixel in bytes 0, 1, of the second in bytes 2, 3), one dp2a per pixel and row is
                 * the horizontal blend (build_chain in borb_api.cu lays the table out) */
                const uint2 e0 = xt[3 * xw], e1 = xt[3 * xw + 1], e2 = xt[3 * xw + 2];
                const uint32_t wa = e0.x & 0xffffu, wb = e0.x >> 16, sa = e0.y & 0xffffu, sb = e0.y >> 16;
                const uint32_t ta0 = __byte_perm(row0[wa], row0[wa + 1], sa), tb0 = __byte_perm(row0[wb], row0[wb + 1], sb);
                const uint32_t ta1 = __byte_perm(row1[wa], row1[wa + 1], sa), tb1 = __byte_perm(row1[wb], row1[wb + 1], sb);
                uint32_t n0[4], n1[4];
                n0[0] = __dp2a_lo(e1.x, ta0, 0u) >> 4; n0[1] = __dp2a_hi(e1.y, ta0, 0u) >> 4;
                n0[2] = __dp2a_lo(e2.x, tb0, 0u) >> 4; n0[3] = __dp2a_hi(e2.y, tb0, 0u) >> 4;
                n1[0] = __dp2a_lo(e1.x, ta1, 0u) >> 4; n1[1] = __dp2a_hi(e1.y, ta1, 0u) >> 4;
                n1[2] = __dp2a_lo(e2.x, tb1, 0u) >> 4; n1[3] = __dp2a_hi(e2.y, tb1, 0u) >> 4;
                uint32_t v = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) v |= ((__umulhi(b0, n0[k]) + __umulhi(b1, n1[k]) + 2u) >> 2) << (8 * k);
                *reinterpret_cast<uint32_t *>(buf + y * pitch + 4 * xw) = v;
            }
            __syncthreads();
        }
        PCH_PROBE(2 + 2 * l);
            /* write the tile out, with the frame where it touches the image edge */
        if (r.ox1 > r.ox0 && r.oy1 > r.oy0) {
            const int ex0 = r.ox0 == 0 ? -BORB_XPAD : r.ox0, ex1 = r.ox1 == L.w ? L.pitch - BORB_XPAD : r.ox1;
            const int ey0 = r.oy0 == 0 ? -BORB_EDGE : r.oy0, ey1 = r.oy1 == L.h ? L.h + BORB_EDGE : r.oy1;
            const int wpr = (ex1 - ex0) >> 2;
            uint8_t *const dst = frame + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
            const uint32_t om = sT_.magic[l][1], osh = (uint32_t)((-r.nx0) & 3) * 8u; /* px is a multiple of 4 */
            for (int i = tid; i < wpr * (ey1 - ey0); i += PCH_NT) {
                const int yy = pch_div(i, om, wpr), py = ey0 + yy, px = ex0 + 4 * (i - yy * wpr);
                const uint8_t *row = buf + (reflect101(py, L.h) - r.ny0) * pitch - r.nx0;
                uint32_t v = 0;
                if (px >= 0 && px + 3 < L.w) { /* four interior columns: two aligned words of the staged row */
                    const uint32_t *w = reinterpret_cast<const uint32_t *>(row + r.nx0 + ((px - r.nx0) & ~3));
                    v = __funnelshift_r(w[0], w[1], osh);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int x = reflect101(min(max(px + j, -BORB_EDGE), L.w - 1 + BORB_EDGE), L.w); /* beyond the frame: don't care */
                        v |= (uint32_t)row[x] << (8 * j);
                    }
                }
                *reinterpret_cast<uint32_t *>(dst + (long long)py * L.pitch + px) = v;
            }
        }
        PCH_PROBE(3 + 2 * l);
        }
}

#endif
