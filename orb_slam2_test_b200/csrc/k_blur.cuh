/*
 * k_blur.cuh -- K5: 7x7 sigma-2 Gaussian of every pyramid level (src/ORBextractor.cc:1375-1377).
 *
 * OpenCV 4.x filters CV_8U with an 8.8 fixed-point kernel; for ksize 7, sigma 2 the quantised taps
 * are [18,34,48,56,48,34,18]/256 per axis (oracle/cvprim.c):
 *   row pass  H = sum k[i] * src            (<= 65280, 16 bit)
 *   col pass  V = sum k[j] * H              (32 bit),   dst = (V + 32768) >> 16
 * The reference blurs an un-padded clone with BORDER_REFLECT_101; the pyramid's own 19-pixel
 * frame holds exactly those reflected values, so halos are read straight from it.
 *
 * A warp owns a 128-column x BORB_BLUR_TH-row strip.  Its source tile (128 + 2 x 16 columns,
 * BORB_BLUR_TH + 6 rows) arrives in shared memory by ONE TMA box load (per-warp mbarrier; v3 read
 * global memory inside the row loop and spent 0.75 ms per 512 frames where this spends 0.55).  The
 * box starts 16 columns left of the strip because TMA boxes start on 16-byte boundaries; parts of
 * the box outside the padded plane are zero-filled and only ever feed pixels outside the level.
 * A thread owns 4 adjacent columns (one aligned 32-bit word of every row).  Walking down the strip
 * it reads three words per row (its own and the two neighbours'; the 160-byte row stride keeps the
 * warp conflict-free), forms the 4 row sums with 8 byte-dot-products (dp4a with the taps packed as
 * bytes: 18,34,48,56 | 48,34,18,0) and keeps the last 7 rows of row sums in registers (the loop is
 * unrolled by 7 so the window is statically indexed); each new row yields one output word.
 * Algorithmic traffic: read 1 B + write 1 B per pixel.
 */
#ifndef K_BLUR_CUH
#define K_BLUR_CUH

#include <cuda.h>

#include "borb_internal.h"

#define BLUR_NT 128 /* 4 warps = 4 consecutive strips of the tile list */

/* Row sums of the 4 columns of word w1, each + 128: the taps sum to 256, so the column pass over
 * (H + 128) carries the + 32768 rounding term for free.  H + 128 <= 65408 still fits 16 bits. */
__device__ __forceinline__ void blur_row_sums(const uint32_t w0, const uint32_t w1, const uint32_t w2, uint32_t h[4])
{
    const uint32_t klo = 18u | (34u << 8) | (48u << 16) | (56u << 24); /* taps 0..3 */
    const uint32_t khi = 48u | (34u << 8) | (18u << 16);               /* taps 4..6 */
    /* output column c+i needs bytes c+i-3 .. c+i+3 */
    h[0] = __dp4a(__byte_perm(w0, w1, 0x4321), klo, __dp4a(__byte_perm(w1, w2, 0x4321), khi, 128u));
    h[1] = __dp4a(__byte_perm(w0, w1, 0x5432), klo, __dp4a(__byte_perm(w1, w2, 0x5432), khi, 128u));
    h[2] = __dp4a(__byte_perm(w0, w1, 0x6543), klo, __dp4a(__byte_perm(w1, w2, 0x6543), khi, 128u));
    h[3] = __dp4a(w1, klo, __dp4a(w2, khi, 128u));
}

/* Column pass on row sums packed in vertical pairs, q = (H[i-1], H[i]) as two 16-bit fields: a
 * 2-way dot product (dp2a) applies two taps per instruction.  q1 = rows r, r+1; q3 = r+2, r+3;
 * q5 = r+4, r+5; q6 = r+5, r+6 (its low half gets tap 0). */
__device__ __forceinline__ uint32_t blur_col(const uint32_t q1, const uint32_t q3, const uint32_t q5, const uint32_t q6)
{
    const uint32_t ka = 18u | (34u << 8) | (48u << 16) | (56u << 24); /* taps 0,1 | 2,3 */
    const uint32_t kb = 48u | (34u << 8) | (0u << 16) | (18u << 24);  /* taps 4,5 | -,6 */
    return __dp2a_lo(q1, ka, __dp2a_hi(q3, ka, __dp2a_lo(q5, kb, __dp2a_hi(q6, kb, 0u))));
}

#define BLUR_BOX_W (BORB_BLUR_TW + 32)
#define BLUR_BOX_H (BORB_BLUR_TH + 6)
#define BLUR_TILE_BYTES ((BLUR_BOX_W * BLUR_BOX_H + 127) & ~127)
#define BLUR_SMEM_BYTES (128 + (BLUR_NT / 32) * BLUR_TILE_BYTES)

struct BlurMaps { /* one TMA descriptor per pyramid level: box BLUR_BOX_W x BLUR_BOX_H x 1 frame */
    CUtensorMap m[BORB_MAXL];
};

__device__ __forceinline__ void blur_row_sums_smem(const uint8_t *row, uint32_t h[4])
{
    const uint32_t *w = reinterpret_cast<const uint32_t *>(row);
    blur_row_sums(w[-1], w[0], w[1], h);
}

__global__ void __launch_bounds__(BLUR_NT)
k_blur_strips(const __grid_constant__ BorbPlan plan_c, const __grid_constant__ BlurMaps maps,
                  const BorbTile *__restrict__ tiles, const int frame0, uint8_t *__restrict__ blur)
{
    extern __shared__ __align__(128) unsigned char blur_smem[];
    const BorbPlan *plan = &plan_c;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t_idx = blockIdx.x * (BLUR_NT / 32) + warp;
    if (t_idx >= plan->n_tiles) return; /* whole warp */
    const BorbTile t = tiles[t_idx];
    const BorbLevel &L = plan->lv[t.level];
    uint8_t *tile = blur_smem + 128 + warp * BLUR_TILE_BYTES;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(blur_smem + 8 * warp);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(BLUR_BOX_W * BLUR_BOX_H) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(tile)), "l"(&maps.m[t.level]), "r"(BORB_XPAD + t.x0 - 16),
                       "r"(BORB_EDGE + t.y0 - 3), "r"(frame0 + (int)blockIdx.y), "r"(bar)
                     : "memory");
    }
    __syncwarp();
    asm volatile("{\n\t.reg .pred p;\n\tBLUR_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra BLUR_DONE;\n\tbra BLUR_WAIT;\n\tBLUR_DONE:\n\t}" ::"r"(bar), "r"(0)
                 : "memory");
    const int c = t.x0 + 4 * lane;
    if (c >= L.w) return;
    const int pitch = L.pitch;
    const int rows = min(BORB_BLUR_TH, L.h - t.y0);
    /* c is a multiple of 4 and the interior starts 32-byte aligned: aligned word stores; the up-to-3
     * bytes past the last column land in our own pitch padding */
    uint8_t *orow = blur + (size_t)(frame0 + blockIdx.y) * (size_t)plan->pyr_bytes + L.off +
                    (size_t)(BORB_EDGE + t.y0) * pitch + BORB_XPAD + c;
    const uint8_t *src = tile + 16 + 4 * lane; /* tile row 0 = level row y0 - 3, this lane's first column */

    /* Q[i % 6] = (H[i-1], H[i]) of tile rows i-1, i; output row r needs Q[r+1], Q[r+3], Q[r+5], Q[r+6] */
    uint32_t Q[6][4], hp[4], hn[4];
    blur_row_sums_smem(src, hp);
#pragma unroll
    for (int i = 1; i < 6; ++i) {
        blur_row_sums_smem(src + i * BLUR_BOX_W, hn);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            Q[i][k] = __byte_perm(hp[k], hn[k], 0x5410);
            hp[k] = hn[k];
        }
    }
    /* groups of 6 rows run unguarded (the box holds BORB_BLUR_TH + 6 rows, BORB_BLUR_TH a multiple
     * of 6; rows past the level are halo or zero fill); only the store is predicated */
    static_assert(BORB_BLUR_TH % 6 == 0, "the row loop is unrolled by 6");
    for (int r0 = 0; r0 < rows; r0 += 6) {
#pragma unroll
        for (int u = 0; u < 6; ++u) {
            blur_row_sums_smem(src + (r0 + u + 6) * BLUR_BOX_W, hn);
            uint32_t V[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                Q[u][k] = __byte_perm(hp[k], hn[k], 0x5410);
                hp[k] = hn[k];
                V[k] = blur_col(Q[(u + 1) % 6][k], Q[(u + 3) % 6][k], Q[(u + 5) % 6][k], Q[u][k]);
            }
            /* V < 2^24: byte 2 of V is (V >> 16) */
            const uint32_t v = __byte_perm(__byte_perm(V[0], V[1], 0x0062), __byte_perm(V[2], V[3], 0x0062), 0x5410);
            if (r0 + u < rows) *reinterpret_cast<uint32_t *>(orow) = v;
            orow += pitch;
        }
    }
}

#endif
