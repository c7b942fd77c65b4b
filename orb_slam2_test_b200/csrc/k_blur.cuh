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
 * v2: no shared memory.  A warp owns a 128-column x BORB_BLUR_TH-row strip; a thread owns 4 adjacent
 * columns (one aligned 32-bit word of every row).  Walking down the strip it loads three words per
 * row (its own and the two neighbours': L1 hits), forms the 4 row sums with 8 byte-dot-products
 * (dp4a with the taps packed as bytes: 18,34,48,56 | 48,34,18,0) and keeps the last 7 rows of row
 * sums in registers (the loop is unrolled by 7 so the window is statically indexed); each new row
 * yields one output word.  Algorithmic traffic: read 1 B + write 1 B per pixel.
 */
#ifndef K_BLUR_CUH
#define K_BLUR_CUH

#include "borb_internal.h"

#define BLUR_NT 128 /* 4 warps = 4 consecutive strips of the tile list */
#ifndef BLUR_PF
#define BLUR_PF 2 /* rows prefetched ahead */
#endif

/* words at columns c-4, c, c+4 of a row (row is 4-byte aligned at column 0; c is a multiple of 4);
 * c2 = min(c + 4, last readable word column of the padded row) */
__device__ __forceinline__ uint3 blur_load(const uint8_t *__restrict__ row, int c, int c2)
{
    return make_uint3(__ldg(reinterpret_cast<const uint32_t *>(row + c - 4)),
                      __ldg(reinterpret_cast<const uint32_t *>(row + c)),
                      __ldg(reinterpret_cast<const uint32_t *>(row + c2)));
}

__device__ __forceinline__ void blur_row_sums(const uint3 w, uint32_t h[4])
{
    const uint32_t w0 = w.x, w1 = w.y, w2 = w.z;
    const uint32_t klo = 18u | (34u << 8) | (48u << 16) | (56u << 24); /* taps 0..3 */
    const uint32_t khi = 48u | (34u << 8) | (18u << 16);               /* taps 4..6 */
    /* output column c+i needs bytes c+i-3 .. c+i+3 */
    h[0] = __dp4a(__byte_perm(w0, w1, 0x4321), klo, __dp4a(__byte_perm(w1, w2, 0x4321), khi, 0u));
    h[1] = __dp4a(__byte_perm(w0, w1, 0x5432), klo, __dp4a(__byte_perm(w1, w2, 0x5432), khi, 0u));
    h[2] = __dp4a(__byte_perm(w0, w1, 0x6543), klo, __dp4a(__byte_perm(w1, w2, 0x6543), khi, 0u));
    h[3] = __dp4a(w1, klo, __dp4a(w2, khi, 0u));
}

__device__ __forceinline__ uint32_t blur_col(const uint32_t a, const uint32_t b, const uint32_t c, const uint32_t d,
                                             const uint32_t e, const uint32_t f, const uint32_t g)
{
    return (18u * (a + g) + 34u * (b + f) + 48u * (c + e) + 56u * d + 32768u) >> 16;
}

__global__ void __launch_bounds__(BLUR_NT)
k_blur_strips(const __grid_constant__ BorbPlan plan_c, const BorbTile *__restrict__ tiles,
              const uint8_t *__restrict__ pyr, uint8_t *__restrict__ blur)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    const int lane = threadIdx.x & 31;
    const int t_idx = blockIdx.x * (BLUR_NT / 32) + (threadIdx.x >> 5);
    if (t_idx >= plan->n_tiles) return;
    const BorbTile t = tiles[t_idx];
    const BorbLevel &L = plan->lv[t.level];
    const int c = t.x0 + 4 * lane;
    if (c >= L.w) return;
    const size_t fo = (size_t)blockIdx.y * (size_t)plan->pyr_bytes + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    const uint8_t *in = pyr + fo;  /* level pixel (0,0) */
    uint8_t *out = blur + fo;
    const int cmax = L.pitch - BORB_XPAD - 4;
    const int rows = min(BORB_BLUR_TH, L.h - t.y0);

    const int c2 = min(c + 4, cmax);
    uint32_t H[7][4];
#pragma unroll
    for (int r = 0; r < 6; ++r) blur_row_sums(blur_load(in + (ptrdiff_t)(t.y0 - 3 + r) * L.pitch, c, c2), H[r]);

    /* the words of the BLUR_PF rows after the one being filtered are already in flight (software
     * prefetch: the kernel is bound by load latency, not by bandwidth) */
    uint3 q[BLUR_PF]; /* q[i] = row y0 + 3 + r + i when output row r is next */
#pragma unroll
    for (int i = 0; i < BLUR_PF; ++i) q[i] = blur_load(in + (ptrdiff_t)(t.y0 + 3 + min(i, rows - 1)) * L.pitch, c, c2);
    const uint8_t *nrow = in + (ptrdiff_t)(t.y0 + 3 + BLUR_PF) * L.pitch;
    for (int r0 = 0; r0 < rows; r0 += 7) {
#pragma unroll
        for (int u = 0; u < 7; ++u) {
            const int r = r0 + u;
            if (r < rows) {
                /* the row entering the window is y0 + r + 3, into slot (6 + u) % 7; the window for
                 * output row r is slots u, u+1, ..., u+6 (mod 7) */
                const uint3 w = q[0];
#pragma unroll
                for (int i = 0; i + 1 < BLUR_PF; ++i) q[i] = q[i + 1];
                if (r + BLUR_PF < rows) q[BLUR_PF - 1] = blur_load(nrow, c, c2);
                nrow += L.pitch;
                blur_row_sums(w, H[(6 + u) % 7]);
                uint32_t v = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    v |= blur_col(H[u % 7][k], H[(u + 1) % 7][k], H[(u + 2) % 7][k], H[(u + 3) % 7][k],
                                  H[(u + 4) % 7][k], H[(u + 5) % 7][k], H[(u + 6) % 7][k]) << (8 * k);
                /* c is a multiple of 4 and the interior starts 32-byte aligned: aligned store; the
                 * up-to-3 bytes past the last column land in our own pitch padding */
                *reinterpret_cast<uint32_t *>(out + (size_t)(t.y0 + r) * L.pitch + c) = v;
            }
        }
    }
}

#endif
