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
 * intra-level synchronisation).  v2: a thread owns 4 adjacent columns (one aligned 32-bit word of
 * the padded row) over a strip of PYR_RS rows: its horizontal taps live in registers for the
 * whole strip, the row taps are warp-uniform, and the horizontal sums of a source row are reused
 * when the next output row needs the same source row (5 times out of 6 at scale 1.2).
 */
#ifndef K_PYRAMID_CUH
#define K_PYRAMID_CUH

#include "borb_internal.h"

#define PYR_NT 128
#define PYR_RS 8       /* output rows per thread */
#define PYR_LPAD 20    /* columns computed left of the image: 19-px frame rounded up to a word */

__device__ __forceinline__ int reflect101(int i, int n)
{
    /* |i| excursions are at most BORB_XPAD < n here */
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

__host__ __device__ __forceinline__ int pyr_words_per_row(int w) { return (w + BORB_EDGE + PYR_LPAD + 3) >> 2; }

/* level 0: copy of the input with the frame (src/ORBextractor.cc:1436-1439); one word per thread */
__global__ void __launch_bounds__(PYR_NT)
k_pyr_level0(const BorbPlan *__restrict__ plan, const uint8_t *__restrict__ img, size_t step,
             size_t frame_stride, uint8_t *__restrict__ pyr)
{
    const BorbLevel &L = plan->lv[0];
    const int wpr = pyr_words_per_row(L.w);
    const int idx = blockIdx.x * PYR_NT + threadIdx.x;
    const int r = idx / wpr;                      /* padded row */
    if (r >= L.prows) return;
    const int c = (idx - r * wpr) * 4 - PYR_LPAD; /* first of 4 level columns */
    const uint8_t *src = img + (size_t)blockIdx.z * frame_stride;
    uint8_t *dst = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes + L.off;
    const int y = reflect101(r - BORB_EDGE, L.h);
    const uint8_t *srow = src + (size_t)y * step;
    uint32_t v = 0;
    if (c >= 0 && c + 3 < L.w && y + 1 < L.h) {
        /* interior word: 4 contiguous source bytes at arbitrary alignment (the input's row step
         * need not be a multiple of 4) from two aligned words */
        const uintptr_t a = reinterpret_cast<uintptr_t>(srow + c);
        const uint32_t *base = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
        const uint32_t lo = __ldg(base), hi = (a & 3) ? __ldg(base + 1) : 0u;
        v = __funnelshift_r(lo, hi, (uint32_t)(a & 3) * 8u);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int x = min(c + k, L.w - 1 + BORB_EDGE); /* columns beyond the frame: don't care */
            x = reflect101(x, L.w);
            v |= (uint32_t)__ldg(srow + x) << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t *>(dst + (size_t)r * L.pitch + BORB_XPAD + c) = v;
}

__device__ __forceinline__ void pyr_hsum(const uint8_t *__restrict__ row, const int sx[4], const int sx1[4],
                                         const int a0[4], const int a1[4], int h[4])
{
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = (int)row[sx[k]] * a0[k] + (int)row[sx1[k]] * a1[k];
}

/* level l > 0 from level l-1 */
__global__ void __launch_bounds__(PYR_NT)
k_pyr_resize(const BorbPlan *__restrict__ plan, int l, const int *__restrict__ tab_i,
             const short *__restrict__ tab_s, uint8_t *__restrict__ pyr)
{
    const BorbLevel &L = plan->lv[l];
    const BorbLevel &S = plan->lv[l - 1];
    const int wpr = pyr_words_per_row(L.w);
    const int idx = blockIdx.x * PYR_NT + threadIdx.x;
    const int strip = idx / wpr;
    const int r_begin = strip * PYR_RS;
    if (r_begin >= L.prows) return;
    const int c = (idx - strip * wpr) * 4 - PYR_LPAD;
    uint8_t *frame = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes;
    const uint8_t *sbase = frame + S.off + (size_t)BORB_EDGE * S.pitch + BORB_XPAD;
    uint8_t *dst = frame + L.off + BORB_XPAD + c;

    /* horizontal taps of this thread's 4 columns */
    int sx[4], sx1[4], a0[4], a1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = min(c + k, L.w - 1 + BORB_EDGE);
        x = reflect101(x, L.w);
        sx[k] = tab_i[L.xofs_off + x];
        sx1[k] = min(sx[k] + 1, S.w - 1);
        a0[k] = tab_s[L.xa_off + 2 * x];
        a1[k] = tab_s[L.xa_off + 2 * x + 1];
    }

    int rowA = -1, rowB = -1;
    int hA[4] = {0, 0, 0, 0}, hB[4] = {0, 0, 0, 0};
    const int r_end = min(r_begin + PYR_RS, L.prows);
    for (int r = r_begin; r < r_end; ++r) {
        const int y = reflect101(r - BORB_EDGE, L.h);
        const int sy = tab_i[L.yofs_off + y];
        const int b0 = tab_s[L.yb_off + 2 * y], b1 = tab_s[L.yb_off + 2 * y + 1];
        const int sy0 = min(max(sy, 0), S.h - 1), sy1 = min(max(sy + 1, 0), S.h - 1);
        int n0[4], n1[4];
        if (sy0 == rowB) {
#pragma unroll
            for (int k = 0; k < 4; ++k) n0[k] = hB[k];
        } else if (sy0 == rowA) {
#pragma unroll
            for (int k = 0; k < 4; ++k) n0[k] = hA[k];
        } else {
            pyr_hsum(sbase + (size_t)sy0 * S.pitch, sx, sx1, a0, a1, n0);
        }
        if (sy1 == rowB) {
#pragma unroll
            for (int k = 0; k < 4; ++k) n1[k] = hB[k];
        } else if (sy1 == rowA) {
#pragma unroll
            for (int k = 0; k < 4; ++k) n1[k] = hA[k];
        } else {
            pyr_hsum(sbase + (size_t)sy1 * S.pitch, sx, sx1, a0, a1, n1);
        }
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int d = (((b0 * (n0[k] >> 4)) >> 16) + ((b1 * (n1[k] >> 4)) >> 16) + 2) >> 2;
            v |= (uint32_t)d << (8 * k);
            hA[k] = n0[k];
            hB[k] = n1[k];
        }
        rowA = sy0;
        rowB = sy1;
        *reinterpret_cast<uint32_t *>(dst + (size_t)r * L.pitch) = v;
    }
}

#endif
