/*
 * k_blur.cuh -- K5: 7x7 sigma-2 Gaussian of every pyramid level (src/ORBextractor.cc:1375-1377).
 *
 * OpenCV 4.x filters CV_8U with an 8.8 fixed-point kernel; for ksize 7, sigma 2 the quantised taps
 * are [18,34,48,56,48,34,18]/256 per axis (oracle/cvprim.c):
 *   row pass  H = sum k[i] * src            (<= 65280, 16 bit)
 *   col pass  V = sum k[j] * H              (32 bit),   dst = (V + 32768) >> 16
 * The reference blurs an un-padded clone with BORDER_REFLECT_101; the pyramid's own 19-pixel
 * frame holds exactly those reflected values, so the tile halo is read straight from it.
 *
 * One CTA per 64x16 output tile of any level (flat tile list): (16+6) x (64+6) source bytes are
 * staged in shared memory, the row pass writes 22 x 64 u16, the column pass emits 4 pixels per
 * thread as one 32-bit store.
 */
#ifndef K_BLUR_CUH
#define K_BLUR_CUH

#include "borb_internal.h"

#define BLUR_SP 72 /* staged row pitch, >= BORB_BLUR_TW + 6 */

__global__ void __launch_bounds__(256)
k_blur_tiles(const BorbPlan *__restrict__ plan, const BorbTile *__restrict__ tiles,
             const uint8_t *__restrict__ pyr, uint8_t *__restrict__ blur)
{
    __shared__ uint8_t src[(BORB_BLUR_TH + 6) * BLUR_SP];
    __shared__ uint16_t hp[(BORB_BLUR_TH + 6) * BORB_BLUR_TW];

    const BorbTile t = tiles[blockIdx.x];
    const BorbLevel &L = plan->lv[t.level];
    const size_t fo = (size_t)blockIdx.y * (size_t)plan->pyr_bytes + L.off;
    /* pointer to level pixel (0,0) */
    const uint8_t *in = pyr + fo + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    uint8_t *out = blur + fo + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    const int tid = threadIdx.x;

    /* stage (x0-3 .. x0+66) x (y0-3 .. y0+18); columns/rows past the frame are clamped (their
     * results are never stored) */
    for (int i = tid; i < (BORB_BLUR_TH + 6) * (BORB_BLUR_TW + 6); i += 256) {
        const int r = i / (BORB_BLUR_TW + 6), c = i - r * (BORB_BLUR_TW + 6);
        const int y = min(t.y0 + r - 3, L.h + BORB_EDGE - 1);
        const int x = min(t.x0 + c - 3, L.w + BORB_EDGE - 1);
        src[r * BLUR_SP + c] = in[(ptrdiff_t)y * L.pitch + x];
    }
    __syncthreads();
    for (int i = tid; i < (BORB_BLUR_TH + 6) * BORB_BLUR_TW; i += 256) {
        const int r = i / BORB_BLUR_TW, c = i - r * BORB_BLUR_TW;
        const uint8_t *q = src + r * BLUR_SP + c;
        hp[i] = (uint16_t)(18 * (q[0] + q[6]) + 34 * (q[1] + q[5]) + 48 * (q[2] + q[4]) + 56 * q[3]);
    }
    __syncthreads();
    {
        const int r = tid >> 4, c4 = (tid & 15) * 4; /* 16 rows x 16 groups of 4 */
        const int y = t.y0 + r, x = t.x0 + c4;
        if (y < L.h && x < L.w) {
            uint32_t v = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint16_t *h = hp + r * BORB_BLUR_TW + c4 + k;
                const uint32_t a = 18u * ((uint32_t)h[0] + h[6 * BORB_BLUR_TW]) +
                                   34u * ((uint32_t)h[BORB_BLUR_TW] + h[5 * BORB_BLUR_TW]) +
                                   48u * ((uint32_t)h[2 * BORB_BLUR_TW] + h[4 * BORB_BLUR_TW]) +
                                   56u * (uint32_t)h[3 * BORB_BLUR_TW];
                v |= ((a + 32768u) >> 16) << (8 * k);
            }
            /* x is a multiple of 4 and the interior starts 32-byte aligned: aligned store; the
             * up-to-3 bytes past the last column land in our own pitch padding */
            *reinterpret_cast<uint32_t *>(out + (size_t)y * L.pitch + x) = v;
        }
    }
}

#endif
