/*
 * k_pyramid.cuh -- K1: scale pyramid level = bilinear resize of the previous level + reflect-101
 * frame, one kernel per level (the level chain is serial: src/ORBextractor.cc:1400-1443 resizes
 * level l from level l-1, not from the input).
 *
 * Arithmetic = OpenCV 4.x cv::resize(INTER_LINEAR) for CV_8UC1 (oracle/cvprim.c):
 *   horizontal  r  = s[x0]*a0 + s[x0+1]*a1            (11-bit weights, int32)
 *   vertical    d  = (((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2
 * The tap tables depend only on (cols, rows) and are built on the host once per geometry.
 *
 * Every thread produces 4 horizontally adjacent bytes of the PADDED row (one aligned 32-bit
 * store); frame pixels are computed at their reflected interior coordinate, so no second pass
 * and no intra-level synchronisation is needed.
 */
#ifndef K_PYRAMID_CUH
#define K_PYRAMID_CUH

#include "borb_internal.h"

__device__ __forceinline__ int reflect101(int i, int n)
{
    /* |i| excursions are at most BORB_XPAD < n here */
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

/* level 0: copy of the input with the frame (src/ORBextractor.cc:1436-1439) */
__global__ void __launch_bounds__(256)
k_pyr_level0(const BorbPlan *__restrict__ plan, const uint8_t *__restrict__ img, size_t step,
             size_t frame_stride, uint8_t *__restrict__ pyr)
{
    const BorbLevel &L = plan->lv[0];
    const int c4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; /* buffer column */
    const int r = blockIdx.y;                                   /* padded row    */
    if (c4 >= L.pitch) return;
    const uint8_t *src = img + (size_t)blockIdx.z * frame_stride;
    uint8_t *dst = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes + L.off;
    const int y = reflect101(r - BORB_EDGE, L.h);
    const uint8_t *srow = src + (size_t)y * step;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = c4 + k - BORB_XPAD;
        x = max(-BORB_XPAD, min(x, L.w - 1 + BORB_XPAD)); /* columns beyond the frame: don't care */
        x = reflect101(x, L.w);
        v |= (uint32_t)__ldg(srow + x) << (8 * k);
    }
    *reinterpret_cast<uint32_t *>(dst + (size_t)r * L.pitch + c4) = v;
}

/* level l > 0 from level l-1 */
__global__ void __launch_bounds__(256)
k_pyr_resize(const BorbPlan *__restrict__ plan, int l, const int *__restrict__ tab_i,
             const short *__restrict__ tab_s, uint8_t *__restrict__ pyr)
{
    const BorbLevel &L = plan->lv[l];
    const BorbLevel &S = plan->lv[l - 1];
    const int c4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int r = blockIdx.y;
    if (c4 >= L.pitch) return;
    uint8_t *frame = pyr + (size_t)blockIdx.z * (size_t)plan->pyr_bytes;
    const uint8_t *sbase = frame + S.off + (size_t)BORB_EDGE * S.pitch + BORB_XPAD;
    uint8_t *dst = frame + L.off;

    const int y = reflect101(r - BORB_EDGE, L.h);
    const int sy = tab_i[L.yofs_off + y];
    const int b0 = tab_s[L.yb_off + 2 * y], b1 = tab_s[L.yb_off + 2 * y + 1];
    const int sy0 = min(max(sy, 0), S.h - 1), sy1 = min(max(sy + 1, 0), S.h - 1);
    const uint8_t *row0 = sbase + (size_t)sy0 * S.pitch;
    const uint8_t *row1 = sbase + (size_t)sy1 * S.pitch;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = c4 + k - BORB_XPAD;
        x = max(-BORB_XPAD, min(x, L.w - 1 + BORB_XPAD));
        x = reflect101(x, L.w);
        const int sx = tab_i[L.xofs_off + x];
        const int sx1 = min(sx + 1, S.w - 1);
        const int a0 = tab_s[L.xa_off + 2 * x], a1 = tab_s[L.xa_off + 2 * x + 1];
        const int r0 = row0[sx] * a0 + row0[sx1] * a1;
        const int r1 = row1[sx] * a0 + row1[sx1] * a1;
        int d = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
        d = min(max(d, 0), 255);
        v |= (uint32_t)d << (8 * k);
    }
    *reinterpret_cast<uint32_t *>(dst + (size_t)r * L.pitch + c4) = v;
}

#endif
