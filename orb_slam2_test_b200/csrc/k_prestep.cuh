/*
 * k_prestep.cuh -- K0: the step before the extractor, fused into one pass:
 *   cv::remap(im, rect, M1, M2, INTER_LINEAR)      Examples/Stereo/stereo_euroc.cc:136-137
 *   cv::cvtColor(im, im, CV_{RGB,BGR,RGBA,BGRA}2GRAY)   src/Tracking.cc:176-201, :213-227, :245-259
 * in that order (the drivers rectify the raw frame, Tracking::GrabImage* then converts it), so raw
 * camera frames are uploaded once and the gray / rectified image the reference hands to
 * ORBextractor::operator() is produced on the device (it becomes pyramid level 0).
 *
 * Arithmetic = OpenCV 4.13.0 (oracle/cvprim_pre.c, pinned against the cv2 wheel):
 *   gray   Y = (B*3735 + G*19235 + R*9798 + 2^14) >> 15
 *   remap  coordinates quantised on upload to 1/32 px: s = cvRound(map*32); tap origin s>>5
 *          (int16), fractions s&31; weights (32-fx|fx)*(32-fy|fy)*32; taps outside the source
 *          read 0 (BORDER_CONSTANT); dst = (sum + 2^14) >> 15, per channel.
 * The map is stored as one u32 (x | y<<16, int16 each) and one u16 (fy<<5 | fx) per output pixel.
 *
 * Traffic per output pixel: C source bytes (x4 taps from L1/L2 when remapping) + 6 map bytes when
 * remapping, 1 byte written: HBM bound.  A thread produces 4 adjacent pixels = one aligned word.
 */
#ifndef K_PRESTEP_CUH
#define K_PRESTEP_CUH

#include <stdint.h>

#define PRE_NT 128
#define PRE_PX (4 * PRE_NT) /* pixels of a row per CTA */

__device__ __forceinline__ uint32_t pre_gray(uint32_t c0, uint32_t c1, uint32_t c2, bool rgb)
{
    const uint32_t b = rgb ? c2 : c0, r = rgb ? c0 : c2;
    return (b * 3735u + c1 * 19235u + r * 9798u + (1u << 14)) >> 15;
}

/* colour -> gray without a map: the row segment is staged through shared memory as aligned words */
template <int C>
__global__ void __launch_bounds__(PRE_NT)
k_pre_gray(const uint8_t *__restrict__ src, size_t step, size_t stride, int cols, int rgb,
           uint8_t *__restrict__ dst, int dpitch, size_t dstride)
{
    __shared__ uint32_t seg[(PRE_PX * C + 8) / 4 + 1];
    const int x0 = blockIdx.x * PRE_PX, y = blockIdx.y, f = blockIdx.z;
    const int npx = min(PRE_PX, cols - x0);
    const uint8_t *row = src + (size_t)f * stride + (size_t)y * step + (size_t)x0 * C;
    const int a = (int)((uintptr_t)row & 3);
    const uint32_t *wsrc = reinterpret_cast<const uint32_t *>(row - a);
    const int nwords = (npx * C + a + 3) >> 2;
    for (int i = threadIdx.x; i < nwords; i += PRE_NT) seg[i] = __ldg(wsrc + i);
    __syncthreads();
    const int px = 4 * threadIdx.x;
    if (px >= npx) return;
    const uint8_t *s = reinterpret_cast<const uint8_t *>(seg) + a + px * C;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (px + k < npx) out |= pre_gray(s[k * C], s[k * C + 1], s[k * C + 2], rgb != 0) << (8 * k);
    *reinterpret_cast<uint32_t *>(dst + (size_t)f * dstride + (size_t)y * dpitch + x0 + px) = out;
}

/* remap (+ gray when C > 1) */
template <int C>
__global__ void __launch_bounds__(PRE_NT)
k_pre_remap(const uint8_t *__restrict__ src, size_t step, size_t stride, int scols, int srows,
            const uint32_t *__restrict__ map_xy, const uint16_t *__restrict__ map_fr, int mpitch,
            int rgb, uint8_t *__restrict__ dst, int dpitch, size_t dstride, int dcols)
{
    const int px = blockIdx.x * PRE_PX + 4 * threadIdx.x, y = blockIdx.y, f = blockIdx.z;
    if (px >= dcols) return;
    const uint8_t *img = src + (size_t)f * stride;
    /* mpitch is a multiple of 4: one 16-byte and one 8-byte load bring the 4 pixels' coordinates */
    const uint4 xy4 = __ldg(reinterpret_cast<const uint4 *>(map_xy + (size_t)y * mpitch + px));
    const uint2 fr4 = __ldg(reinterpret_cast<const uint2 *>(map_fr + (size_t)y * mpitch + px));
    const uint32_t xy[4] = {xy4.x, xy4.y, xy4.z, xy4.w};
    const uint32_t fr[4] = {fr4.x & 0xffffu, fr4.x >> 16, fr4.y & 0xffffu, fr4.y >> 16};
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (px + k >= dcols) break;
        const int ix = (int)(short)(xy[k] & 0xffffu), iy = (int)(short)(xy[k] >> 16);
        const uint32_t fx = fr[k] & 31u, fy = fr[k] >> 5;
        const uint32_t w00 = (32u - fx) * (32u - fy) * 32u, w01 = fx * (32u - fy) * 32u;
        const uint32_t w10 = (32u - fx) * fy * 32u, w11 = fx * fy * 32u;
        const bool x0 = (unsigned)ix < (unsigned)scols, x1 = (unsigned)(ix + 1) < (unsigned)scols;
        const bool y0 = (unsigned)iy < (unsigned)srows, y1 = (unsigned)(iy + 1) < (unsigned)srows;
        const uint8_t *p = img + (ptrdiff_t)iy * (ptrdiff_t)step + (ptrdiff_t)ix * C;
        uint32_t v[C > 1 ? 3 : 1];
#pragma unroll
        for (int c = 0; c < (C > 1 ? 3 : 1); ++c) {
            uint32_t acc = 0;
            if (y0 && x0) acc += __ldg(p + c) * w00;
            if (y0 && x1) acc += __ldg(p + C + c) * w01;
            if (y1 && x0) acc += __ldg(p + step + c) * w10;
            if (y1 && x1) acc += __ldg(p + step + C + c) * w11;
            v[c] = (acc + (1u << 14)) >> 15;
        }
        out |= (C > 1 ? pre_gray(v[0], v[C > 1 ? 1 : 0], v[C > 1 ? 2 : 0], rgb != 0) : v[0]) << (8 * k);
    }
    *reinterpret_cast<uint32_t *>(dst + (size_t)f * dstride + (size_t)y * dpitch + px) = out;
}

#endif
