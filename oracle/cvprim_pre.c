/*
 * oracle/cvprim_pre.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatements of the two OpenCV primitives of the step BEFORE the extractor (SURVEY.md
 * 8(f)-2), bit-exact against the cv2 4.13.0 wheel (tests/test_oracle_primitives.py):
 *   cvp_cvt_gray      <- cv::cvtColor(im, im, CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY)
 *                        src/Tracking.cc:176-201 (and :213-227, :245-259)
 *   cvp_remap_linear  <- cv::remap(im, rect, M1, M2, cv::INTER_LINEAR) with CV_32F maps
 *                        Examples/Stereo/stereo_euroc.cc:136-137
 *
 * cvtColor, 8-bit: Y = (B*3735 + G*19235 + R*9798 + 2^14) >> 15 (15-bit coefficients, exhaustively
 * equal to the wheel over all 2^24 colours).
 * remap, 8-bit, float maps, INTER_LINEAR, BORDER_CONSTANT(0): coordinates are first quantised to
 * 1/32 pixel, s = cvRound(map * 32) (float product, round half to even); integer part s >> 5
 * saturated to int16, fraction s & 31; the four weights are (32-fx|fx)*(32-fy|fy)*32 (sum 2^15,
 * always exact so the table's sum correction never fires); taps outside the source read 0;
 * dst = (sum + 2^14) >> 15.  Multi-channel images are interpolated per channel.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#include "cvprim.h"

void cvp_cvt_gray(const uint8_t *src, size_t sstep, int cols, int rows, int channels, int rgb_order,
                  uint8_t *dst, size_t dstep)
{
    const int ib = rgb_order ? 2 : 0, ir = rgb_order ? 0 : 2;
    for (int y = 0; y < rows; ++y) {
        const uint8_t *s = src + (size_t)y * sstep;
        uint8_t *d = dst + (size_t)y * dstep;
        for (int x = 0; x < cols; ++x, s += channels)
            d[x] = channels == 1 ? s[0]
                                 : (uint8_t)((s[ib] * 3735 + s[1] * 19235 + s[ir] * 9798 + (1 << 14)) >> 15);
    }
}

static inline int sat16(int v) { return v < -32768 ? -32768 : v > 32767 ? 32767 : v; }

void cvp_remap_linear(const uint8_t *src, size_t sstep, int scols, int srows, int channels,
                      const float *map_x, const float *map_y, size_t map_step /* floats per row */,
                      uint8_t *dst, size_t dstep, int dcols, int drows)
{
    for (int y = 0; y < drows; ++y) {
        const float *mx = map_x + (size_t)y * map_step, *my = map_y + (size_t)y * map_step;
        uint8_t *d = dst + (size_t)y * dstep;
        for (int x = 0; x < dcols; ++x) {
            const int sx = (int)lrintf(mx[x] * 32.f), sy = (int)lrintf(my[x] * 32.f);
            const int fx = sx & 31, fy = sy & 31;
            const int ix = sat16(sx >> 5), iy = sat16(sy >> 5);
            const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32;
            const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
            const int x0 = ix >= 0 && ix < scols, x1 = ix + 1 >= 0 && ix + 1 < scols;
            const int y0 = iy >= 0 && iy < srows, y1 = iy + 1 >= 0 && iy + 1 < srows;
            for (int c = 0; c < channels; ++c) {
                const uint8_t *p = src + (ptrdiff_t)iy * (ptrdiff_t)sstep + (ptrdiff_t)ix * channels + c;
                int acc = 0;
                if (y0 && x0) acc += p[0] * w00;
                if (y0 && x1) acc += p[channels] * w01;
                if (y1 && x0) acc += p[sstep] * w10;
                if (y1 && x1) acc += p[sstep + channels] * w11;
                d[x * channels + c] = (uint8_t)((acc + (1 << 14)) >> 15);
            }
        }
    }
}
