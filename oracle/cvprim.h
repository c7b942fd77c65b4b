/*
 * oracle/cvprim.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatements (plain C, raw buffers) of the five OpenCV primitives the reference's
 * ORB extractor calls.  OpenCV is an un-vendored, un-pinned dependency of the reference
 * (/root/reference/CMakeLists.txt:33-39); the anchor adopted here is OpenCV 4.13.0 as shipped in
 * the `cv2` wheel, and every function below is checked bit-exact against that wheel by
 * tests/test_oracle_primitives.py.
 *
 * Reference call sites (file:line relative to /root/reference):
 *   cvp_resize_linear_u8   <- cv::resize(..., INTER_LINEAR)            src/ORBextractor.cc:1429
 *   cvp_border_reflect101  <- cv::copyMakeBorder(BORDER_REFLECT_101)   src/ORBextractor.cc:1432,1438
 *   cvp_fast9_16           <- cv::FAST(img, kps, th, true)             src/ORBextractor.cc:1069,1074
 *   cvp_gaussian7_sigma2   <- cv::GaussianBlur(7x7, 2, 2, REFLECT_101) src/ORBextractor.cc:1377
 *   cvp_fast_atan2         <- cv::fastAtan2                            src/ORBextractor.cc:110
 *   cvp_round_f / _d       <- cvRound                                  src/ORBextractor.cc:87,124,129,472
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may use anything in
 * oracle/.  The shipped CUDA path never links or calls it.
 */
#ifndef ORACLE_CVPRIM_H
#define ORACLE_CVPRIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* keypoint record produced by cvp_fast9_16: window-relative integer coords + score */
typedef struct cvp_kp {
    int x, y;
    int score;
} cvp_kp;

/* round-half-to-even, as cvRound compiles to cvtss2si / cvtsd2si */
int cvp_round_f(float v);
int cvp_round_d(double v);

/* Build the per-axis tap tables of the 8-bit bilinear resize: for d in [0,dn):
 * ofs[d] = first source index, coef[2d], coef[2d+1] = 11-bit fixed-point weights. */
void cvp_resize_axis_table(int sn, int dn, int *ofs, short *coef);

/* dst(dw x dh) = bilinear resize of src(sw x sh), 8-bit single channel. */
void cvp_resize_linear_u8(const uint8_t *src, size_t sstep, int sw, int sh,
                          uint8_t *dst, size_t dstep, int dw, int dh);

/* Fill the `b`-pixel frame around the w x h interior located at buf + b*step + b with the
 * reflect-101 extension of that interior (interior itself untouched). */
void cvp_border_reflect101(uint8_t *buf, size_t step, int w, int h, int b);

/* FAST-9/16 with 3x3 non-maximum suppression on a w x h window.  Writes up to `cap` keypoints
 * in row-major order and returns the number found (may exceed cap; only cap are written). */
int cvp_fast9_16(const uint8_t *img, size_t step, int w, int h, int threshold,
                 cvp_kp *out, int cap);

/* corner score of one pixel (max threshold for which it is still a FAST-9 corner), given that
 * the result is only meaningful when >= threshold; exposed for unit tests. */
int cvp_fast_corner_score(const uint8_t *p, size_t step, int threshold);

/* 7x7 Gaussian, sigma 2, fixed-point kernel [18,34,48,56,48,34,18]/256 per axis,
 * reflect-101 borders.  src and dst may alias. */
void cvp_gaussian7_sigma2(const uint8_t *src, size_t sstep, uint8_t *dst, size_t dstep,
                          int w, int h);

/* degrees in [0,360), scalar float32 polynomial, every operation individually rounded */
float cvp_fast_atan2(float y, float x);

/* -- the step before the extractor (oracle/cvprim_pre.c) -------------------- */

/* cv::cvtColor to gray for 8-bit 1/3/4-channel rows; rgb_order != 0: channel 0 is R (CV_RGB2GRAY,
 * CV_RGBA2GRAY), else channel 0 is B.  One channel: copy. */
void cvp_cvt_gray(const uint8_t *src, size_t sstep, int cols, int rows, int channels, int rgb_order,
                  uint8_t *dst, size_t dstep);

/* cv::remap(src, dst, map_x, map_y, INTER_LINEAR), CV_32FC1 maps, BORDER_CONSTANT 0, 8-bit with
 * `channels` interleaved channels.  map_step in floats. */
void cvp_remap_linear(const uint8_t *src, size_t sstep, int scols, int srows, int channels,
                      const float *map_x, const float *map_y, size_t map_step,
                      uint8_t *dst, size_t dstep, int dcols, int drows);

#ifdef __cplusplus
}
#endif
#endif
