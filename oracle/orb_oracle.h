/*
 * oracle/orb_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C-ABI of the CPU oracle: a from-scratch restatement of the reference's ORB extraction path
 * (/root/reference/src/ORBextractor.cc:83-157, 432-951, 966-1128, 1310-1443) on flat arrays,
 * with the canonical quadtree tie-break.  It is pinned two ways:
 *   - its five image primitives (cvprim.c) bit-exact against the cv2 4.13.0 wheel;
 *   - its end-to-end output bit-exact (incl. order) against the reference's own ORBextractor.cc
 *     compiled in place (oracle/_ref/libreforb_mono.so) on seeded frames.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it.
 */
#ifndef ORACLE_ORB_ORACLE_H
#define ORACLE_ORB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbo_kp { /* == cv::KeyPoint, 28 bytes */
    float x, y, size, angle, response;
    int octave, class_id;
} orbo_kp;

typedef struct orbo_cand { /* one FAST candidate, coords relative to (16,16) of its level */
    int x, y, score;
} orbo_cand;

void *orbo_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
void orbo_destroy(void *h);

/* whole path; returns keypoint count (written: min(n, cap)), -1 for the empty-image early
 * return, -2 for geometry the reference cannot handle (nIni == 0 -> division by zero). */
int orbo_extract(void *h, const uint8_t *img, int cols, int rows, size_t step,
                 orbo_kp *kps, uint8_t *desc, int cap);

/* constructor-derived tables */
int orbo_levels(void *h);
void orbo_table(void *h, int which, float *out); /* 0 scale, 1 inv scale, 2 sigma2, 3 inv sigma2 */
void orbo_quotas(void *h, int *out);
void orbo_umax(void *h, int *out);               /* 16 ints */

/* per-stage results of the last orbo_extract call */
int orbo_level_size(void *h, int level, int *cols, int *rows);
int orbo_pyramid_copy(void *h, int level, int border, uint8_t *out); /* (cols+2b)*(rows+2b) */
int orbo_blurred_copy(void *h, int level, uint8_t *out);             /* cols*rows */
int orbo_candidates(void *h, int level, orbo_cand *out, int cap);    /* returns count */
int orbo_level_keypoints(void *h, int level, orbo_kp *out, int cap); /* level coords, with angle */

/* stand-alone quadtree distribution (src/ORBextractor.cc:668-951) for adversarial unit tests:
 * candidates (xs, ys, resp) in input order; writes the indices of the survivors in output order;
 * returns their count, or -2 when nIni == 0. */
int orbo_distribute(const int *xs, const int *ys, const int *resp, int n,
                    int min_x, int max_x, int min_y, int max_y, int n_target,
                    int *out_idx, int cap);

/* stand-alone rotated-BRIEF of one keypoint on a (blurred) plane; angle in degrees */
void orbo_describe(const uint8_t *img, size_t step, int x, int y, float angle_deg, uint8_t *desc32);
/* stand-alone intensity-centroid angle */
float orbo_ic_angle(const uint8_t *img, size_t step, int x, int y);
/* the 512 pattern points as (x,y) int8 pairs -- derived table, see orb_pattern.inc */
const signed char *orbo_pattern(void);

#ifdef __cplusplus
}
#endif
#endif
