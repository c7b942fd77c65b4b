/*
 * oracle/cvprim.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See cvprim.h.
 *
 * Restates the published OpenCV 4.x algorithms for the five image primitives on the reference's
 * ORB-extraction path.  Nothing here is copied from the reference tree (which does not contain
 * OpenCV); the arithmetic is pinned by tests against the cv2 4.13.0 wheel.
 *
 * Build with -ffp-contract=off: cvp_fast_atan2 and the resize tap tables must round every
 * floating-point operation individually.
 */
#include "cvprim.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ rounding */

int cvp_round_f(float v) { return (int)lrintf(v); } /* default FE_TONEAREST = half-to-even */
int cvp_round_d(double v) { return (int)lrint(v); }

static inline int floor_f(float v)
{
    int i = (int)v;
    return i - (v < (float)i);
}

static inline short sat_s16(int v) { return (short)(v < -32768 ? -32768 : v > 32767 ? 32767 : v); }

/* ------------------------------------------------------------------ resize
 * cv::resize INTER_LINEAR, CV_8UC1: taps from a double-precision centre-aligned map cast to
 * float, weights quantised to 11 bits; horizontal pass keeps 19-bit products, vertical pass is
 * the 8-bit special case ((b*(r>>4))>>16 summed, +2, >>2). */

void cvp_resize_axis_table(int sn, int dn, int *ofs, short *coef)
{
    const double scale = 1.0 / ((double)dn / (double)sn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = floor_f(f);
        f -= (float)s;
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= sn - 1) { f = 0.f; s = sn - 1; }
        ofs[d] = s;
        coef[2 * d] = sat_s16(cvp_round_f((1.f - f) * 2048.f));
        coef[2 * d + 1] = sat_s16(cvp_round_f(f * 2048.f));
    }
}

/* vertical table: OpenCV does not clamp the weights, it clips the ROW INDEX instead */
static void resize_row_table(int sn, int dn, int *ofs, short *coef)
{
    const double scale = 1.0 / ((double)dn / (double)sn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = floor_f(f);
        f -= (float)s;
        ofs[d] = s;
        coef[2 * d] = sat_s16(cvp_round_f((1.f - f) * 2048.f));
        coef[2 * d + 1] = sat_s16(cvp_round_f(f * 2048.f));
    }
}

static inline int clipi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

void cvp_resize_linear_u8(const uint8_t *src, size_t sstep, int sw, int sh,
                          uint8_t *dst, size_t dstep, int dw, int dh)
{
    int *xofs = (int *)malloc(sizeof(int) * (size_t)(dw + dh));
    int *yofs = xofs + dw;
    short *xa = (short *)malloc(sizeof(short) * 2 * (size_t)(dw + dh));
    short *yb = xa + 2 * dw;
    int *rows = (int *)malloc(sizeof(int) * 2 * (size_t)dw);
    int *r0 = rows, *r1 = rows + dw;
    int have0 = -1, have1 = -1; /* which source rows r0/r1 currently hold */

    cvp_resize_axis_table(sw, dw, xofs, xa);
    resize_row_table(sh, dh, yofs, yb);

    for (int dy = 0; dy < dh; ++dy) {
        const int sy0 = clipi(yofs[dy], 0, sh - 1);
        const int sy1 = clipi(yofs[dy] + 1, 0, sh - 1);
        /* reuse the previous iteration's second row when the window slides by one */
        if (have1 == sy0 && have0 != sy0) {
            int *t = r0; r0 = r1; r1 = t;
            have0 = have1; have1 = -1;
        }
        if (have0 != sy0) {
            const uint8_t *s = src + (size_t)sy0 * sstep;
            for (int dx = 0; dx < dw; ++dx) {
                const int sx = xofs[dx];
                const int sx1 = sx + 1 < sw ? sx + 1 : sx;
                r0[dx] = s[sx] * xa[2 * dx] + s[sx1] * xa[2 * dx + 1];
            }
            have0 = sy0;
        }
        if (sy1 == sy0) {
            memcpy(r1, r0, sizeof(int) * (size_t)dw);
            have1 = sy1;
        } else if (have1 != sy1) {
            const uint8_t *s = src + (size_t)sy1 * sstep;
            for (int dx = 0; dx < dw; ++dx) {
                const int sx = xofs[dx];
                const int sx1 = sx + 1 < sw ? sx + 1 : sx;
                r1[dx] = s[sx] * xa[2 * dx] + s[sx1] * xa[2 * dx + 1];
            }
            have1 = sy1;
        }
        const int b0 = yb[2 * dy], b1 = yb[2 * dy + 1];
        uint8_t *d = dst + (size_t)dy * dstep;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            d[dx] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
    free(rows);
    free(xa);
    free(xofs);
}

/* ------------------------------------------------------------------ border */

static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * (n - 1) - i;
    }
    return i;
}

void cvp_border_reflect101(uint8_t *buf, size_t step, int w, int h, int b)
{
    uint8_t *in = buf + (size_t)b * step + b;
    /* left / right strips of the interior rows */
    for (int y = 0; y < h; ++y) {
        uint8_t *row = in + (size_t)y * step;
        for (int i = 1; i <= b; ++i) {
            row[-i] = row[reflect101(-i, w)];
            row[w - 1 + i] = row[reflect101(w - 1 + i, w)];
        }
    }
    /* top / bottom rows are copies of full (already widened) rows */
    for (int i = 1; i <= b; ++i) {
        memcpy(buf + (size_t)(b - i) * step, buf + (size_t)(b + reflect101(-i, h)) * step,
               (size_t)(w + 2 * b));
        memcpy(buf + (size_t)(b + h - 1 + i) * step,
               buf + (size_t)(b + reflect101(h - 1 + i, h)) * step, (size_t)(w + 2 * b));
    }
}

/* ------------------------------------------------------------------ FAST-9/16
 * Bresenham ring of radius 3, 16 pixels, listed clockwise from (0,3).  A pixel is a corner at
 * threshold t when 9 contiguous ring pixels are all darker than I-t or all brighter than I+t.
 * Its score is the largest t for which that still holds (OpenCV cornerScore<16>), computed for
 * corners only; 3x3 non-maximum suppression is strict and treats everything that is not a
 * corner, and the 3-pixel frame of the window, as score 0. */

static const int ring_dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int ring_dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

static int score_from_diffs(const int *d /* 25 entries, d[k+16]==d[k] */, int threshold)
{
    int a0 = threshold;
    for (int k = 0; k < 16; k += 2) {
        int a = imin(d[k + 1], d[k + 2]);
        a = imin(a, d[k + 3]);
        if (a <= a0) continue;
        a = imin(a, d[k + 4]);
        a = imin(a, d[k + 5]);
        a = imin(a, d[k + 6]);
        a = imin(a, d[k + 7]);
        a = imin(a, d[k + 8]);
        a0 = imax(a0, imin(a, d[k]));
        a0 = imax(a0, imin(a, d[k + 9]));
    }
    int b0 = -a0;
    for (int k = 0; k < 16; k += 2) {
        int b = imax(d[k + 1], d[k + 2]);
        b = imax(b, d[k + 3]);
        if (b >= b0) continue;
        b = imax(b, d[k + 4]);
        b = imax(b, d[k + 5]);
        b = imax(b, d[k + 6]);
        b = imax(b, d[k + 7]);
        b = imax(b, d[k + 8]);
        b0 = imin(b0, imax(b, d[k]));
        b0 = imin(b0, imax(b, d[k + 9]));
    }
    return -b0 - 1;
}

int cvp_fast_corner_score(const uint8_t *p, size_t step, int threshold)
{
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 25; ++k)
        d[k] = v - p[(ptrdiff_t)ring_dy[k & 15] * (ptrdiff_t)step + ring_dx[k & 15]];
    return score_from_diffs(d, threshold);
}

int cvp_fast9_16(const uint8_t *img, size_t step, int w, int h, int threshold,
                 cvp_kp *out, int cap)
{
    if (w < 7 || h < 7) return 0;
    threshold = clipi(threshold, 0, 255);

    ptrdiff_t off[25];
    for (int k = 0; k < 25; ++k)
        off[k] = (ptrdiff_t)ring_dy[k & 15] * (ptrdiff_t)step + ring_dx[k & 15];

    /* class table indexed by (ring - centre + 255): 1 = darker than I-t, 2 = brighter than I+t */
    uint8_t cls[512];
    for (int i = -255; i <= 255; ++i)
        cls[i + 255] = (uint8_t)(i < -threshold ? 1 : i > threshold ? 2 : 0);

    /* score plane with a one-pixel zero apron so NMS needs no bounds checks */
    const int sw = w + 2;
    uint8_t *score = (uint8_t *)calloc((size_t)sw * (size_t)(h + 2), 1);
    int ncorner = 0;

    for (int y = 3; y < h - 3; ++y) {
        const uint8_t *row = img + (size_t)y * step;
        uint8_t *srow = score + (size_t)(y + 1) * sw + 1;
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t *p = row + x;
            const int v = p[0];
            const uint8_t *tab = cls + 255 - v;
            /* necessary conditions on opposite ring pairs (a 9-arc covers one of each pair) */
            int m = tab[p[off[0]]] | tab[p[off[8]]];
            if (!m) continue;
            m &= tab[p[off[2]]] | tab[p[off[10]]];
            m &= tab[p[off[4]]] | tab[p[off[12]]];
            m &= tab[p[off[6]]] | tab[p[off[14]]];
            if (!m) continue;
            m &= tab[p[off[1]]] | tab[p[off[9]]];
            m &= tab[p[off[3]]] | tab[p[off[11]]];
            m &= tab[p[off[5]]] | tab[p[off[13]]];
            m &= tab[p[off[7]]] | tab[p[off[15]]];
            if (!m) continue;

            int corner = 0;
            if (m & 1) { /* ring darker than centre - t */
                const int vt = v - threshold;
                int run = 0;
                for (int k = 0; k < 25; ++k) {
                    if (p[off[k]] < vt) { if (++run > 8) { corner = 1; break; } }
                    else run = 0;
                }
            }
            if (!corner && (m & 2)) {
                const int vt = v + threshold;
                int run = 0;
                for (int k = 0; k < 25; ++k) {
                    if (p[off[k]] > vt) { if (++run > 8) { corner = 1; break; } }
                    else run = 0;
                }
            }
            if (corner) {
                int d[25];
                for (int k = 0; k < 25; ++k) d[k] = v - p[off[k]];
                srow[x] = (uint8_t)score_from_diffs(d, threshold);
                ++ncorner;
            }
        }
    }

    int n = 0;
    if (ncorner) {
        for (int y = 3; y < h - 3; ++y) {
            const uint8_t *s = score + (size_t)(y + 1) * sw + 1;
            for (int x = 3; x < w - 3; ++x) {
                const int c = s[x];
                if (!c) continue;
                /* a corner always has score >= threshold, and a zero score only occurs at
                 * threshold 0 where OpenCV's strict '>' against zero neighbours drops it too */
                if (c > s[x - 1] && c > s[x + 1] &&
                    c > s[x - sw - 1] && c > s[x - sw] && c > s[x - sw + 1] &&
                    c > s[x + sw - 1] && c > s[x + sw] && c > s[x + sw + 1]) {
                    if (n < cap) { out[n].x = x; out[n].y = y; out[n].score = c; }
                    ++n;
                }
            }
        }
    }
    free(score);
    return n;
}

/* ------------------------------------------------------------------ Gaussian 7x7 sigma 2
 * OpenCV >= 3.4 filters CV_8U with a bit-exact 8.8 fixed-point kernel; for ksize 7, sigma 2 the
 * quantised taps are 18,34,48,56,48,34,18 (sum 256).  Row pass -> 16-bit, column pass -> 32-bit,
 * result (v + 2^15) >> 16. */

static const int gk7[7] = {18, 34, 48, 56, 48, 34, 18};

void cvp_gaussian7_sigma2(const uint8_t *src, size_t sstep, uint8_t *dst, size_t dstep,
                          int w, int h)
{
    /* full intermediate plane keeps the code obviously alias-safe (src may equal dst) */
    uint16_t *hp = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)w * (size_t)h);
    int xi[7];
    for (int y = 0; y < h; ++y) {
        const uint8_t *s = src + (size_t)y * sstep;
        uint16_t *o = hp + (size_t)y * w;
        const int lo = w > 6 ? 3 : w, hi = w > 6 ? w - 3 : w;
        for (int x = 0; x < w; ++x) {
            if (x >= lo && x < hi) {
                const uint8_t *q = s + x;
                o[x] = (uint16_t)(18 * (q[-3] + q[3]) + 34 * (q[-2] + q[2]) +
                                  48 * (q[-1] + q[1]) + 56 * q[0]);
            } else {
                int acc = 0;
                for (int i = 0; i < 7; ++i) {
                    xi[i] = reflect101(x + i - 3, w);
                    acc += gk7[i] * s[xi[i]];
                }
                o[x] = (uint16_t)acc;
            }
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t *r[7];
        for (int j = 0; j < 7; ++j) r[j] = hp + (size_t)reflect101(y + j - 3, h) * w;
        uint8_t *d = dst + (size_t)y * dstep;
        for (int x = 0; x < w; ++x) {
            uint32_t v = 18u * ((uint32_t)r[0][x] + r[6][x]) + 34u * ((uint32_t)r[1][x] + r[5][x]) +
                         48u * ((uint32_t)r[2][x] + r[4][x]) + 56u * (uint32_t)r[3][x];
            d[x] = (uint8_t)((v + 32768u) >> 16);
        }
    }
    free(hp);
}

/* ------------------------------------------------------------------ fastAtan2 (scalar) */

float cvp_fast_atan2(float y, float x)
{
    const float rad2deg = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = 0.9997878412794807f * rad2deg;
    const float p3 = -0.3258083974640975f * rad2deg;
    const float p5 = 0.1555786518463281f * rad2deg;
    const float p7 = -0.04432655554792128f * rad2deg;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}
