/*
 * oracle/frame_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the per-keypoint step after the extractor:
 *   fo_undistort_points  <- cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) as called by
 *                           Frame::UndistortKeyPoints (src/Frame.cc:542-572) and
 *                           Frame::ComputeImageBounds (:574-609); OpenCV 4.13.0 arithmetic, pinned
 *                           bit-exact against the cv2 wheel by tests/test_oracle_frame.py
 *   fo_image_bounds      <- Frame::ComputeImageBounds (:574-609)
 *   fo_assign_grid       <- Frame::AssignFeaturesToGrid (:292-307) with Frame::PosInGrid (:510-520)
 *                           and the grid constants of the constructor (:135-136), include/Frame.h:37-38
 * Compiled with -ffp-contract=off: OpenCV's baseline build has no FMA.
 */
#include <math.h>
#include <stdint.h>

#define GRID_COLS 64
#define GRID_ROWS 48

/* cam: fx fy cx cy k1 k2 p1 p2 k3 (floats, as mK / mDistCoef hold them) */
void fo_undistort_points(const float *xy, int n, const float *cam, float *out)
{
    const double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    double k[12] = {0};
    k[0] = cam[4]; k[1] = cam[5]; k[2] = cam[6]; k[3] = cam[7]; k[4] = cam[8];
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; ++i) {
        const double u = xy[2 * i], v = xy[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) { /* TermCriteria(MAX_ITER, 5, 0.01) */
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) {
                x = (u - cx) * ifx;
                y = (v - cy) * ify;
                break;
            }
            const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - dx) * icdist;
            y = (y0 - dy) * icdist;
        }
        const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
        out[2 * i] = (float)(xx * ww);
        out[2 * i + 1] = (float)(yy * ww);
    }
}

/* bounds: mnMinX mnMaxX mnMinY mnMaxY */
void fo_image_bounds(int cols, int rows, const float *cam, float *bounds)
{
    if (cam[4] != 0.0f) {
        const float c[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
        float m[8];
        fo_undistort_points(c, 4, cam, m);
        bounds[0] = fminf(m[0], m[4]);
        bounds[1] = fmaxf(m[2], m[6]);
        bounds[2] = fminf(m[1], m[3]);
        bounds[3] = fmaxf(m[5], m[7]);
    } else {
        bounds[0] = 0.f; bounds[1] = (float)cols; bounds[2] = 0.f; bounds[3] = (float)rows;
    }
}

/* mvKeysUn (only pt changes) */
void fo_undistort_keypoints(const float *xy, int n, const float *cam, float *out)
{
    if (cam[4] == 0.0f) {
        for (int i = 0; i < 2 * n; ++i) out[i] = xy[i];
        return;
    }
    fo_undistort_points(xy, n, cam, out);
}

/* cell (i, j) = start[i*48 + j] .. start[i*48 + j + 1] of idx: mGrid[i][j] in push_back order.
 * Returns the number of keypoints that landed inside the grid. */
int fo_assign_grid(const float *xy_un, int n, const float *bounds, int32_t *start /* 3073 */, int32_t *idx /* n */)
{
    const float inv_w = (float)GRID_COLS / (bounds[1] - bounds[0]);
    const float inv_h = (float)GRID_ROWS / (bounds[3] - bounds[2]);
    for (int c = 0; c <= GRID_COLS * GRID_ROWS; ++c) start[c] = 0;
    for (int pass = 0; pass < 2; ++pass) {
        for (int i = 0; i < n; ++i) {
            const int px = (int)roundf((xy_un[2 * i] - bounds[0]) * inv_w);
            const int py = (int)roundf((xy_un[2 * i + 1] - bounds[2]) * inv_h);
            if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
            const int c = px * GRID_ROWS + py;
            if (pass == 0) start[c + 1]++;
            else idx[start[c]++] = i;
        }
        if (pass == 0)
            for (int c = 0; c < GRID_COLS * GRID_ROWS; ++c) start[c + 1] += start[c];
        else { /* the fill advanced start[c] to the end of cell c: shift back */
            for (int c = GRID_COLS * GRID_ROWS; c > 0; --c) start[c] = start[c - 1];
            start[0] = 0;
        }
    }
    return start[GRID_COLS * GRID_ROWS];
}
