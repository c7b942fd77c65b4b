/*
 * k_frame.cuh -- K8: the per-keypoint step AFTER the extractor:
 *   Frame::UndistortKeyPoints   src/Frame.cc:542-572  (cv::undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK))
 *   Frame::ComputeImageBounds   src/Frame.cc:574-609  (the four image corners through the same call)
 *   Frame::AssignFeaturesToGrid src/Frame.cc:292-307 + PosInGrid :510-520 (64 x 48 cells)
 *
 * cv::undistortPoints (OpenCV 4.13.0; pinned in oracle/frame_oracle.c): float point -> double,
 * x = (u-cx)/fx via ifx = 1/fx, five fixed-point iterations of the inverse Brown model
 *   icdist = (1 + ((k7 r2 + k6) r2 + k5) r2) / (1 + ((k3c r2 + k2c) r2 + k1c) r2)     (k5..k7 = 0 here)
 *   x = (x0 - deltaX) * icdist,  deltaX = 2 p1 x y + p2 (r2 + 2 x^2)  (+ zero thin-prism terms)
 * then re-projection with P = K and a final (float) cast.  All in IEEE double, every operation
 * individually rounded (__d*_rn: no FMA contraction), so the result is the bit pattern OpenCV gives.
 * The reference skips the whole thing when k1 == 0 (mvKeysUn = mvKeys, bounds = image rectangle).
 *
 * One CTA per frame: corners -> bounds (recomputed per launch: 4 points), keypoints -> undistorted
 * coordinates + cell; counting sort into the 64x48 grid (cell-major, mGrid[i][j] order, indices
 * ascending inside a cell as push_back in index order gives).
 */
#ifndef K_FRAME_CUH
#define K_FRAME_CUH

#include "borb_internal.h"
#include "k_describe.cuh"
#include "borb_cta.cuh"

#define FR_NT_THROUGHPUT 256 /* many frames per call: several CTAs per SM */
#define FR_NT_LATENCY 1024   /* a few frames: one keypoint per thread, the double-precision chains run side by side */
#define FR_GRID_COLS 64 /* FRAME_GRID_COLS, include/Frame.h:38 */
#define FR_GRID_ROWS 48 /* FRAME_GRID_ROWS, include/Frame.h:37 */
#define FR_CELLS (FR_GRID_COLS * FR_GRID_ROWS)

struct BorbCamera {
    float fx, fy, cx, cy;
    float k[5]; /* k1 k2 p1 p2 k3 (mDistCoef order; k3 = 0 when the settings give four) */
};

__device__ __forceinline__ void fr_undistort(const BorbCamera &cam, float u, float v, float *ox, float *oy)
{
    const double fx = cam.fx, fy = cam.fy, cx = cam.cx, cy = cam.cy;
    const double k1 = cam.k[0], k2 = cam.k[1], p1 = cam.k[2], p2 = cam.k[3], k3 = cam.k[4];
    const double ifx = __ddiv_rn(1.0, fx), ify = __ddiv_rn(1.0, fy);
    const double ud = u, vd = v;
    double x = __dmul_rn(__dsub_rn(ud, cx), ifx), y = __dmul_rn(__dsub_rn(vd, cy), ify);
    const double x0 = x, y0 = y;
#define M(a, b) __dmul_rn(a, b)
#define A(a, b) __dadd_rn(a, b)
    for (int j = 0; j < 5; ++j) {
        const double r2 = A(M(x, x), M(y, y));
        /* numerator 1 + ((k7*r2 + k6)*r2 + k5)*r2 with k5..k7 = 0 is exactly 1 */
        const double icdist = __ddiv_rn(1.0, A(1.0, M(A(M(A(M(k3, r2), k2), r2), k1), r2)));
        if (icdist < 0) {
            x = x0;
            y = y0;
            break;
        }
        /* 2*k[2]*x*y + k[3]*(r2 + 2*x*x) + k[8]*r2 + k[9]*r2*r2, the last two terms exactly 0 */
        const double dx = A(M(M(M(2.0, p1), x), y), M(p2, A(r2, M(M(2.0, x), x))));
        const double dy = A(M(p1, A(r2, M(M(2.0, y), y))), M(M(M(2.0, p2), x), y));
        x = M(__dsub_rn(x0, dx), icdist);
        y = M(__dsub_rn(y0, dy), icdist);
    }
    /* xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1) = 1 */
    *ox = (float)A(M(fx, x), cx);
    *oy = (float)A(M(fy, y), cy);
#undef M
#undef A
}

template <int FR_NT>
__global__ void __launch_bounds__(FR_NT)
k_frame_undistort_grid(const BorbPlan *__restrict__ plan, BorbCamera cam, const BorbKpRec *__restrict__ kp,
                       const int *__restrict__ n_kp, float2 *__restrict__ un_xy, int *__restrict__ cell_start,
                       int *__restrict__ cell_idx, int *__restrict__ cell_of, float *__restrict__ bounds_out)
{
    __shared__ int cnt[FR_CELLS];
    __shared__ int cur[FR_CELLS];
    __shared__ int part[32];
    __shared__ float corner[4][2];
    __shared__ float bnd[6]; /* minX maxX minY maxY invW invH */
    const int frame = blockIdx.x, tid = threadIdx.x;
    const int kp_total = plan->kp_total;
    const size_t fo = (size_t)frame * kp_total;
    const int n = n_kp[frame];
    const bool distorted = cam.k[0] != 0.0f;

    for (int c = tid; c < FR_CELLS; c += FR_NT) cnt[c] = 0;
    if (tid < 4) {
        const float u = (tid & 1) ? (float)plan->cols : 0.f, v = (tid & 2) ? (float)plan->rows : 0.f;
        if (distorted) fr_undistort(cam, u, v, &corner[tid][0], &corner[tid][1]);
        else { corner[tid][0] = u; corner[tid][1] = v; }
    }
    __syncthreads();
    if (tid == 0) {
        bnd[0] = fminf(corner[0][0], corner[2][0]);
        bnd[1] = fmaxf(corner[1][0], corner[3][0]);
        bnd[2] = fminf(corner[0][1], corner[1][1]);
        bnd[3] = fmaxf(corner[2][1], corner[3][1]);
        bnd[4] = __fdiv_rn((float)FR_GRID_COLS, __fsub_rn(bnd[1], bnd[0]));
        bnd[5] = __fdiv_rn((float)FR_GRID_ROWS, __fsub_rn(bnd[3], bnd[2]));
        if (bounds_out)
            for (int i = 0; i < 4; ++i) bounds_out[frame * 4 + i] = bnd[i];
    }
    __syncthreads();

    /* pass 1: undistort, cell of every keypoint, histogram */
    for (int i = tid; i < n; i += FR_NT) {
        const BorbKpRec k = kp[fo + i];
        float x = k.x, y = k.y;
        if (distorted) fr_undistort(cam, k.x, k.y, &x, &y);
        un_xy[fo + i] = make_float2(x, y);
        const int px = (int)roundf(__fmul_rn(__fsub_rn(x, bnd[0]), bnd[4]));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(y, bnd[2]), bnd[5]));
        int cell = -1;
        if (px >= 0 && px < FR_GRID_COLS && py >= 0 && py < FR_GRID_ROWS) {
            cell = px * FR_GRID_ROWS + py;
            atomicAdd(&cnt[cell], 1);
        }
        cell_of[fo + i] = cell;
    }
    __syncthreads();

    /* pass 2: exclusive scan of the 3072 counts (FR_CELLS / FR_NT consecutive cells per thread) */
    const int per = FR_CELLS / FR_NT;
    int sum = 0;
    for (int c = 0; c < per; ++c) sum += cnt[tid * per + c];
    int total;
    int base = cta_exscan<FR_NT>(sum, part, &total);
    int *start = cell_start + (size_t)frame * (FR_CELLS + 1);
    for (int c = 0; c < per; ++c) {
        start[tid * per + c] = base;
        cur[tid * per + c] = base;
        base += cnt[tid * per + c];
    }
    if (tid == 0) start[FR_CELLS] = total;
    __syncthreads();

    /* pass 3: scatter */
    for (int i = tid; i < n; i += FR_NT) {
        const int c = cell_of[fo + i];
        if (c >= 0) cell_idx[fo + atomicAdd(&cur[c], 1)] = i;
    }
    __syncthreads();
    /* order inside each cell: ascending keypoint index (cells hold a handful of entries) */
    for (int c = tid * per; c < (tid + 1) * per; ++c) {
        const int s = start[c], e = s + cnt[c];
        for (int a = s + 1; a < e; ++a) {
            const int v = cell_idx[fo + a];
            int b = a - 1;
            while (b >= s && cell_idx[fo + b] > v) { cell_idx[fo + b + 1] = cell_idx[fo + b]; --b; }
            cell_idx[fo + b + 1] = v;
        }
    }
}

#endif
