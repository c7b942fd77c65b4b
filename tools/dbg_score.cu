// debug harness: candidate formulations of the FAST corner score on the device vs the host
// bisection reference (fast_score_px).  ptxas 12.9 miscompiled OpenCV's interleaved min/max chain
// for sm_100a; this checks which formulations are safe.   nvcc -arch=sm_100a -O3 ...
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "../orb_slam2_test_b200/csrc/k_fast.cuh"

// A: sliding-window minimum (window 9 over the circular 16), int32, then a max tree
__device__ __forceinline__ int scoreA(const int *e)
{
    int m2[16], m4[16], m8[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m8[k] = min(m4[k], m4[(k + 4) & 15]);
    int M = -100000;
#pragma unroll
    for (int k = 0; k < 16; ++k) M = max(M, min(m8[k], e[(k + 8) & 15]));
    return M - 1;
}
// B: same in float (exact for |e| <= 255)
__device__ __forceinline__ int scoreB(const int *e)
{
    float f[16], m2[16], m4[16], m8[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) f[k] = (float)e[k];
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = fminf(f[k], f[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = fminf(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m8[k] = fminf(m4[k], m4[(k + 4) & 15]);
    float M = -100000.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) M = fmaxf(M, fminf(m8[k], f[(k + 8) & 15]));
    return (int)M - 1;
}
// C: 16-bit pairs (e[j], e[j+8]) with the DPX-style s16x2 min/max intrinsics
__device__ __forceinline__ unsigned swp(unsigned v) { return __byte_perm(v, 0, 0x1032); }
__device__ __forceinline__ int scoreC(const int *e)
{
    unsigned P[8], Q[8], R[8], S[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) P[j] = ((unsigned)e[j] & 0xffffu) | ((unsigned)e[j + 8] << 16);
#pragma unroll
    for (int j = 0; j < 8; ++j) Q[j] = __vmins2(P[j], j + 1 < 8 ? P[j + 1] : swp(P[j + 1 - 8]));
#pragma unroll
    for (int j = 0; j < 8; ++j) R[j] = __vmins2(Q[j], j + 2 < 8 ? Q[j + 2] : swp(Q[j + 2 - 8]));
#pragma unroll
    for (int j = 0; j < 8; ++j) S[j] = __vmins2(R[j], j + 4 < 8 ? R[j + 4] : swp(R[j + 4 - 8]));
    unsigned M = 0x80008000u;
#pragma unroll
    for (int j = 0; j < 8; ++j) M = __vmaxs2(M, __vmins2(S[j], swp(P[j])));
    const int a = (short)(M & 0xffff), b = (short)(M >> 16);
    return max(a, b) - 1;
}

template <int V> __global__ void k(const uint8_t *img, int w, int h, int t_lo, int *out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x < 3 || x >= w - 3 || y < 3 || y >= h - 3) return;
    const uint8_t *p = img + y * w + x;
    const int v = p[0];
    int best = 0;
    for (int side = 0; side < 2; ++side) {
        int e[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) e[k] = side ? (int)FAST_RING(p, w, k) - v : v - (int)FAST_RING(p, w, k);
        const int s = V == 0 ? scoreA(e) : V == 1 ? scoreB(e) : scoreC(e);
        if (s >= t_lo) best = s;
    }
    out[y * w + x] = best;
}

int main()
{
    const int w = 640, h = 480;
    std::vector<uint8_t> img(w * h);
    uint8_t *d; int *o;
    cudaMalloc(&d, w * h); cudaMalloc(&o, w * h * 4);
    int total_bad = 0;
    for (int pattern = 0; pattern < 3; ++pattern) {
        srand(1 + pattern);
        for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) {
            int v = pattern == 0 ? ((x / 9 + y / 7) % 2) * 90 + 60 + rand() % 12
                  : pattern == 1 ? rand() % 256
                                 : 128 + (rand() % 41) - 20 + ((x / 5 + y / 6) % 2) * 30;
            img[y * w + x] = (uint8_t)v;
        }
        cudaMemcpy(d, img.data(), w * h, cudaMemcpyHostToDevice);
        for (int t_lo : {7, 20, 1}) for (int var = 0; var < 3; ++var) {
            cudaMemset(o, 0, w * h * 4);
            if (var == 0) k<0><<<dim3((w + 127) / 128, h), 128>>>(d, w, h, t_lo, o);
            if (var == 1) k<1><<<dim3((w + 127) / 128, h), 128>>>(d, w, h, t_lo, o);
            if (var == 2) k<2><<<dim3((w + 127) / 128, h), 128>>>(d, w, h, t_lo, o);
            std::vector<int> out(w * h);
            cudaMemcpy(out.data(), o, w * h * 4, cudaMemcpyDeviceToHost);
            int bad = 0, n = 0;
            for (int y = 3; y < h - 3; y++) for (int x = 3; x < w - 3; x++) {
                const int r = fast_score_px(img.data() + y * w + x, w, t_lo);
                n += r > 0;
                if (r != out[y * w + x]) { if (bad < 3) printf("  mismatch (%d,%d) host %d dev %d\n", x, y, r, out[y * w + x]); bad++; }
            }
            printf("pattern %d t_lo %2d variant %c: corners %d mismatches %d (%s)\n", pattern, t_lo, "ABC"[var], n, bad, cudaGetErrorString(cudaGetLastError()));
            total_bad += bad;
        }
    }
    printf("TOTAL mismatches %d\n", total_bad);
    return 0;
}
