/*
 * borb_cta.cuh -- small CTA-wide primitives shared by the single-CTA-per-frame kernels.
 */
#ifndef BORB_CTA_CUH
#define BORB_CTA_CUH

/* exclusive prefix sum of one int per thread over a CTA of NT threads (NT a multiple of 32, <= 1024):
 * shuffle scan inside each warp, then over the warp totals.  wsum: 32 ints of shared memory.
 * Returns the exclusive prefix; *total = sum over the CTA. */
template <int NT>
__device__ __forceinline__ int cta_exscan(int v, int *wsum, int *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < NT / 32 ? wsum[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        wsum[lane] = w; /* inclusive over warps */
    }
    __syncthreads();
    const int base = warp ? wsum[warp - 1] : 0;
    *total = wsum[NT / 32 - 1];
    return base + inc - v;
}

#endif
