/*
 * borb_cta.cuh -- small CTA-wide primitives shared by the single-CTA-per-frame kernels.
 */
#ifndef BORB_CTA_CUH
#define BORB_CTA_CUH

/* Programmatic dependent launch (the single-frame shape, borb_api.cu enqueue_chain): a kernel launched with the
 * programmatic-serialization attribute may start while the kernel before it in the stream still runs; it does what
 * needs none of that kernel's results (shared-memory set-up, barrier init, constant tables), then waits here.  The wait
 * returns when the preceding grid has COMPLETED and its writes are visible, so it is also what makes the chain
 * transitive: every kernel of the chain calls it, before its first read of anything a predecessor wrote.  In a
 * kernel launched the ordinary way both instructions do nothing. */
__device__ __forceinline__ void borb_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
/* A pointer to data the preceding grid wrote, to be read only after borb_pdl_wait(): loads through a
 * `const T *__restrict__` kernel parameter are invariant loads to the compiler, which moves them freely -- also above the
 * wait's asm statement (seen in the SASS of k_orient_describe: the counts were loaded at the top of the kernel).  The
 * value that comes out of this asm is opaque and volatile asm statements keep their order, so no load through the
 * returned pointer can be scheduled before the wait. */
template <typename T>
__device__ __forceinline__ T *borb_pdl_after_wait(T *p)
{
    asm volatile("" : "+l"(p) : : "memory");
    __builtin_assume(__isGlobal(p)); /* the asm hides the address space too: without this the loads are generic LD */
    return p;
}
/* lets the NEXT kernel of the stream be scheduled once every CTA of this grid has got here (or exited) */
__device__ __forceinline__ void borb_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

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
