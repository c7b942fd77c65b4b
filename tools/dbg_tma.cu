// Stand-alone check of the TMA window load used by k_fast_cells (3-D map over a padded plane,
// box WP x ROWS x 1, unaligned x coordinate, out-of-range parts zero-filled).
// build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/dbg_tma tools/dbg_tma.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>

struct Maps { CUtensorMap m[4]; };

__global__ void k_bar_only(uint8_t *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(smem + 64);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}"
                 ::"r"(bar), "r"(0) : "memory");
    out[threadIdx.x] = 1;
}

template <int WP, int ROWS>
__global__ void k(const __grid_constant__ Maps maps, int level, int x, int y, int z, uint8_t *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(smem + 64);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem + 128);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(WP * ROWS) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"(&maps.m[level]), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}"
                 ::"r"(bar), "r"(0) : "memory");
    for (int i = threadIdx.x; i < WP * ROWS; i += blockDim.x) out[i] = smem[128 + i];
}

int main(int argc, char **argv)
{
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const int pitch = 1280, rows = 100, frames = 2, WP = 48, ROWS = 38;
    std::vector<uint8_t> h((size_t)pitch * rows * frames);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 7 + i / pitch) & 0xff);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size());
    cudaMalloc(&o, WP * ROWS);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                            const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    Maps maps;
    memset(&maps, 0, sizeof maps);
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * rows};
    const cuuint32_t box[3] = {WP, ROWS, 1}, es[3] = {1, 1, 1};
    CUresult r = ((Enc)fn)(&maps.m[1], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    int x = 45, y = 70, z = 1; /* rows 70..107: the last 8 are out of range */
    if (mode == 1) { k_bar_only<<<1, 64, 256>>>(o); printf("bar only: %s\n", cudaGetErrorString(cudaDeviceSynchronize())); return 0; }
    if (mode == 2) { x = 48; y = 10; }   /* aligned, fully inside */
    if (mode == 3) { x = 45; y = 10; }   /* unaligned x, inside */
    if (mode == 4) { x = 48; y = 70; }   /* aligned, rows out of range */
    if (mode == 5) { x = 1264; y = 10; } /* columns out of range */
    if (mode == 6) { x = 44; y = 10; }   /* 4-aligned */
    if (mode == 7) { x = 40; y = 10; }   /* 8-aligned */
    k<WP, ROWS><<<1, 64, 128 + WP * ROWS>>>(maps, 1, x, y, z, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    std::vector<uint8_t> g(WP * ROWS);
    cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < ROWS; ++rr)
        for (int c = 0; c < WP; ++c) {
            const int yy = y + rr, xx = x + c;
            const uint8_t want = (yy < rows && xx < pitch) ? h[((size_t)z * rows + yy) * pitch + xx] : 0;
            bad += g[rr * WP + c] != want;
        }
    printf("mismatches: %d\n", bad);
    return bad != 0 || e != cudaSuccess;
}
