// How fast can one host thread move a 466 KB frame from pageable memory into a page-locked staging buffer?
// nvcc -O3 -o gpurun_out/packbench tools/packbench.cu -Xcompiler -mavx2
#include <cuda_runtime.h>
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>
static double now() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void copy_nt(uint8_t *d, const uint8_t *s, size_t n)
{
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        __m256i a = _mm256_loadu_si256((const __m256i *)(s + i)), b = _mm256_loadu_si256((const __m256i *)(s + i + 32));
        __m256i c = _mm256_loadu_si256((const __m256i *)(s + i + 64)), e = _mm256_loadu_si256((const __m256i *)(s + i + 96));
        _mm256_stream_si256((__m256i *)(d + i), a); _mm256_stream_si256((__m256i *)(d + i + 32), b);
        _mm256_stream_si256((__m256i *)(d + i + 64), c); _mm256_stream_si256((__m256i *)(d + i + 96), e);
    }
    _mm_sfence();
    if (i < n) memcpy(d + i, s + i, n - i);
}
static void copy_avx(uint8_t *d, const uint8_t *s, size_t n)
{
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        __m256i a = _mm256_loadu_si256((const __m256i *)(s + i)), b = _mm256_loadu_si256((const __m256i *)(s + i + 32));
        __m256i c = _mm256_loadu_si256((const __m256i *)(s + i + 64)), e = _mm256_loadu_si256((const __m256i *)(s + i + 96));
        _mm256_store_si256((__m256i *)(d + i), a); _mm256_store_si256((__m256i *)(d + i + 32), b);
        _mm256_store_si256((__m256i *)(d + i + 64), c); _mm256_store_si256((__m256i *)(d + i + 96), e);
    }
    if (i < n) memcpy(d + i, s + i, n - i);
}
static void copy_movsb(uint8_t *d, const uint8_t *s, size_t n) { asm volatile("rep movsb" : "+D"(d), "+S"(s), "+c"(n)::"memory"); }
int main()
{
    const size_t n = 1241 * 376;
    uint8_t *src = (uint8_t *)malloc(n + 64), *pin = nullptr, *dev = nullptr;
    for (size_t i = 0; i < n; ++i) src[i] = (uint8_t)(i * 7);
    cudaHostAlloc((void **)&pin, n + 4096, cudaHostAllocDefault);
    cudaMalloc((void **)&dev, n);
    cudaStream_t st; cudaStreamCreate(&st);
    const char *names[] = {"memcpy", "avx2 cached stores", "avx2 streaming stores", "rep movsb"};
    for (int with_dma = 0; with_dma < 2; ++with_dma)
        for (int m = 0; m < 4; ++m) {
            std::vector<double> t;
            for (int it = 0; it < 400; ++it) {
                const double a = now();
                if (m == 0) memcpy(pin, src, n); else if (m == 1) copy_avx(pin, src, n); else if (m == 2) copy_nt(pin, src, n); else copy_movsb(pin, src, n);
                const double b = now();
                if (with_dma) { cudaMemcpyAsync(dev, pin, n, cudaMemcpyHostToDevice, st); cudaStreamSynchronize(st); }
                const double c = now();
                if (it >= 50) t.push_back(with_dma ? c - a : b - a);
            }
            std::sort(t.begin(), t.end());
            printf("%s %-22s p50 %.1f us (%.1f GB/s if copy only)\n", with_dma ? "pack + H2D + sync:" : "pack only:        ", names[m], t[t.size() / 2], n / t[t.size() / 2] / 1e3);
        }
    return 0;
}
