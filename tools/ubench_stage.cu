// ubench_stage.cu -- what limits pageable -> pinned -> device staging of a few MB?
//   nvcc -O3 -Xcompiler -mavx2 -o build/ubench_stage tools/ubench_stage.cu && build/ubench_stage
// For each size: (1) the DMA alone from pinned memory nobody touched recently, (2) the DMA right after the CPU
// wrote the pinned buffer with glibc memcpy (lines dirty in the CPU caches), (3) the same with non-temporal
// stores (lines go to DRAM, nothing to snoop), and the CPU copy rates themselves.
#include <cuda_runtime.h>
#include <immintrin.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void copy_nt(void* dst, const void* src, size_t n) {   // dst 32-byte aligned, n % 32 == 0
  const __m256i* s = static_cast<const __m256i*>(src);
  __m256i* d = static_cast<__m256i*>(dst);
  for (size_t i = 0; i < n / 32; ++i) _mm256_stream_si256(d + i, _mm256_loadu_si256(s + i));
  _mm_sfence();
}

int main() {
  const size_t sizes[] = {(size_t)1 << 20, (size_t)4 << 20, (size_t)21 << 20, (size_t)64 << 20};
  cudaStream_t st;
  cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  for (size_t n : sizes) {
    char *src = static_cast<char*>(malloc(n)), *pin = nullptr, *dev = nullptr;
    memset(src, 3, n);
    cudaHostAlloc(reinterpret_cast<void**>(&pin), n, cudaHostAllocDefault);
    cudaMalloc(reinterpret_cast<void**>(&dev), n);
    memset(pin, 1, n);
    cudaMemcpyAsync(dev, pin, n, cudaMemcpyHostToDevice, st);
    cudaStreamSynchronize(st);
    double best[3] = {1e9, 1e9, 1e9}, cpu[3] = {0, 1e9, 1e9};
    for (int rep = 0; rep < 7; ++rep)
      for (int mode = 0; mode < 3; ++mode) {
        double t0 = now_ms();
        if (mode == 1) memcpy(pin, src, n);
        if (mode == 2) copy_nt(pin, src, n);
        double t1 = now_ms();
        cudaMemcpyAsync(dev, pin, n, cudaMemcpyHostToDevice, st);
        cudaStreamSynchronize(st);
        double t2 = now_ms();
        if (t2 - t1 < best[mode]) best[mode] = t2 - t1;
        if (mode && t1 - t0 < cpu[mode]) cpu[mode] = t1 - t0;
      }
    printf("%3zu MB  DMA alone %.3f ms (%.1f GB/s) | after memcpy: cpu %.3f ms (%.1f GB/s) DMA %.3f ms (%.1f GB/s) | after NT copy: cpu %.3f ms (%.1f GB/s) DMA %.3f ms (%.1f GB/s)\n",
           n >> 20, best[0], n / best[0] / 1e6, cpu[1], n / cpu[1] / 1e6, best[1], n / best[1] / 1e6, cpu[2], n / cpu[2] / 1e6,
           best[2], n / best[2] / 1e6);
    cudaFree(dev); cudaFreeHost(pin); free(src);
  }
  return 0;
}
