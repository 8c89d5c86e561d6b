// ubench_stage_cvt.cu -- the cascade's float32 rows: convert to uint8 ON THE HOST while staging (4 B read, 1 B written to
// pinned memory, 1 B over PCIe) against staging the floats and converting on the device (4 B read, 4 B written, 4 B DMA).
//   nvcc -O3 -Xcompiler -mavx2 -o build/ubench_stage_cvt tools/ubench_stage_cvt.cu && build/ubench_stage_cvt
#include <cuda_runtime.h>
#include <immintrin.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
// n % 32 == 0; (uint8)((int)v + 128)
static void cvt(const float* s, uint8_t* d, size_t n) {
  const __m256i k128 = _mm256_set1_epi32(128), kff = _mm256_set1_epi32(255), perm = _mm256_setr_epi32(0, 4, 1, 5, 2, 6, 3, 7);
  for (size_t i = 0; i < n; i += 32) {
    __m256i a = _mm256_and_si256(_mm256_add_epi32(_mm256_cvttps_epi32(_mm256_loadu_ps(s + i)), k128), kff);
    __m256i b = _mm256_and_si256(_mm256_add_epi32(_mm256_cvttps_epi32(_mm256_loadu_ps(s + i + 8)), k128), kff);
    __m256i c = _mm256_and_si256(_mm256_add_epi32(_mm256_cvttps_epi32(_mm256_loadu_ps(s + i + 16)), k128), kff);
    __m256i e = _mm256_and_si256(_mm256_add_epi32(_mm256_cvttps_epi32(_mm256_loadu_ps(s + i + 24)), k128), kff);
    __m256i q = _mm256_packus_epi16(_mm256_packus_epi32(a, b), _mm256_packus_epi32(c, e));
    _mm256_storeu_si256(reinterpret_cast<__m256i*>(d + i), _mm256_permutevar8x32_epi32(q, perm));
  }
}
constexpr int MAXT = 8;
constexpr size_t PIN = (size_t)4 << 20;
struct LaneT { cudaStream_t st; char* pin[2]; cudaEvent_t done[2]; };
static LaneT L[MAXT];

// mode 0: floats staged as they are (sub bytes of floats per sub-chunk); mode 1: converted while staged (sub bytes of u8)
static double run(const float* src, char* dev, size_t nfloat, int T, size_t sub, int mode) {
  const double t0 = now_ms();
  auto work = [&](int t) {
    const size_t per = (nfloat / T + 1023) / 1024 * 1024, f0 = per * t < nfloat ? per * t : nfloat,
                 f1 = f0 + per < nfloat ? f0 + per : nfloat;
    const size_t step = mode ? sub : sub / 4;   // floats per sub-chunk
    int turn = 0; bool used[2] = {false, false};
    for (size_t f = f0; f < f1; f += step) {
      const size_t m = f1 - f < step ? f1 - f : step;
      const int b = turn; turn ^= 1;
      if (used[b]) cudaEventSynchronize(L[t].done[b]);
      if (mode) {
        cvt(src + f, reinterpret_cast<uint8_t*>(L[t].pin[b]), m);
        cudaMemcpyAsync(dev + f, L[t].pin[b], m, cudaMemcpyHostToDevice, L[t].st);
      } else {
        memcpy(L[t].pin[b], src + f, m * 4);
        cudaMemcpyAsync(dev + f * 4, L[t].pin[b], m * 4, cudaMemcpyHostToDevice, L[t].st);
      }
      cudaEventRecord(L[t].done[b], L[t].st);
      used[b] = true;
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  for (int t = 0; t < T; ++t) cudaStreamSynchronize(L[t].st);
  return now_ms() - t0;
}

int main() {
  for (int t = 0; t < MAXT; ++t) {
    cudaStreamCreateWithFlags(&L[t].st, cudaStreamNonBlocking);
    for (int b = 0; b < 2; ++b) {
      cudaHostAlloc(reinterpret_cast<void**>(&L[t].pin[b]), PIN, cudaHostAllocDefault);
      memset(L[t].pin[b], 0, PIN);
      cudaEventCreateWithFlags(&L[t].done[b], cudaEventDisableTiming);
    }
  }
  const size_t nfloat = (size_t)128 << 20;   // 512 MB of floats = one matrix of C5
  float* src = static_cast<float*>(malloc(nfloat * 4));
  for (size_t i = 0; i < nfloat; ++i) src[i] = (float)((int)(i * 2654435761u >> 24) - 128);
  char* dev = nullptr;
  cudaMalloc(reinterpret_cast<void**>(&dev), nfloat * 4);
  // check the conversion on a slice
  std::vector<uint8_t> chk(1024);
  cvt(src, chk.data(), 1024);
  for (int i = 0; i < 1024; ++i) if (chk[i] != (uint8_t)((int)src[i] + 128)) { printf("cvt mismatch at %d\n", i); return 1; }
  for (int mode = 0; mode < 2; ++mode)
    for (int T : {1, 2, 4, 6, 8})
      for (size_t sub : {(size_t)1 << 20, (size_t)4 << 20}) {
        double best = 1e9;
        for (int rep = 0; rep < 4; ++rep) { const double ms = run(src, dev, nfloat, T, sub, mode); if (ms < best) best = ms; }
        printf("512 MB of float rows, %s, T=%d sub=%zu MB: %.2f ms = %.1f GB/s of floats\n",
               mode ? "converted to u8 while staged" : "staged as floats               ", T, sub >> 20, best, nfloat * 4 / best / 1e6);
      }
  return 0;
}
