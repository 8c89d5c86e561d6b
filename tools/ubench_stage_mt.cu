// ubench_stage_mt.cu -- the library's lane scheme in isolation: T host threads, each with a copy stream and two pinned
// buffers, move a pageable array of N bytes to the device in sub-chunks (memcpy -> cudaMemcpyAsync -> event).
//   nvcc -O3 -o build/ubench_stage_mt tools/ubench_stage_mt.cu && build/ubench_stage_mt
// Prints, per (N, T, sub-chunk, one shared stream or one per thread), the wall time until every byte is on the device.
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
constexpr int MAXT = 8;
constexpr size_t PIN = (size_t)4 << 20;
struct LaneT { cudaStream_t st; char* pin[2]; cudaEvent_t done[2]; };
static LaneT L[MAXT];
static cudaStream_t shared_st;

static double run(const char* src, char* dev, size_t n, int T, size_t sub, bool one_stream, double* cpu_done) {
  double t0 = now_ms();
  std::vector<double> tq(T);
  auto work = [&](int t) {
    const size_t grain = 256 << 10, ng = (n + grain - 1) / grain;
    const size_t b0 = ng * t / T * grain, b1 = t + 1 == T ? n : ng * (t + 1) / T * grain;
    cudaStream_t st = one_stream ? shared_st : L[t].st;
    int turn = 0; bool used[2] = {false, false};
    for (size_t off = b0; off < b1; off += sub) {
      const size_t m = b1 - off < sub ? b1 - off : sub;
      const int b = turn; turn ^= 1;
      if (used[b]) cudaEventSynchronize(L[t].done[b]);
      memcpy(L[t].pin[b], src + off, m);
      cudaMemcpyAsync(dev + off, L[t].pin[b], m, cudaMemcpyHostToDevice, st);
      cudaEventRecord(L[t].done[b], st);
      used[b] = true;
    }
    tq[t] = now_ms();
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  double last = 0;
  for (int t = 0; t < T; ++t) last = tq[t] > last ? tq[t] : last;
  *cpu_done = last - t0;
  for (int t = 0; t < T; ++t) cudaStreamSynchronize(L[t].st);
  cudaStreamSynchronize(shared_st);
  return now_ms() - t0;
}

int main() {
  cudaStreamCreateWithFlags(&shared_st, cudaStreamNonBlocking);
  for (int t = 0; t < MAXT; ++t) {
    cudaStreamCreateWithFlags(&L[t].st, cudaStreamNonBlocking);
    for (int b = 0; b < 2; ++b) {
      cudaHostAlloc(reinterpret_cast<void**>(&L[t].pin[b]), PIN, cudaHostAllocDefault);
      memset(L[t].pin[b], 0, PIN);
      cudaEventCreateWithFlags(&L[t].done[b], cudaEventDisableTiming);
    }
  }
  const size_t sizes[] = {(size_t)8 << 20, (size_t)21 << 20, (size_t)256 << 20};
  for (size_t n : sizes) {
    char* src = static_cast<char*>(malloc(n)); char* dev = nullptr;
    memset(src, 3, n);
    cudaMalloc(reinterpret_cast<void**>(&dev), n);
    for (int one = 0; one < 2; ++one)
      for (int T : {1, 2, 4, 6, 8})
        for (size_t sub : {(size_t)256 << 10, (size_t)1 << 20, (size_t)4 << 20}) {
          double best = 1e9, cpu = 0;
          for (int rep = 0; rep < 6; ++rep) {
            double c; const double ms = run(src, dev, n, T, sub, one != 0, &c);
            if (ms < best) { best = ms; cpu = c; }
          }
          printf("%3zu MB  T=%d sub=%4zu KB %s  total %.3f ms (%.1f GB/s)  last sub-chunk queued at %.3f ms\n", n >> 20, T, sub >> 10,
                 one ? "shared stream" : "own streams  ", best, n / best / 1e6, cpu);
        }
    cudaFree(dev); free(src);
  }
  return 0;
}
