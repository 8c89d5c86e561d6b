// ubench_gather.cu -- the ceiling K4 (probe_rerank_kernel) works against: random 128-byte row gathers.
// 8 lanes read one 128-byte row (16 B each), row ids come from a sequential id list (like K4's staged candidate
// lists), 8 rows in flight per 8-lane group, 4 groups per warp, 128 threads per CTA, 8 CTAs per SM (64 registers) --
// K4's shape with everything but the gathers removed.  Reports effective GB/s for a table that nearly fits the L2
// (128 MB, C5's train set) and one that does not (1 GB).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_gather tools/ubench_gather.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

__global__ void __launch_bounds__(128, 8) gather(const uint4* __restrict__ table, const int* __restrict__ ids, long long per_group,
                                                 unsigned* __restrict__ sink) {
  const int lane = threadIdx.x & 31, l8 = lane & 7;
  const long long group = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int* my = ids + group * per_group;
  unsigned acc = 0;
  for (long long i = 0; i < per_group; i += 8) {
    const int4 a = *reinterpret_cast<const int4*>(my + i), b = *reinterpret_cast<const int4*>(my + i + 4);
    const int id[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldg(table + (size_t)id[u] * 8 + l8);
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345u) sink[0] = acc;
}

int main() {
  const int sms = 148, ctas = sms * 8, groups = ctas * 16;
  const long long per_group = 64 * 60;          // ~ K4: 60 candidates x 64 queries per group
  const long long total = (long long)groups * per_group;
  for (size_t rows : {(size_t)1 << 20, (size_t)8 << 20}) {
    uint4* table; int* ids; unsigned* sink;
    cudaMalloc(&table, rows * 128); cudaMemset(table, 1, rows * 128);
    cudaMalloc(&ids, total * 4); cudaMalloc(&sink, 4);
    std::vector<int> h(total);
    uint64_t s = 88172645463325252ull;
    for (long long i = 0; i < total; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % rows); }
    cudaMemcpy(ids, h.data(), total * 4, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0); gather<<<ctas, 128>>>(table, ids, per_group, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    printf("table %4zu MB: %lld random 128-byte gathers in %.3f ms = %.0f GB/s effective (%s)\n", rows * 128 >> 20, total, best,
           total * 128.0 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
    cudaFree(table); cudaFree(ids); cudaFree(sink);
  }
  return 0;
}
