// ubench_lds_sad.cu -- does shared-memory operand traffic slow the VABSDIFF4 pipe?
// 256 VABSDIFF4.U8.ACC per iteration (32 independent chains x 8 rounds) with L
// shared-memory loads of width W bytes interleaved; the loaded registers become
// SAD operands so nothing is dead.  Reports SAD thread-ops/s relative to L = 0.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_lds_sad tools/ubench_lds_sad.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

// W: bytes per load (4, 8, 16); L: loads per 256 SADs; BCAST: all lanes same address
template <int W, int L, bool BCAST, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) k(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int iters) {
  __shared__ __align__(16) uint32_t sm[8192];
  const int t = threadIdx.x;
  for (int i = t; i < 8192; i += blockDim.x) sm[i] = in[i & 1023];
  __syncthreads();
  uint32_t acc[32], a[16], b[8];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = in[(t + i) & 1023];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = in[(t * 3 + i) & 1023];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = in[(t * 5 + i) & 1023];
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm) + (BCAST ? 0 : (t & 31) * 16);
  constexpr int GAP = L > 0 ? 256 / L : 1 << 30;   // SADs between loads
  for (int it = 0; it < iters; ++it) {
    const uint32_t ad = base + ((it & 7) << 9);
#pragma unroll
    for (int s = 0; s < 256; ++s) {
      if (L > 0 && s % GAP == 0) {
        const int q = s / GAP;   // load number
        if (W == 16) {
          uint4 v = lds128(ad + (q & 15) * 1024);
          const int o = (q & 3) * 4;
          a[o] = v.x; a[o + 1] = v.y; a[o + 2] = v.z; a[o + 3] = v.w;
        } else if (W == 8) {
          uint2 v = lds64(ad + (q & 15) * 1024);
          const int o = (q & 7) * 2;
          a[o] = v.x; a[o + 1] = v.y;
        } else {
          a[q & 15] = lds32(ad + (q & 15) * 1024);
        }
      }
      const int i = s & 31;
      acc[i] = sad4(a[(s * 7 + 5) & 15], b[(s >> 2) & 7], acc[i]);
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) r ^= acc[i];
  out[blockIdx.x * blockDim.x + t] = r;
}

template <int W, int L, bool BCAST, int WARPS>
double run(const char* name, const uint32_t* din, uint32_t* dout, int sms, double ref) {
  const int iters = 4000;
  k<W, L, BCAST, WARPS><<<sms, WARPS * 32>>>(din, dout, 100);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    cudaEventRecord(e0);
    k<W, L, BCAST, WARPS><<<sms, WARPS * 32>>>(din, dout, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return 0; }
  const double gops = (double)iters * 256 * WARPS * 32 * sms / (best * 1e-3) / 1e9;
  printf("%-44s %9.1f G sad4/s  %5.1f%% of no-load\n", name, gops, ref > 0 ? 100 * gops / ref : 100.0);
  return gops;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  uint32_t *din, *dout;
  cudaMalloc(&din, 4096); cudaMalloc(&dout, (size_t)sms * 1024 * 4);
  uint32_t h[1024]; for (int i = 0; i < 1024; ++i) h[i] = 0x9E3779B9u * (i + 1);
  cudaMemcpy(din, h, 4096, cudaMemcpyHostToDevice);
  printf("8 warps / SM (2 per scheduler)\n");
  double r8 = run<16, 0, false, 8>("no loads", din, dout, sms, 0);
  run<16, 8, false, 8>("LDS.128 x8 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<16, 16, false, 8>("LDS.128 x16 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<16, 32, false, 8>("LDS.128 x32 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<16, 16, true, 8>("LDS.128 x16 /256 SAD (broadcast)", din, dout, sms, r8);
  run<8, 32, false, 8>("LDS.64 x32 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<4, 64, false, 8>("LDS.32 x64 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<4, 16, false, 8>("LDS.32 x16 /256 SAD (distinct addr)", din, dout, sms, r8);
  run<4, 64, true, 8>("LDS.32 x64 /256 SAD (broadcast)", din, dout, sms, r8);
  printf("16 warps / SM (4 per scheduler)\n");
  double r16 = run<16, 0, false, 16>("no loads", din, dout, sms, 0);
  run<16, 16, false, 16>("LDS.128 x16 /256 SAD (distinct addr)", din, dout, sms, r16);
  run<16, 32, false, 16>("LDS.128 x32 /256 SAD (distinct addr)", din, dout, sms, r16);
  printf("4 warps / SM (1 per scheduler)\n");
  double r4 = run<16, 0, false, 4>("no loads", din, dout, sms, 0);
  run<16, 16, false, 4>("LDS.128 x16 /256 SAD (distinct addr)", din, dout, sms, r4);
  return 0;
}
