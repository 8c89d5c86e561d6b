// ubench_tma_box.cu -- how long does TMA take to fill one 64-row x 128-byte stage
//   (a) as 8 boxes of 16 B x 64 rows (linear [chunk][row][16 B] layout), vs
//   (b) as 1 box of 128 B x 64 rows (SWIZZLE_128B), vs
//   (c) as 1 box through a permuted 3-D map (16 B, 64 rows, 8 chunks)?
// One CTA per SM, one thread issues, everybody streams the same 200k x 128 u8
// matrix like K1 does.  Reports average cycles from first issue to completion.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_tma_box tools/ubench_tma_box.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}

template <int MODE>
__global__ void __launch_bounds__(32, 1) k(const __grid_constant__ CUtensorMap lin, const __grid_constant__ CUtensorMap swz,
                                           const __grid_constant__ CUtensorMap perm, int ntiles, int depth,
                                           long long* cyc_out, uint32_t* sink) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bars[8];
  if (threadIdx.x != 0) return;
  for (int s = 0; s < 8; ++s)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[s])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  uint32_t acc = 0;
  long long t0 = clock64();
  // keep `depth` stages in flight (like K1's ring), wait for the oldest, re-issue
  for (int t = 0; t < ntiles + depth; ++t) {
    if (t >= depth) {
      const int s = (t - depth) % depth;
      const uint32_t par = ((t - depth) / depth) & 1;
      while (!try_wait(smem_u32(&bars[s]), par)) {}
      acc += smem[s * 8192 + (t & 1023)];
    }
    if (t < ntiles) {
      const int s = t % depth;
      const uint32_t bar = smem_u32(&bars[s]), dst = smem_u32(smem + s * 8192);
      const int row0 = ((t * 148 + blockIdx.x) * 64) % 199936;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(8192) : "memory");
      if (MODE == 0) {
        for (int c = 0; c < 8; ++c)
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                       ::"r"(dst + c * 1024), "l"(&lin), "r"(c * 16), "r"(row0), "r"(bar) : "memory");
      } else if (MODE == 1) {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(dst), "l"(&swz), "r"(0), "r"(row0), "r"(bar) : "memory");
      } else {
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"(&perm), "r"(0), "r"(row0), "r"(0), "r"(bar) : "memory");
      }
    }
  }
  long long t1 = clock64();
  cyc_out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x] = acc;
}

typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int rows = 200000, dim = 128;
  uint8_t* d; cudaMalloc(&d, (size_t)rows * dim); cudaMemset(d, 7, (size_t)rows * dim);
  void* sym; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
  Enc enc = (Enc)sym;
  CUtensorMap lin, swz, perm;
  cuuint32_t es2[2] = {1, 1}, es3[3] = {1, 1, 1};
  {
    cuuint64_t gd[2] = {(cuuint64_t)dim, (cuuint64_t)rows}, gs[1] = {(cuuint64_t)dim};
    cuuint32_t b1[2] = {16, 64}, b2[2] = {128, 64};
    CUresult r1 = enc(&lin, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gd, gs, b1, es2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = enc(&swz, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gd, gs, b2, es2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode lin %d swz %d\n", (int)r1, (int)r2);
  }
  bool have_perm;
  {
    cuuint64_t gd[3] = {16, (cuuint64_t)rows, 8}, gs[2] = {(cuuint64_t)dim, 16};
    cuuint32_t b[3] = {16, 64, 8};
    CUresult r = enc(&perm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, gd, gs, b, es3, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode permuted 3-D map: %d\n", (int)r);
    have_perm = (r == CUDA_SUCCESS);
    if (!have_perm) perm = lin;
  }
  long long* dc; uint32_t* ds; cudaMalloc(&dc, 148 * 8); cudaMalloc(&ds, 148 * 4);
  const int ntiles = 2000;
  for (int depth = 1; depth <= 4; depth *= 2) {
    for (int mode = 0; mode < 3; ++mode) {
      if (mode == 2 && !have_perm) continue;
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) { cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); k<0><<<148, 32, 65536>>>(lin, swz, perm, ntiles, depth, dc, ds); }
        if (mode == 1) { cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); k<1><<<148, 32, 65536>>>(lin, swz, perm, ntiles, depth, dc, ds); }
        if (mode == 2) { cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); k<2><<<148, 32, 65536>>>(lin, swz, perm, ntiles, depth, dc, ds); }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(e)); return 1; }
      }
      long long h[148]; cudaMemcpy(h, dc, sizeof(h), cudaMemcpyDeviceToHost);
      double s = 0; for (int i = 0; i < 148; ++i) s += (double)h[i];
      const char* nm[3] = {"8 x (16 B x 64 rows) linear", "1 x (128 B x 64 rows) swizzle128", "1 x permuted 3-D (16 B, 64 rows, 8)"};
      printf("depth %d  %-36s %8.0f cycles per 8 KB stage\n", depth, nm[mode], s / 148 / ntiles);
    }
  }
  return 0;
}
