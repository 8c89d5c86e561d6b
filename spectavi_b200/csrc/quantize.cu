// quantize.cu -- K7: the descriptor quantiser that feeds the matchers, on the device
// (SURVEY.md 8f rank 2).  Reference: spectavi/feature.py:384-407
// normalize_to_ubyte_and_multiple_16_dim (numpy, float32 in -> float32 out):
//     x0   = x - np.mean(x, axis=0, keepdims=True)
//     norm = max(max(x0, axis=0), -min(x0, axis=0))
//     x0   = np.round(x0 / norm * 128) ; clip to [-128, 127] ; zero-pad columns to a multiple of 16
// and the step the callers apply next, (x + 128).astype('uint8')
// (example/ex01_essential_estimation.py:96-97; CascadingHashNn.h:236-239).
//
// Bit-exact against numpy: np.mean over axis 0 of a C-contiguous float32 matrix is a plain
// row-by-row float32 accumulation (probed on numpy 2.3: equal to the sequential loop on every
// column) followed by one division by float32(rows), so Q1 keeps that order -- one thread per
// column, rows in sequence; the loads are unrolled and independent, only the add chain is serial
// (4 cycles per row: ~2 ms at 1 M rows, 20 us at the castle pair's 10 k).  Q2 (column max / min
// of x - mean) is order-independent; Q3 applies IEEE float32 sub, div, mul and round-half-even.
// A constant column gives 0/0 = NaN there and here (float32 output; the uint8 output is then
// unspecified, as numpy's NaN -> uint8 cast is).
#include "spv_common.h"

#include <cfloat>

namespace spv {
namespace {

constexpr int Q1_UNROLL = 16;

// one thread per column; blockDim.x = 32 columns per CTA so that columns spread over SMs
__global__ void __launch_bounds__(32) colmean_kernel(const float* __restrict__ x, int rows, int dim,
                                                     float* __restrict__ mean) {
  const int col = blockIdx.x * 32 + threadIdx.x;
  if (col >= dim) return;
  const float* p = x + col;
  float acc = 0.f;
  int i = 0;
  for (; i + Q1_UNROLL <= rows; i += Q1_UNROLL) {
    float v[Q1_UNROLL];
#pragma unroll
    for (int u = 0; u < Q1_UNROLL; ++u) v[u] = __ldg(p + (size_t)(i + u) * dim);
#pragma unroll
    for (int u = 0; u < Q1_UNROLL; ++u) acc = __fadd_rn(acc, v[u]);   // the reference's order
  }
  for (; i < rows; ++i) acc = __fadd_rn(acc, __ldg(p + (size_t)i * dim));
  mean[col] = __fdiv_rn(acc, (float)rows);
}

// order-preserving float <-> uint map for atomicMax / atomicMin on non-NaN floats
__device__ __forceinline__ unsigned f2o(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float o2f(unsigned o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}

// hi[col] = max_i (x[i][col] - mean[col]), lo[col] = min_i (...), kept as ordered uints
__global__ void __launch_bounds__(256) colrange_kernel(const float* __restrict__ x, int rows, int dim,
                                                       const float* __restrict__ mean, unsigned* __restrict__ hi,
                                                       unsigned* __restrict__ lo, int rows_per_block) {
  const int r0 = blockIdx.x * rows_per_block, r1 = min(rows, r0 + rows_per_block);
  for (int col = threadIdx.x; col < dim; col += blockDim.x) {
    const float m = mean[col];
    float mx = -FLT_MAX, mn = FLT_MAX;
    bool any_nan = false;
    for (int i = r0; i < r1; ++i) {
      const float d = __fsub_rn(__ldg(x + (size_t)i * dim + col), m);
      any_nan |= (d != d);
      mx = fmaxf(mx, d);
      mn = fminf(mn, d);
    }
    if (r1 > r0) {
      // np.max / np.min propagate NaN
      atomicMax(hi + col, any_nan ? 0xFFFFFFFFu : f2o(mx));
      atomicMin(lo + col, any_nan ? 0u : f2o(mn));
    }
  }
}

__global__ void __launch_bounds__(256) quantize_kernel(const float* __restrict__ x, int rows, int dim, int new_dim,
                                                       const float* __restrict__ mean, const unsigned* __restrict__ hi,
                                                       const unsigned* __restrict__ lo, float* __restrict__ out_f32,
                                                       uint8_t* __restrict__ out_u8) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (size_t)rows * new_dim) return;
  const int col = (int)(e % new_dim);
  const size_t row = e / new_dim;
  float q = 0.f;   // zero padding (feature.py:404-406)
  if (col < dim) {
    const unsigned oh = hi[col], ol = lo[col];
    const float nan = __uint_as_float(0x7FC00000u);
    const float h = oh == 0xFFFFFFFFu ? nan : o2f(oh), l = ol == 0u ? nan : o2f(ol);
    // norm = np.max(np.stack([hi, -lo]), axis=0): NaN propagates
    const float nl = -l;
    const float norm = (h != h || nl != nl) ? nan : fmaxf(h, nl);
    const float x0 = __fsub_rn(__ldg(x + row * dim + col), mean[col]);
    q = rintf(__fmul_rn(__fdiv_rn(x0, norm), 128.f));   // np.round = round half to even
    if (q > 127.f) q = 127.f;                            // NaN fails both comparisons and stays NaN
    if (q < -128.f) q = -128.f;
  }
  if (out_f32) out_f32[e] = q;
  if (out_u8) out_u8[e] = (uint8_t)((int)q + 128);       // (x + 128).astype('uint8'), ex01:96-97
}

}  // namespace

size_t normalize_ws_bytes(int dim) { return (size_t)3 * (dim > 0 ? dim : 0) * sizeof(float); }

int normalize_new_dim(int dim) { return (dim + 15) / 16 * 16; }   // feature.py:403

cudaError_t normalize_rows(const float* dx, int rows, int dim, float* dout_f32, uint8_t* dout_u8, void* ws, size_t ws_bytes,
                           cudaStream_t stream) {
  if (rows <= 0 || dim <= 0) return cudaSuccess;
  if (ws_bytes < normalize_ws_bytes(dim)) return cudaErrorInvalidValue;
  float* mean = static_cast<float*>(ws);
  unsigned* hi = reinterpret_cast<unsigned*>(mean + dim);
  unsigned* lo = hi + dim;
  cudaError_t e;
  if ((e = cudaMemsetAsync(hi, 0x00, (size_t)dim * 4, stream)) != cudaSuccess) return e;   // below every ordered value
  if ((e = cudaMemsetAsync(lo, 0xFF, (size_t)dim * 4, stream)) != cudaSuccess) return e;   // above every ordered value
  colmean_kernel<<<(dim + 31) / 32, 32, 0, stream>>>(dx, rows, dim, mean);
  const int rpb = 256;
  colrange_kernel<<<(rows + rpb - 1) / rpb, 256, 0, stream>>>(dx, rows, dim, mean, hi, lo, rpb);
  const int new_dim = normalize_new_dim(dim);
  const size_t total = (size_t)rows * new_dim;
  quantize_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(dx, rows, dim, new_dim, mean, hi, lo, dout_f32, dout_u8);
  count_launch(3);
  return cudaGetLastError();
}

}  // namespace spv
