// ratio.cu -- K6: the ratio test and order-preserving compaction of the surviving
// matches, on the key buffer the matchers leave in HBM.
//
// Reference: example/ex01_essential_estimation.py:102-105
//     ratio    = nn_dist[:, 1] / nn_dist[:, 0].astype('float64')
//     pass_idx = ratio >= args.min_ratio
//     idx0, _  = nn_idx.T ;  xd = x[idx0[pass_idx]] ;  yd = y[pass_idx]
// i.e. an IEEE float64 division of the second-best by the best distance (x/0 = inf
// passes, 0/0 = nan fails; a "no neighbour" distance takes part with its sentinel
// value: INT_MAX for the int32 matcher, 2147483648.0f after CascadingHashNn.h:244's
// int -> float cast) and a boolean-mask selection, which keeps query order.  The
// match list is (idx0[pass], arange(yrows)[pass]).
//
// Three launches, all HBM-bound (16 B read per query row twice, 16 B written per
// match): per-block pass counts, one-block exclusive scan of the counts, scatter.
#include "spv_common.h"

#include <mutex>

namespace spv {
namespace {

constexpr int RT_THREADS = 256;

__device__ __forceinline__ bool ratio_pass(unsigned long long k0, unsigned long long k1, double min_ratio, bool dist_f32) {
  // "no neighbour" is the all-ones key; the reference's arrays hold INT_MAX there (BruteForceNnL1K2.h:99-103)
  const int d0 = k0 == KEY_NONE ? 0x7FFFFFFF : (int)(unsigned)(k0 >> 32);
  const int d1 = k1 == KEY_NONE ? 0x7FFFFFFF : (int)(unsigned)(k1 >> 32);
  // int32 / float64 (numpy promotes the int32 numerator), or float32 / float64
  const double den = dist_f32 ? (double)(float)d0 : (double)d0;
  const double num = dist_f32 ? (double)(float)d1 : (double)d1;
  return __ddiv_rn(num, den) >= min_ratio;   // nan compares false, +inf true
}

__global__ void __launch_bounds__(RT_THREADS)
ratio_count_kernel(const unsigned long long* __restrict__ keys, int yrows, double min_ratio, int dist_f32,
                   unsigned char* __restrict__ mask, unsigned* __restrict__ block_count) {
  __shared__ unsigned wsum[RT_THREADS / 32];
  const int row = blockIdx.x * RT_THREADS + threadIdx.x;
  bool pass = false;
  if (row < yrows) {
    pass = ratio_pass(keys[row], keys[(size_t)yrows + row], min_ratio, dist_f32 != 0);
    if (mask) mask[row] = pass ? 1 : 0;
  }
  const unsigned b = __ballot_sync(0xffffffffu, pass);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned s = 0;
#pragma unroll
    for (int w = 0; w < RT_THREADS / 32; ++w) s += wsum[w];
    block_count[blockIdx.x] = s;
  }
}

// one block: exclusive scan of nblocks counts in place, total -> *count
__global__ void __launch_bounds__(1024) ratio_scan_kernel(unsigned* __restrict__ block_count, int nblocks,
                                                          unsigned* __restrict__ count) {
  __shared__ unsigned wex[32];      // exclusive offset of every warp inside the current 1024-chunk
  __shared__ unsigned chunk_total;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned carry = 0;               // sum of all earlier chunks (same value in every thread)
  for (int base = 0; base < nblocks; base += 1024) {
    const int i = base + threadIdx.x;
    const unsigned v = i < nblocks ? block_count[i] : 0u;
    unsigned inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) wex[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const unsigned w = wex[lane];
      unsigned winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += t;
      }
      wex[lane] = winc - w;
      if (lane == 31) chunk_total = winc;
    }
    __syncthreads();
    if (i < nblocks) block_count[i] = carry + wex[warp] + inc - v;
    carry += chunk_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = carry;
}

__global__ void __launch_bounds__(RT_THREADS)
ratio_scatter_kernel(const unsigned long long* __restrict__ keys, int yrows, double min_ratio, int dist_f32,
                     const unsigned* __restrict__ block_offset, uint64_t* __restrict__ matches) {
  __shared__ unsigned wsum[RT_THREADS / 32];
  const int row = blockIdx.x * RT_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool pass = false;
  unsigned long long k0 = KEY_NONE;
  if (row < yrows) {
    k0 = keys[row];
    pass = ratio_pass(k0, keys[(size_t)yrows + row], min_ratio, dist_f32 != 0);
  }
  const unsigned b = __ballot_sync(0xffffffffu, pass);
  if (lane == 0) wsum[warp] = __popc(b);
  __syncthreads();
  unsigned off = block_offset[blockIdx.x];
  for (int w = 0; w < warp; ++w) off += wsum[w];
  if (pass) {
    const unsigned at = off + __popc(b & ((1u << lane) - 1u));
    const unsigned lo = (unsigned)k0;
    // idx0[pass]: a "no neighbour" index is (size_t)-1 in the reference's output
    matches[2 * (size_t)at] = (k0 == KEY_NONE) ? ~0ull : (uint64_t)lo;
    matches[2 * (size_t)at + 1] = (uint64_t)row;
  }
}

}  // namespace

size_t ratio_ws_bytes(int yrows) {
  const size_t nblocks = ((size_t)(yrows > 0 ? yrows : 0) + RT_THREADS - 1) / RT_THREADS;
  return (nblocks + 1) * sizeof(unsigned);
}

cudaError_t ratio_matches(const unsigned long long* keys, int yrows, double min_ratio, bool dist_f32, unsigned char* mask,
                          uint64_t* matches, unsigned* count, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (yrows <= 0) return cudaMemsetAsync(count, 0, sizeof(unsigned), stream);
  if (ws_bytes < ratio_ws_bytes(yrows)) return cudaErrorInvalidValue;
  const int nblocks = (yrows + RT_THREADS - 1) / RT_THREADS;
  unsigned* bc = static_cast<unsigned*>(ws);
  ratio_count_kernel<<<nblocks, RT_THREADS, 0, stream>>>(keys, yrows, min_ratio, dist_f32 ? 1 : 0, mask, bc);
  ratio_scan_kernel<<<1, 1024, 0, stream>>>(bc, nblocks, count);
  ratio_scatter_kernel<<<nblocks, RT_THREADS, 0, stream>>>(keys, yrows, min_ratio, dist_f32 ? 1 : 0, bc, matches);
  count_launch(3);
  return cudaGetLastError();
}

// ---- the variant that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  The per-block counts
// live in a file-scope device array instead of a caller-supplied workspace: RATIO_STATIC_MAX_ROWS query rows at most,
// (2^22 / 256 + 1) * 4 B = 65 540 B, held by every GPU that loads the library until the process exits.  Every call on
// a GPU shares the array, so calls that overlap on that GPU (different streams) race; calls on one stream are ordered.
// Nothing in it has to start at zero: ratio_count_kernel writes every entry the other two kernels read.
constexpr int RATIO_STATIC_MAX_ROWS = 1 << 22;
static __device__ __align__(256) unsigned g_ratio_static_counts[RATIO_STATIC_MAX_ROWS / RT_THREADS + 1];

int ratio_static_max_rows() { return RATIO_STATIC_MAX_ROWS; }

cudaError_t ratio_matches_static(const unsigned long long* keys, int yrows, double min_ratio, bool dist_f32,
                                 unsigned char* mask, uint64_t* matches, unsigned* count, cudaStream_t stream) {
  if (yrows > RATIO_STATIC_MAX_ROWS) return cudaErrorInvalidValue;
  // In host code the array's NAME is a host-side stand-in, not a device address: the address is asked for once per
  // device, with that device current (the first call on a device is the set-up; it allocates nothing either, but a
  // caller that captures a graph should make one call before the capture starts).
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static unsigned* addr[MAX_DEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEV) return cudaErrorInvalidDevice;
  unsigned* bc;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_ratio_static_counts);
      if (e != cudaSuccess) return e;
      addr[dev] = static_cast<unsigned*>(p);
    }
    bc = addr[dev];
  }
  return ratio_matches(keys, yrows, min_ratio, dist_f32, mask, matches, count, bc, sizeof(g_ratio_static_counts), stream);
}

}  // namespace spv
