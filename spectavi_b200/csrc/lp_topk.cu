// lp_topk.cu -- K1f: all-pairs sum |x - y|^p + per-query top-k over float32 / int32
// rows, hand-written for sm_100a.  Replaces BruteForceNn::find_neighbours
// (reference src/BruteForceNn.h:37-120) behind nn_bruteforce / nn_bruteforcei.
//
// Arithmetic is the reference's, operation for operation, so distances are
// BIT-EXACT and not merely close: diff is a float (:68) -- for int32 rows the
// integer difference converted to float --, the term is |d| (p == 1), d*d
// (p == 2), sqrtf|d| (p == .5) or (float)pow((double)|d|, p) (:70-78), truncated
// toward zero for the int32 instantiation (Scalar pval, :69), and the terms are
// added one by one in column order (:79) in Scalar arithmetic: every thread sums
// its 8 x 8 pairs sequentially over the columns with non-fused IEEE operations
// (__fadd_rn / __fmul_rn: the reference is built without -march, so no FMA).
// Only pow() may differ from glibc's in the last place (CUDA's is <= 2 ulp in
// double before the rounding to float).
//
// Selection is the lexicographic (dist, idx) top-k.  The reference's heap gives
// the same answer except when a candidate TIES the current k-th distance, where
// it pushes without popping and later stores to column -1 (:102-108,:111-118;
// undefined behaviour); such a candidate is rejected here.
// The early-exit prune (:82-86) is exact for mu == 0 and never changes the
// result; mu != 0 (an order-dependent approximation) is emulated by lp_topk_mu_kernel at the end of this file.
//
// Bound: FP32 CUDA-core pipe, 2 (p=1) or 3 (p=2) dependent-free ops per element.
//
// FILT = true is the filter overload (BruteForceNn.h:37-49: unordered_set filter_x /
// filter_y), in the batched form its one caller needs (KMedians::find_nearest_neighbours,
// src/KMedians.h:260-295: one filtered search per query cluster): every CTA takes a tile
// {ypos, ycnt, xpos, xcnt} -- up to 128 query rows named by ysel[ypos ..) against the
// candidate rows named by xsel[xpos .. xpos + xcnt) -- so ALL clusters run in one launch.
// Rows are gathered through the lists while staging; keys carry the real train row id.
#include "spv_common.h"

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <type_traits>

namespace spv {
namespace {

constexpr int TB = 128;          // query rows and train rows per CTA tile
constexpr int KC = 16;           // columns staged per step
constexpr int PITCH = TB + 4;    // shared-memory row pitch (floats)
constexpr int NT = 256;          // 16 x 16 threads, 8 x 8 register tile each
constexpr int QCAP = 1024;       // candidate queue entries per drain pass
constexpr uint32_t THR_NONE = 0xFFFFFFFFu;

enum PMode { P_POW = 0, P_ONE = 1, P_TWO = 2, P_HALF = 3 };

// Packed FP32 through explicit PTX.  ptxas (12.9) contracts `mul.rn.f32x2` + `add.rn.f32x2` into one FFMA2 although
// both carry an explicit rounding modifier (it also sees through fma(d, d, -0) and through the __fmul2_rn /
// __fadd2_rn intrinsics) -- measured: FFMA2 in the p = 2 kernel and last-place differences against the reference,
// which is built without FMA.  The scalar `mul.rn.f32` IS respected, so the square is two scalar FMULs on the
// unpacked halves between the packed subtraction and the packed accumulation: 4 instructions per two
// pair-columns instead of 6, bit-exact (profiles/r02_sass_opcodes.txt: no FFMA2 / FFMA in these kernels).
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long add2_rn(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long sub2_rn(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// acc(2 lanes) += term(x(2 lanes) - y): |d| for P1, d * d otherwise; every operation IEEE-rn per lane
template <bool P1>
__device__ __forceinline__ void pair_step(float& a0, float& a1, float x0, float x1, float y) {
  unsigned long long d = sub2_rn(pk2(x0, x1), pk2(y, y));
  if (P1) {
    float d0, d1;
    upk2(d, d0, d1);
    d = pk2(fabsf(d0), fabsf(d1));
  } else {
    float d0, d1;
    upk2(d, d0, d1);
    asm("mul.rn.f32 %0, %0, %0;" : "+f"(d0));
    asm("mul.rn.f32 %0, %0, %0;" : "+f"(d1));
    d = pk2(d0, d1);
  }
  upk2(add2_rn(pk2(a0, a1), d), a0, a1);
}

template <int PM>
__device__ __forceinline__ float term_f(float d, double p) {
  if (PM == P_ONE) return fabsf(d);
  if (PM == P_TWO) return __fmul_rn(d, d);
  if (PM == P_HALF) return sqrtf(fabsf(d));
  return (float)pow((double)fabsf(d), p);
}

struct LpParams {
  const void* x;
  const void* y;
  int xrows, yrows, dim, k;
  double p;
  int nxt;                         // train tiles in total
  int splits;                      // train-side splits (gridDim.y)
  unsigned long long* partial;     // [splits][yrows][k] sorted keys, KEY_NONE padded
  // FILT only
  const uint32_t* ysel;            // query row ids, group after group
  const uint32_t* xsel;            // candidate row ids, group after group
  const int4* tiles;               // per blockIdx.x: {ypos, ycnt <= TB, xpos, xcnt}
  int nysel;                       // entries in ysel (row stride of partial)
  // FLOOR only (k > LP_MAX_K runs in passes of LP_MAX_K): a candidate must be GREATER than floor[query row],
  // the last key the previous pass delivered; KEY_NONE = the row is exhausted
  const unsigned long long* floor;
};

// Train rows of a thread's 8 x 8 register tile: columns j = 0..3 are tile rows 4 tx + j, columns 4..7 are
// 64 + 4 tx + (j - 4), so that the lanes of a quarter-warp read one contiguous 128 B span per LDS.128.
// (With 8 consecutive rows per thread the 32 B lane stride makes each operand load a 2-way bank conflict.
// Removing it left the kernel time unchanged, 4.96 ms at 20 k x 20 k x 132: the loop is bound by the FP32
// pipe -- ncu: math-pipe throttle and dispatch stalls with ~1.9 eligible warps per scheduler -- not by LDS.)
__device__ __forceinline__ int xrow(int tx, int j) { return (j & 4) * 16 + tx * 4 + (j & 3); }

// key = order-preserving 32 bits of the distance << 32 | train index
__device__ __forceinline__ uint32_t dist_bits(float d) { return __float_as_uint(d); }   // d >= 0
__device__ __forceinline__ uint32_t dist_bits(int d) { return (uint32_t)d; }            // d >= 0

// Two CTAs per SM (128 registers, ~170 B of spills around the prefetch registers): with one CTA the two
// warps per scheduler left 40 % of the issue slots empty on fixed-latency waits, barriers and LDS
// (ncu: issue active 60 -> 78 %, fma pipe 51 -> 65 %, 20 k x 20 k x 132 p = 1: 6.3 -> 4.9 ms).
template <typename T, int PM, bool FILT, bool FLOOR>
__global__ void __launch_bounds__(NT, 2) lp_topk_kernel(const LpParams P) {
  static_assert(!(FILT && FLOOR), "the filter overload runs single-pass (k <= LP_MAX_K)");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* ys = reinterpret_cast<T*>(smem_raw);                     // [KC][PITCH]
  T* xs = ys + KC * PITCH;                                    // [KC][PITCH]
  uint32_t* thr = reinterpret_cast<uint32_t*>(xs + KC * PITCH);   // [TB] k-th best distance bits
  int* cnt = reinterpret_cast<int*>(thr + TB);                // [TB] entries in each row's list
  unsigned long long* qkey = reinterpret_cast<unsigned long long*>(cnt + TB);   // [QCAP]
  unsigned short* qrow = reinterpret_cast<unsigned short*>(qkey + QCAP);        // [QCAP]
  int* qn = reinterpret_cast<int*>(qrow + QCAP);              // [4]
  uint32_t* yid = reinterpret_cast<uint32_t*>(qn + 4);        // [TB] FILT: query row ids of this tile
  uint32_t* xid = yid + TB;                                   // [TB] FILT: train row ids of the current train tile
  unsigned long long* list = reinterpret_cast<unsigned long long*>(xid + TB);   // [TB][k] ascending

  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int k = P.k, dim = P.dim;
  int y0, ycnt, xbase, xcnt, nxt, prow;   // prow: row stride of P.partial
  if (FILT) {
    const int4 tl = P.tiles[blockIdx.x];
    y0 = tl.x; ycnt = tl.y; xbase = tl.z; xcnt = tl.w;
    nxt = (xcnt + TB - 1) / TB; prow = P.nysel;
  } else {
    y0 = blockIdx.x * TB; ycnt = min(TB, P.yrows - y0); xbase = 0; xcnt = P.xrows;
    nxt = P.nxt; prow = P.yrows;
  }
  const T* __restrict__ gx = reinterpret_cast<const T*>(P.x);
  const T* __restrict__ gy = reinterpret_cast<const T*>(P.y);
  // contiguous range of train tiles of this split, scanned in ascending order
  const int t0 = (int)((long long)nxt * blockIdx.y / P.splits);
  const int t1 = (int)((long long)nxt * (blockIdx.y + 1) / P.splits);

  if (tid < TB) {
    thr[tid] = THR_NONE; cnt[tid] = 0;
    if (FILT) yid[tid] = tid < ycnt ? P.ysel[y0 + tid] : 0u;
  }
  __syncthreads();

  for (int t = t0; t < t1; ++t) {
    const int x0 = t * TB;
    if (FILT) {   // the previous tile's last reader of xid is behind the __syncthreads_or that ended its selection
      if (tid < TB) xid[tid] = x0 + tid < xcnt ? P.xsel[xbase + x0 + tid] : 0u;
      __syncthreads();
    }
    T acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = (T)0;

    // KC columns of both tiles per step, transposed to [column][row] in shared memory; the next step's
    // global loads are issued into registers before this step's arithmetic and land behind it
    T py[(TB * KC) / NT], px[(TB * KC) / NT];
    auto fetch = [&](int c0) {
#pragma unroll
      for (int q = 0; q < (TB * KC) / NT; ++q) {
        const int e = tid + NT * q, col = e % KC, row = e / KC;
        const bool cok = c0 + col < dim;
        py[q] = (T)0; px[q] = (T)0;
        if (cok && row < ycnt) py[q] = gy[(size_t)(FILT ? yid[row] : (uint32_t)(y0 + row)) * dim + c0 + col];
        if (cok && x0 + row < xcnt) px[q] = gx[(size_t)(FILT ? xid[row] : (uint32_t)(x0 + row)) * dim + c0 + col];
      }
    };
    fetch(0);
    for (int c0 = 0; c0 < dim; c0 += KC) {
#pragma unroll
      for (int q = 0; q < (TB * KC) / NT; ++q) {
        const int e = tid + NT * q, col = e % KC, row = e / KC;
        ys[col * PITCH + row] = py[q];
        xs[col * PITCH + row] = px[q];
      }
      __syncthreads();
      if (c0 + KC < dim) fetch(c0 + KC);
      const int kc = min(KC, dim - c0);
      // (reading column c + 1's operands while column c is summed was measured: no gain -- the loop is
      // bound by FADD issue, ncu: fma pipe 51 % of peak at 2 FADD per pair-column, not by LDS latency)
      // a warp owns 16 query rows; warps whose rows all lie beyond the tile's last query row (ragged last tile,
      // small clusters of the filter overload: 45 of 128 rows) skip the arithmetic, not the staging or barriers
      // (filter overload only: in the dense kernel the extra predicate cost 6 % -- measured -- for one ragged tile)
      if (!FILT || (tid >> 5) * 16 < ycnt)
#pragma unroll 4
      for (int c = 0; c < kc; ++c) {
        T ya[8], xb[8];
        const float4* yp = reinterpret_cast<const float4*>(ys + c * PITCH + ty * 8);
        const float4* xp = reinterpret_cast<const float4*>(xs + c * PITCH + tx * 4);
        *reinterpret_cast<float4*>(&ya[0]) = yp[0]; *reinterpret_cast<float4*>(&ya[4]) = yp[1];
        *reinterpret_cast<float4*>(&xb[0]) = xp[0]; *reinterpret_cast<float4*>(&xb[4]) = xp[TB / 8];   // + 64 rows
        if constexpr (std::is_same<T, float>::value && (PM == P_ONE || PM == P_TWO)) {
          // packed FP32 (SASS FADD2 / FMUL2: two IEEE-rn operations per instruction, each lane rounded exactly as the
          // scalar instruction rounds, so the sums stay bit-identical): the accumulators of two adjacent train rows
          // share a register pair, the query value is the instruction's scalar-broadcast operand
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              pair_step<PM == P_ONE>(acc[i][2 * jj], acc[i][2 * jj + 1], xb[2 * jj], xb[2 * jj + 1], ya[i]);   // BruteForceNn.h:68-79
            }
        } else
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if constexpr (std::is_same<T, float>::value) {
              const float d = __fsub_rn(xb[j], ya[i]);                        // BruteForceNn.h:68
              acc[i][j] = __fadd_rn(acc[i][j], term_f<PM>(d, P.p));          // :70-79
            } else {
              const float d = (float)(xb[j] - ya[i]);                         // int difference, then float (:68)
              if (PM == P_POW) acc[i][j] += __double2int_rz(pow((double)fabsf(d), P.p));   // Scalar pval truncates (:69)
              else acc[i][j] += __float2int_rz(term_f<PM>(d, P.p));
            }
          }
      }
      __syncthreads();
    }

    // ---- selection: candidates (dist <= the row's k-th best so far; the row owner compares whole keys)
    // go through a queue.  On a CTA's first train tile no row has a k-th best yet and all 128 x 128
    // pairs would queue up; instead each half-warp (the 16 threads sharing a row) first derives a bound:
    // the k-th smallest of its 16 per-thread minima is >= the row's k-th smallest distance.
    uint32_t thb[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) thb[i] = thr[ty * 8 + i];
    if (!FLOOR && t == t0 && k <= 16) {
      const int lane = tid & 31, hb = lane & 16;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint32_t mn = THR_NONE;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (x0 + xrow(tx, j) < xcnt) mn = min(mn, dist_bits(acc[i][j]));
        int rank = 0;
#pragma unroll
        for (int o = 0; o < 16; ++o) {
          const uint32_t v = __shfl_sync(0xFFFFFFFFu, mn, hb | o);
          rank += (v < mn) || (v == mn && o < tx);
        }
        const unsigned has = (__ballot_sync(0xFFFFFFFFu, rank == k - 1) >> hb) & 0xFFFFu;   // exactly one lane
        const uint32_t bound = __shfl_sync(0xFFFFFFFFu, mn, hb | (__ffs(has) - 1));
        if (thb[i] == THR_NONE) thb[i] = bound;     // THR_NONE again when fewer than k threads hold a valid pair
      }
    }
    unsigned long long pend = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        bool c = (!FILT || ty * 8 + i < ycnt) && (x0 + xrow(tx, j) < xcnt) && dist_bits(acc[i][j]) <= thb[i];   // THR_NONE admits all
        if (FLOOR) {   // later passes of a k > LP_MAX_K search: only what the earlier passes have not delivered
          const int qr = y0 + ty * 8 + i;
          const unsigned long long fl = qr < P.yrows ? P.floor[qr] : KEY_NONE;
          c = c && ((((unsigned long long)dist_bits(acc[i][j]) << 32) | (uint32_t)(x0 + xrow(tx, j))) > fl);
        }
        pend |= (unsigned long long)c << (i * 8 + j);
      }
    }
    while (__syncthreads_or(pend != 0)) {
      if (tid == 0) qn[0] = 0;
      __syncthreads();
      {   // one queue reservation per warp
        const int lane = tid & 31, mine = __popcll(pend);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
          if (lane >= o) incl += v;
        }
        int base = 0;
        if (lane == 31 && incl) base = atomicAdd(&qn[0], incl);
        int slot = __shfl_sync(0xFFFFFFFFu, base, 31) + incl - mine;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if ((pend >> (i * 8 + j)) & 1) {
              if (slot < QCAP) {
                qkey[slot] = ((unsigned long long)dist_bits(acc[i][j]) << 32) |
                             (FILT ? xid[xrow(tx, j)] : (uint32_t)(x0 + xrow(tx, j)));
                qrow[slot] = (unsigned short)(ty * 8 + i);
                pend &= ~(1ull << (i * 8 + j));
              }
              ++slot;
            }
      }
      __syncthreads();
      const int n = min(qn[0], QCAP);
      if (tid < TB) {               // thread r owns row r's list
        unsigned long long* L = list + (size_t)tid * k;
        int m = cnt[tid];
        for (int e = 0; e < n; ++e) {
          if (qrow[e] != tid) continue;
          const unsigned long long key = qkey[e];
          if (m == k) {
            if (key >= L[k - 1]) continue;
            --m;                     // drop the current k-th
          }
          int pos = m;
          while (pos > 0 && L[pos - 1] > key) { L[pos] = L[pos - 1]; --pos; }
          L[pos] = key;
          ++m;
        }
        cnt[tid] = m;
        thr[tid] = (m == k) ? (uint32_t)(L[k - 1] >> 32) : THR_NONE;
      }
      __syncthreads();
      // what is still pending may have been beaten meanwhile
      if (pend) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t th = thr[ty * 8 + i];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (((pend >> (i * 8 + j)) & 1) && dist_bits(acc[i][j]) > th) pend &= ~(1ull << (i * 8 + j));
        }
      }
    }
  }

  // publish this split's sorted lists
  if (tid < TB && tid < ycnt) {
    unsigned long long* out = P.partial + ((size_t)blockIdx.y * prow + y0 + tid) * k;
    const int m = cnt[tid];
    for (int q = 0; q < k; ++q) out[q] = q < m ? list[(size_t)tid * k + q] : KEY_NONE;
  }
}

// one thread per query row: S sorted lists -> the k smallest keys, written in the
// reference's layout: ascending, RIGHT-aligned when fewer than k exist
// (BruteForceNn.h:111-118 fills from column k-1 downwards; the columns before
// keep the caller's zeros)
// ysel != nullptr (filtered): list position `row` belongs to output row ysel[row]; a row's unfilled
// leading columns and every row outside the filter keep the caller's values (idx / dist are in-out).
// Passes (k > LP_MAX_K): this pass's k entries go to columns col0 .. of a kout-wide output whose row holds
// ntotal entries in all (right-aligned as a whole), and the last key delivered becomes the next pass's floor.
template <typename D>
__global__ void lp_merge_kernel(const unsigned long long* __restrict__ partial, int splits, int yrows, int k,
                                uint64_t* __restrict__ idx, D* __restrict__ dist,
                                const uint32_t* __restrict__ ysel = nullptr, int kout = 0, int col0 = -1,
                                unsigned long long* __restrict__ floor_out = nullptr) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= yrows) return;
  const size_t orow = ysel ? ysel[row] : (uint32_t)row;
  if (kout <= 0) kout = k;
  // count what exists (at most k are needed)
  int total = 0;
  for (int s = 0; s < splits && total < k; ++s) {
    const unsigned long long* L = partial + ((size_t)s * yrows + row) * k;
    for (int q = 0; q < k && L[q] != KEY_NONE; ++q) ++total;
  }
  const int n = total < k ? total : k;
  unsigned long long last = 0;
  bool have_last = false;
  for (int out = 0; out < n; ++out) {
    // smallest key strictly greater than the previous pick (keys are distinct: idx is unique)
    unsigned long long best = KEY_NONE;
    for (int s = 0; s < splits; ++s) {
      const unsigned long long* L = partial + ((size_t)s * yrows + row) * k;
      for (int q = 0; q < k; ++q) {
        const unsigned long long v = L[q];
        if (v == KEY_NONE) break;
        if (have_last && v <= last) continue;
        if (v < best) best = v;
        break;   // lists ascend: the first admissible entry is this list's minimum
      }
    }
    last = best; have_last = true;
    const int col = col0 >= 0 ? col0 + out : k - n + out;
    idx[orow * kout + col] = best & 0xFFFFFFFFull;
    const uint32_t b = (uint32_t)(best >> 32);
    if constexpr (std::is_same<D, float>::value) dist[orow * kout + col] = __uint_as_float(b);
    else dist[orow * kout + col] = (int32_t)b;
  }
  if (floor_out) floor_out[row] = (n == k) ? last : KEY_NONE;
}

size_t lp_smem_bytes(int k) {
  return (size_t)2 * KC * PITCH * 4 + TB * 4 + TB * 4 + QCAP * 8 + QCAP * 2 + 16 + 2 * TB * 4 + (size_t)TB * k * 8;
}

template <typename T, int PM, bool FILT, bool FLOOR>
cudaError_t launch_lp(const LpParams& P, dim3 grid, size_t smem, cudaStream_t stream) {
  cudaError_t e = cudaFuncSetAttribute(lp_topk_kernel<T, PM, FILT, FLOOR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  lp_topk_kernel<T, PM, FILT, FLOOR><<<grid, NT, smem, stream>>>(P);
  count_launch();
  return cudaGetLastError();
}

template <typename T, bool FILT = false, bool FLOOR = false>
cudaError_t launch_lp_p(const LpParams& P, dim3 grid, size_t smem, cudaStream_t stream) {
  if (P.p == 1.0) return launch_lp<T, P_ONE, FILT, FLOOR>(P, grid, smem, stream);     // BruteForceNn.h:70-78: exact comparisons on m_p
  if (P.p == 2.0) return launch_lp<T, P_TWO, FILT, FLOOR>(P, grid, smem, stream);
  if (P.p == 0.5) return launch_lp<T, P_HALF, FILT, FLOOR>(P, grid, smem, stream);
  return launch_lp<T, P_POW, FILT, FLOOR>(P, grid, smem, stream);
}

}  // namespace

// Train-side splits: CTAs are (query tile, split) pairs and a CTA runs its train tiles in sequence, so
// time ~ waves of CTAs x train tiles per CTA (+ a fixed cost per CTA: query staging, bound bootstrap,
// publishing).  Pick the split count that minimises it -- 157 query tiles on 296 CTA slots are 1 wave
// of 157 train tiles unsplit, but 8 full waves of 11 at 15 splits -- under a cap on the partial-list scratch.
static int pick_splits(int ny_tiles, int nxt, int sms, size_t rows_times_k) {
  if (ny_tiles < 1 || nxt < 1) return 1;
  int smax = nxt < 64 ? nxt : 64;
  const size_t cap = ((size_t)1 << 27) / (rows_times_k ? rows_times_k : 1);   // <= 1 GiB of 8-byte keys
  if ((size_t)smax > cap) smax = cap < 1 ? 1 : (int)cap;
  int best = 1;
  double best_cost = 1e300;
  const int slots = sms * 2;   // __launch_bounds__(NT, 2): two CTAs per SM
  for (int s = 1; s <= smax; ++s) {
    const long long ctas = (long long)ny_tiles * s;
    const long long waves = (ctas + slots - 1) / slots;
    const int per = (nxt + s - 1) / s;
    const double cost = (double)waves * (per + 0.5);
    if (cost < best_cost * 0.98) { best_cost = cost; best = s; }
  }
  return best;
}

int lp_splits(int xrows, int yrows, int k, int sms) {
  return pick_splits((yrows + TB - 1) / TB, (xrows + TB - 1) / TB, sms, (size_t)(yrows > 0 ? yrows : 0) * k);
}

// k > LP_MAX_K runs in passes of LP_MAX_K (the sorted per-row lists live in shared memory); the scratch is one
// pass's partial lists plus one floor key per query row
size_t lp_partial_bytes(int xrows, int yrows, int k, int sms) {
  const int kp = k < LP_MAX_K ? k : LP_MAX_K;
  const size_t lists = (size_t)lp_splits(xrows, yrows, kp, sms) * (size_t)yrows * kp * sizeof(unsigned long long);
  return lists + (k > LP_MAX_K ? (size_t)(yrows > 0 ? yrows : 0) * sizeof(unsigned long long) : 0);
}

size_t lp_ws_bytes(int xrows, int yrows, int dim, int k, int sms) {
  const size_t a = lp_partial_bytes(xrows, yrows, k, sms) + 8, b = lp_tma_ws_bytes(xrows, yrows, dim, k, sms);
  return a > b ? a : b;
}

cudaError_t lp_topk(const void* dx, const void* dy, bool is_int, int xrows, int yrows, int dim, int k, double p,
                    void* ws, size_t ws_bytes, uint64_t* didx, void* ddist, cudaStream_t stream) {
  unsigned long long* partial = static_cast<unsigned long long*>(ws);
  if (yrows <= 0) return cudaSuccess;
  if (k < 1 || dim < 1) return cudaErrorInvalidValue;
  // the caller's arrays start as zeros (numpy), and unfilled columns stay zero
  cudaError_t e = cudaMemsetAsync(didx, 0, (size_t)yrows * k * sizeof(uint64_t), stream);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(ddist, 0, (size_t)yrows * k * 4, stream);
  if (e != cudaSuccess || xrows <= 0) return e;
  int dev = 0, sms = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  if (!is_int) {
    // float32 rows: the TMA-fed kernel (lp_topk_tma.cu) when the caller's scratch admits its transposed copies
    const char* force = getenv("SPECTAVI_K1F");
    const size_t need = lp_tma_ws_bytes(xrows, yrows, dim, k, sms);
    if (!(force && strcmp(force, "v1") == 0) && need && ws_bytes >= need) {
      unsigned long long* tp = nullptr;
      int tsplits = 1;
      e = lp_topk_tma((const float*)dx, (const float*)dy, xrows, yrows, dim, k, p, ws, ws_bytes, sms, &tp, &tsplits, stream);
      if (e != cudaSuccess) return e;
      lp_merge_kernel<float><<<(yrows + 127) / 128, 128, 0, stream>>>(tp, tsplits, yrows, k, didx, (float*)ddist);
      count_launch();
      return cudaGetLastError();
    }
  }
  if (ws_bytes < lp_partial_bytes(xrows, yrows, k, sms)) return cudaErrorInvalidValue;
  LpParams P;
  P.x = dx; P.y = dy; P.xrows = xrows; P.yrows = yrows; P.dim = dim; P.k = k; P.p = p;
  P.nxt = (xrows + TB - 1) / TB;
  P.partial = partial;
  P.ysel = P.xsel = nullptr; P.tiles = nullptr; P.nysel = 0; P.floor = nullptr;
  if (k <= LP_MAX_K) {
    P.splits = lp_splits(xrows, yrows, k, sms);
    const dim3 grid((yrows + TB - 1) / TB, P.splits);
    const size_t smem = lp_smem_bytes(k);
    e = is_int ? launch_lp_p<int>(P, grid, smem, stream) : launch_lp_p<float>(P, grid, smem, stream);
    if (e != cudaSuccess) return e;
    if (is_int) lp_merge_kernel<int32_t><<<(yrows + 127) / 128, 128, 0, stream>>>(partial, P.splits, yrows, k, didx, (int32_t*)ddist);
    else lp_merge_kernel<float><<<(yrows + 127) / 128, 128, 0, stream>>>(partial, P.splits, yrows, k, didx, (float*)ddist);
    count_launch();
    return cudaGetLastError();
  }
  // k > LP_MAX_K (the reference takes any k, BruteForceNn.h:93-118): passes of LP_MAX_K over the whole train set,
  // each admitting only keys above the last one the previous pass delivered.  A row holds min(k, xrows) entries,
  // right-aligned as a whole (:111-118).
  const int ntotal = k < xrows ? k : xrows;
  P.splits = lp_splits(xrows, yrows, LP_MAX_K, sms);
  unsigned long long* floor = partial + (size_t)P.splits * yrows * LP_MAX_K;
  const dim3 grid((yrows + TB - 1) / TB, P.splits);
  for (int done = 0; done < ntotal; done += LP_MAX_K) {
    const int kp = ntotal - done < LP_MAX_K ? ntotal - done : LP_MAX_K;
    P.k = kp;
    P.floor = done ? floor : nullptr;
    const size_t smem = lp_smem_bytes(kp);
    if (done == 0) e = is_int ? launch_lp_p<int>(P, grid, smem, stream) : launch_lp_p<float>(P, grid, smem, stream);
    else e = is_int ? launch_lp_p<int, false, true>(P, grid, smem, stream) : launch_lp_p<float, false, true>(P, grid, smem, stream);
    if (e != cudaSuccess) return e;
    const int col0 = k - ntotal + done;
    if (is_int) lp_merge_kernel<int32_t><<<(yrows + 127) / 128, 128, 0, stream>>>(partial, P.splits, yrows, kp, didx, (int32_t*)ddist, nullptr, k, col0, floor);
    else lp_merge_kernel<float><<<(yrows + 127) / 128, 128, 0, stream>>>(partial, P.splits, yrows, kp, didx, (float*)ddist, nullptr, k, col0, floor);
    count_launch();
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// ---- mu != 0: the reference's approximate search, emulated operation for operation ----------------------------
// BruteForceNn.h:80-86 prunes a train row as soon as `distp + mu * (dim - icol - 1) > worst`: an early exit that,
// for mu != 0, also drops rows that WOULD have entered the top k -- and which rows depends on the order they were
// visited in and on the running worst distance, so the only way to return the reference's result is to walk the
// train rows in its order.  One thread per query row, sequential over train rows and columns, the k best in a
// sorted array (only the heap's top and size are ever observed, oracle/spv_oracle.c); float accumulation with
// unfused IEEE operations.  A search of n x n rows costs n * dim sequential steps per thread -- fine for the shapes
// an approximation knob is used on, and the exact matcher (mu == 0) never comes here.
template <typename T>
__global__ void lp_topk_mu_kernel(const T* __restrict__ x, const T* __restrict__ y, int xrows, int yrows, int dim, int k, double p,
                                  float mu_f, unsigned long long* __restrict__ sets, uint64_t* __restrict__ idx,
                                  T* __restrict__ dist) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= yrows) return;
  unsigned long long* S = sets + (size_t)r * (k + 1);     // ascending (dist bits << 32 | idx); dist >= 0 so bits order = value order
  const T* yr = y + (size_t)r * dim;
  const int pm = (p == 1.0) ? P_ONE : (p == 2.0) ? P_TWO : (p == 0.5) ? P_HALF : P_POW;
  const int mu_i = (int)mu_f;                             // Scalar m_mu with Scalar = int (BruteForceNn.h:28)
  int n = 0;
  bool have_worst = false;
  T worst = (T)0;
  for (int c = 0; c < xrows; ++c) {
    const T* xr = x + (size_t)c * dim;
    T acc = (T)0;
    bool pruned = false;
    for (int j = 0; j < dim; ++j) {
      if constexpr (std::is_same<T, float>::value) {
        const float d = __fsub_rn(xr[j], yr[j]);
        float t;
        if (pm == P_ONE) t = fabsf(d);
        else if (pm == P_TWO) t = __fmul_rn(d, d);
        else if (pm == P_HALF) t = sqrtf(fabsf(d));
        else t = (float)pow((double)fabsf(d), p);
        acc = __fadd_rn(acc, t);
        if (have_worst && __fadd_rn(acc, __fmul_rn(mu_f, (float)(dim - (j + 1)))) > worst) { pruned = true; break; }
      } else {
        const float d = (float)(xr[j] - yr[j]);
        if (pm == P_POW) acc += __double2int_rz(pow((double)fabsf(d), p));
        else acc += __float2int_rz(pm == P_ONE ? fabsf(d) : pm == P_TWO ? __fmul_rn(d, d) : sqrtf(fabsf(d)));
        if (have_worst && acc + mu_i * (dim - (j + 1)) > worst) { pruned = true; break; }
      }
    }
    if (pruned) continue;
    // :93-108 (a candidate that is not strictly better than a full heap's worst is rejected: the reference's
    // push-without-pop there is undefined behaviour)
    if (have_worst && !(acc < worst)) continue;
    if (have_worst && n >= k) --n;                        // pop the largest
    const unsigned long long key = ((unsigned long long)dist_bits(acc) << 32) | (uint32_t)c;
    int pos = n;
    while (pos > 0 && S[pos - 1] > key) { S[pos] = S[pos - 1]; --pos; }
    S[pos] = key;
    ++n;
    if (have_worst || n == k) {
      have_worst = true;
      const uint32_t b = (uint32_t)(S[n - 1] >> 32);
      if constexpr (std::is_same<T, float>::value) worst = __uint_as_float(b);
      else worst = (int)b;
    }
  }
  for (int q = 0; q < n; ++q) {                           // :111-118, right-aligned
    const int col = k - n + q;
    idx[(size_t)r * k + col] = S[q] & 0xFFFFFFFFull;
    const uint32_t b = (uint32_t)(S[q] >> 32);
    if constexpr (std::is_same<T, float>::value) dist[(size_t)r * k + col] = __uint_as_float(b);
    else dist[(size_t)r * k + col] = (int)b;
  }
}

size_t lp_mu_ws_bytes(int yrows, int k) { return (size_t)(yrows > 0 ? yrows : 0) * (size_t)(k + 1) * sizeof(unsigned long long) + 8; }

cudaError_t lp_topk_mu(const void* dx, const void* dy, bool is_int, int xrows, int yrows, int dim, int k, double p, float mu,
                       void* ws, size_t ws_bytes, uint64_t* didx, void* ddist, cudaStream_t stream) {
  if (yrows <= 0) return cudaSuccess;
  if (k < 1 || dim < 1 || ws_bytes < lp_mu_ws_bytes(yrows, k)) return cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(didx, 0, (size_t)yrows * k * sizeof(uint64_t), stream);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(ddist, 0, (size_t)yrows * k * 4, stream);
  if (e != cudaSuccess || xrows <= 0) return e;
  unsigned long long* sets = static_cast<unsigned long long*>(ws);
  if (is_int) lp_topk_mu_kernel<int><<<(yrows + 63) / 64, 64, 0, stream>>>((const int*)dx, (const int*)dy, xrows, yrows, dim, k, p, mu, sets, didx, (int*)ddist);
  else lp_topk_mu_kernel<float><<<(yrows + 63) / 64, 64, 0, stream>>>((const float*)dx, (const float*)dy, xrows, yrows, dim, k, p, mu, sets, didx, (float*)ddist);
  count_launch();
  return cudaGetLastError();
}

// ---- the filter overload, batched over groups (see the header)
int lp_filtered_tile_rows() { return TB; }

int lp_filtered_splits(int ntiles, int max_xcnt, int nysel, int k, int sms) {
  return pick_splits(ntiles, (max_xcnt + TB - 1) / TB, sms, (size_t)(nysel > 0 ? nysel : 0) * k);
}

size_t lp_filtered_partial_bytes(int ntiles, int max_xcnt, int nysel, int k, int sms) {
  return (size_t)lp_filtered_splits(ntiles, max_xcnt, nysel, k, sms) * (size_t)nysel * k * sizeof(unsigned long long);
}

cudaError_t lp_topk_filtered(const float* dx, const float* dy, int dim, int k, double p, const int4* dtiles, int ntiles,
                             const uint32_t* dysel, int nysel, const uint32_t* dxsel, int max_xcnt,
                             unsigned long long* partial, uint64_t* didx, float* ddist, cudaStream_t stream) {
  if (ntiles <= 0 || nysel <= 0) return cudaSuccess;
  if (k < 1 || k > LP_MAX_K || dim < 1) return cudaErrorInvalidValue;
  int dev = 0, sms = 0;
  cudaError_t e;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  LpParams P;
  P.x = dx; P.y = dy; P.xrows = 0; P.yrows = 0; P.dim = dim; P.k = k; P.p = p; P.nxt = 0;
  P.splits = lp_filtered_splits(ntiles, max_xcnt, nysel, k, sms);
  P.partial = partial;
  P.ysel = dysel; P.xsel = dxsel; P.tiles = dtiles; P.nysel = nysel; P.floor = nullptr;
  const dim3 grid(ntiles, P.splits);
  if ((e = launch_lp_p<float, true>(P, grid, lp_smem_bytes(k), stream)) != cudaSuccess) return e;
  lp_merge_kernel<float><<<(nysel + 127) / 128, 128, 0, stream>>>(partial, P.splits, nysel, k, didx, ddist, dysel);
  count_launch();
  return cudaGetLastError();
}

}  // namespace spv
