// lp_topk_tma.cu -- K1f, the TMA-fed form of lp_topk_kernel for float32 rows (the dense matcher: no filter
// lists, k <= LP_MAX_K in one pass).  Same arithmetic as lp_topk.cu, operation for operation (BruteForceNn.h:61-108:
// float difference, |d| / d*d / sqrtf|d| / pow, terms added one by one in column order), so distances stay
// BIT-EXACT; what changes is how the operands reach the FP32 pipe.
//
// lp_topk_kernel stages 16-column panels through registers (16 LDG.32 + 16 transposing STS.32 per thread and
// panel, two __syncthreads per panel, the stores 2-way bank-conflicting) at two CTAs of 128 registers per SM.
// ncu r01: fma pipe 63 %, stall_barrier 0.76 and short_scoreboard 0.41 per issue, 2.4e8 shared-memory bank
// conflicts, ~170 B of spills.  Here -- K1's skeleton (l1_top2.cu):
//   * a transposition pre-pass writes both matrices column-major ([dim][rows padded], a few microseconds of HBM
//     time) so that ONE cp.async.bulk.tensor.2d per operand drops a 16-column x 128-row panel into shared memory
//     already in the [column][row] order the register tile reads (TMA cannot transpose 4-byte elements, and the
//     packed FADD2 form needs the accumulators of ADJACENT ROWS of one column in one register pair);
//   * one producer warp issues the TMA loads into a ring of (x panel, y panel) stages with full / empty
//     mbarriers; the 8 compute warps never touch global memory, never execute __syncthreads in the column loop,
//     and read operands as conflict-free LDS.128 (16 lanes x 16 B contiguous on the train side, broadcast on the
//     query side);
//   * one CTA of 9 warps per SM: up to 224 registers per thread, no spills;
//   * ragged edges cost nothing: TMA zero-fills rows / columns outside the matrix, a zero column adds +0.0 to
//     every accumulator (exact), padded rows are masked where candidates are selected.
// Selection (per-row sorted lists, first-tile bound) is warp-local: a warp owns its 16 query rows outright, so there is
// no CTA-wide barrier anywhere after start-up and the warps drift apart like K1's.
#include "spv_common.h"

#include <cuda.h>
#include <cstdint>

namespace spv {
namespace {

constexpr int TB = 128;          // query rows and train rows per CTA tile
constexpr int KC = 32;           // columns per stage (32 measured against 16 below)
constexpr int NCW = 8;           // compute warps
constexpr int NCT = NCW * 32;    // compute threads
constexpr int NT = NCT + 32;     // + the producer warp
constexpr uint32_t THR_NONE = 0xFFFFFFFFu;
constexpr int STAGE_FLOATS = 2 * KC * TB;   // x panel then y panel, [col][row]
constexpr int LP_TMA_MAX_K = 64;            // lists (1 KB per k) + 64 KB of parked tiles + >= 3 stages must fit 200 KB

enum PMode { P_POW = 0, P_ONE = 1, P_TWO = 2, P_HALF = 3 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity, uint32_t hint_ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(hint_ns)
      : "memory");
  return ok != 0;
}
// bounded: a protocol bug must trap, never hang the device
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t timeouts = 0;
  while (!mbar_try_wait(bar, parity, 20000u)) {
    if (++timeouts > (1u << 19)) {
      printf("lp_topk_tma: mbarrier wait timed out (block %d,%d thread %d)\n", blockIdx.x, blockIdx.y, threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ void tma_load_panel(uint32_t dst, const CUtensorMap* map, int row0, int col0, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(row0), "r"(col0), "r"(bar)
      : "memory");
}
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

// train rows of a thread's register tile, as in lp_topk.cu: columns 0..3 are tile rows 4 tx + j, 4..7 are 64 + 4 tx + (j - 4)
__device__ __forceinline__ int xrow(int tx, int j) { return (j & 4) * 16 + tx * 4 + (j & 3); }

// rows x dim (row-major) -> dim x pitch (column-major, pitch a multiple of 4 floats so that TMA's 16-byte
// stride rule holds); 32 x 32 tiles through shared memory, both sides coalesced
__global__ void transpose_f32_kernel(const float* __restrict__ in, int rows, int dim, float* __restrict__ out, size_t pitch) {
  __shared__ float t[32][33];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    t[i][threadIdx.x] = (r < rows && c < dim) ? in[(size_t)r * dim + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < dim && r < rows) out[(size_t)c * pitch + r] = t[threadIdx.x][i];
  }
}

struct TmaParams {
  int xrows, yrows, dim, k;
  double p;
  int nxt, splits, nst;
  unsigned long long* partial;   // [splits][yrows][k]
  uint32_t* gthr;                // [yrows] smallest k-th-best distance (bits) any train split has reached so far
};

template <int PM>
__global__ void __launch_bounds__(NT, 1)
lp_topk_tma_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmy, const TmaParams P) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int nst = P.nst, k = P.k, dim = P.dim;
  float* ring = reinterpret_cast<float*>(smem_raw);                                  // [nst][x panel | y panel][KC][TB]
  uint32_t* thr = reinterpret_cast<uint32_t*>(ring + (size_t)nst * STAGE_FLOATS);   // [TB] k-th best distance bits
  int* cnt = reinterpret_cast<int*>(thr + TB);                                       // [TB] entries in each row's list
  uint32_t* teff = reinterpret_cast<uint32_t*>(cnt + TB);                            // [TB] min(own, shared) threshold of the tile
  unsigned long long* list = reinterpret_cast<unsigned long long*>(teff + TB);       // [TB][k] ascending
  float4* scratch = reinterpret_cast<float4*>(list + (size_t)TB * k);                // [NCW][16][32]: parked register tiles
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(scratch + NCW * 512);   // full[nst], empty[nst]
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * nst;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, NCW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const int y0 = blockIdx.x * TB, ycnt = min(TB, P.yrows - y0), xcnt = P.xrows;
  const int t0 = (int)((long long)P.nxt * blockIdx.y / P.splits);
  const int t1 = (int)((long long)P.nxt * (blockIdx.y + 1) / P.splits);
  const int npanels = (dim + KC - 1) / KC;

  if (warp == NCW) {
    // ===================== producer: one thread issues all TMA loads, train tile after train tile, panel after panel
    if (lane == 0) {
      const uint32_t ring_s = smem_u32(ring);
      int stage = 0;
      uint32_t phase = 0;
      for (int t = t0; t < t1; ++t)
        for (int pn = 0; pn < npanels; ++pn) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          mbar_expect_tx(bar_full + 8 * stage, STAGE_FLOATS * 4);
          const uint32_t dst = ring_s + (uint32_t)stage * (STAGE_FLOATS * 4);
          tma_load_panel(dst, &tmx, t * TB, pn * KC, bar_full + 8 * stage);
          tma_load_panel(dst + KC * TB * 4, &tmy, y0, pn * KC, bar_full + 8 * stage);
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
    }
    return;
  }

  // ===================== compute warps
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  if (lane < 16) { thr[warp * 16 + lane] = THR_NONE; cnt[warp * 16 + lane] = 0; }   // this warp's rows
  __syncwarp();
  int stage = 0;
  uint32_t phase = 0;

  for (int t = t0; t < t1; ++t) {
    const int x0 = t * TB;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    // The train set is split over gridDim.y CTAs per query tile, each scanning its share cold: without help every
    // split re-pays the threshold warm-up (ncu: ~8 candidates per warp and tile at 10 splits where one scan would
    // produce ~1).  So the splits share, per query row, the smallest k-th-best distance any of them has reached
    // (a candidate farther than that cannot be among the final k; ties go through and the merge compares whole keys).
    // The load is issued here and consumed after the column loop.
    uint32_t gth = THR_NONE;
    if (lane < 16 && warp * 16 + lane < ycnt) gth = __ldcg(P.gthr + y0 + warp * 16 + lane);

    for (int pn = 0; pn < npanels; ++pn) {
      mbar_wait(bar_full + 8 * stage, phase);
      const float* xs = ring + (size_t)stage * STAGE_FLOATS;
      const float* ys = xs + KC * TB;
      // (columns past dim are TMA zero fill and would add +0.0, but a constant trip count was measured slower: at
      // dim = 132 the last panel holds 4 real columns, and 12 idle ones are 8 % more arithmetic)
      const int kc = min(KC, dim - pn * KC);
#pragma unroll 4
      for (int c = 0; c < kc; ++c) {
        float ya[8], xb[8];
        const float4* yp = reinterpret_cast<const float4*>(ys + c * TB + ty * 8);
        const float4* xp = reinterpret_cast<const float4*>(xs + c * TB + tx * 4);
        *reinterpret_cast<float4*>(&ya[0]) = yp[0]; *reinterpret_cast<float4*>(&ya[4]) = yp[1];
        *reinterpret_cast<float4*>(&xb[0]) = xp[0]; *reinterpret_cast<float4*>(&xb[4]) = xp[TB / 8];   // + 64 rows
        if (PM == P_ONE || PM == P_TWO) {
          // packed FP32 (FADD2 / FMUL2): two adjacent train rows per register pair, the query value broadcast
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              pair_step<PM == P_ONE>(acc[i][2 * jj], acc[i][2 * jj + 1], xb[2 * jj], xb[2 * jj + 1], ya[i]);   // BruteForceNn.h:68-79
            }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float d = __fsub_rn(xb[j], ya[i]);
              acc[i][j] = __fadd_rn(acc[i][j], term_f<PM>(d, P.p));
            }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8 * stage);   // this warp has read the panels: the stage may be refilled
      if (++stage == nst) { stage = 0; phase ^= 1; }
    }

    // ---- selection, warp-local and COMPACT.  A warp owns 16 query rows outright (two groups of 16 lanes, 8 rows each,
    // every lane 8 train rows), so a row's candidates all come from ONE warp and its sorted list needs no CTA-wide
    // barrier: warps drift apart -- one selects while the others accumulate.  The common case costs ~80 instructions:
    // per row the minimum of the lane's 8 distances against the row's k-th best.  Only when some lane has a hit does
    // the warp park its accumulators in shared memory and walk the candidates in a small dynamic loop (one ballot bit
    // at a time, the group's first lane inserting into the row's list).  (First form of this kernel: CTA-wide queue +
    // 4 barriers per tile, slower than the register-staged kernel; second: fully unrolled per-(i, j) ballots, 2 300
    // instructions per tile and 10 k instructions of code -- ncu: 36 % of the compute warps' time outside the FADD2
    // loop, stall_no_instruction 15 %.)
    if (lane < 16) teff[warp * 16 + lane] = min(thr[warp * 16 + lane], gth);
    __syncwarp();
    uint32_t thb[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) thb[i] = teff[ty * 8 + i];
    const bool xin = x0 + TB <= xcnt;                    // no zero-filled train rows in this tile (CTA-uniform)
    if (t == t0 && k <= 16 && xin) {
      // first train tile: no row has a k-th best yet; the k-th smallest of the 16 per-lane minima of a row bounds
      // its k-th distance from above
      const int hb = lane & 16;
#pragma unroll 1
      for (int i = 0; i < 8; ++i) {
        float mf = acc[0][0];
#pragma unroll
        for (int ii = 0; ii < 8; ++ii)
          if (ii == i) {
            mf = fminf(fminf(fminf(acc[ii][0], acc[ii][1]), fminf(acc[ii][2], acc[ii][3])),
                       fminf(fminf(acc[ii][4], acc[ii][5]), fminf(acc[ii][6], acc[ii][7])));
          }
        const uint32_t mn = __float_as_uint(mf);
        int rank = 0;
#pragma unroll 4
        for (int o = 0; o < 16; ++o) {
          const uint32_t v = __shfl_sync(0xFFFFFFFFu, mn, hb | o);
          rank += (v < mn) || (v == mn && o < tx);
        }
        const unsigned has = (__ballot_sync(0xFFFFFFFFu, rank == k - 1) >> hb) & 0xFFFFu;   // exactly one lane
        const uint32_t bound = __shfl_sync(0xFFFFFFFFu, mn, hb | (__ffs(has) - 1));
#pragma unroll
        for (int ii = 0; ii < 8; ++ii)
          if (ii == i && thb[ii] == THR_NONE) thb[ii] = bound;
      }
    }
    unsigned rowhit = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float m = fminf(fminf(fminf(acc[i][0], acc[i][1]), fminf(acc[i][2], acc[i][3])),
                            fminf(fminf(acc[i][4], acc[i][5]), fminf(acc[i][6], acc[i][7])));
      rowhit |= (unsigned)(__float_as_uint(m) <= thb[i]) << i;          // THR_NONE admits all; NaN sums never qualify
    }
    if (__any_sync(0xFFFFFFFFu, rowhit != 0)) {
      // park the register tile: element e = 8 i + j of lane l at float4 slot (e / 4) * 32 + l (conflict-free)
      float4* sc = scratch + warp * 512;
#pragma unroll
      for (int c = 0; c < 16; ++c)
        sc[c * 32 + lane] = make_float4(acc[c >> 1][(c & 1) * 4], acc[c >> 1][(c & 1) * 4 + 1], acc[c >> 1][(c & 1) * 4 + 2],
                                        acc[c >> 1][(c & 1) * 4 + 3]);
      __syncwarp();
      const float* scf = reinterpret_cast<const float*>(sc);
      // which of the lane's pairs qualify: per ROW a warp-uniform test first (typically one or two rows of a tile have a
      // hit anywhere in the warp), then 8 register compares; zero-filled train rows only exist in the last tile
      unsigned jvalid = 0xFFu;
      if (!xin) {
        jvalid = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) jvalid |= (unsigned)(x0 + xrow(tx, j) < xcnt) << j;
      }
      unsigned long long pend = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (!__any_sync(0xFFFFFFFFu, (rowhit >> i) & 1u)) continue;
        unsigned mk = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) mk |= (unsigned)(__float_as_uint(acc[i][j]) <= thb[i]) << j;
        if (ty * 8 + i >= ycnt) mk = 0;
        pend |= (unsigned long long)(mk & jvalid) << (8 * i);
      }
#pragma unroll 1
      for (;;) {
        const unsigned b = __ballot_sync(0xFFFFFFFFu, pend != 0);
        if (!b) break;
        const int src = __ffs(b) - 1;
        const int e = __shfl_sync(0xFFFFFFFFu, __ffsll((long long)pend) - 1, src);
        if (lane == src) pend &= pend - 1;
        if (lane == (src & 16)) {                       // the first lane of the candidate's 16-lane group owns the row's list
          const uint32_t v = __float_as_uint(scf[((e >> 2) * 32 + src) * 4 + (e & 3)]);
          const unsigned long long key = ((unsigned long long)v << 32) | (uint32_t)(x0 + xrow(src & 15, e & 7));
          const int row = (warp * 2 + (src >> 4)) * 8 + (e >> 3);
          unsigned long long* L = list + (size_t)row * k;
          int m = cnt[row];
          bool take = true;
          if (m == k) {
            if (key >= L[k - 1]) take = false;          // beaten meanwhile by an earlier candidate of this tile
            else --m;                                   // drop the current k-th
          }
          if (take) {
            int pos = m;
            while (pos > 0 && L[pos - 1] > key) { L[pos] = L[pos - 1]; --pos; }
            L[pos] = key;
            ++m;
            cnt[row] = m;
            if (m == k) {
              const uint32_t kth = (uint32_t)(L[k - 1] >> 32);
              thr[row] = kth;
              if (row < ycnt) atomicMin(P.gthr + y0 + row, kth);   // tell the other train splits
            }
          }
        }
      }
      __syncwarp();
    }
  }

  // publish this split's sorted lists: every warp its own 16 rows
  __syncwarp();
  if (lane < 16) {
    const int row = warp * 16 + lane;
    if (row < ycnt) {
      unsigned long long* out = P.partial + ((size_t)blockIdx.y * P.yrows + y0 + row) * k;
      const int m = cnt[row];
      for (int q = 0; q < k; ++q) out[q] = q < m ? list[(size_t)row * k + q] : KEY_NONE;
    }
  }
}

size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }
size_t pitch_of(int rows) { return ((size_t)(rows > 0 ? rows : 1) + 3) & ~(size_t)3; }

// CTAs are (query tile, train split) pairs, one per SM at a time: time ~ waves x train tiles per CTA (+ a fixed cost per CTA)
int pick_splits_tma(int ny_tiles, int nxt, int sms, size_t rows_times_k) {
  if (ny_tiles < 1 || nxt < 1) return 1;
  int smax = nxt < 64 ? nxt : 64;
  const size_t cap = ((size_t)1 << 27) / (rows_times_k ? rows_times_k : 1);   // <= 1 GiB of 8-byte keys
  if ((size_t)smax > cap) smax = cap < 1 ? 1 : (int)cap;
  int best = 1;
  double best_cost = 1e300;
  for (int s = 1; s <= smax; ++s) {
    const long long ctas = (long long)ny_tiles * s;
    const long long waves = (ctas + sms - 1) / sms;
    const int per = (nxt + s - 1) / s;
    const double cost = (double)waves * (per + 0.5);
    if (cost < best_cost * 0.98) { best_cost = cost; best = s; }
  }
  return best;
}

size_t fixed_smem(int k) { return (size_t)TB * 4 * 3 + (size_t)TB * k * 8 + (size_t)NCW * 512 * 16; }

template <int PM>
cudaError_t launch(const CUtensorMap& tmx, const CUtensorMap& tmy, const TmaParams& P, dim3 grid, size_t smem, cudaStream_t stream) {
  cudaError_t e = cudaFuncSetAttribute(lp_topk_tma_kernel<PM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  lp_topk_tma_kernel<PM><<<grid, NT, smem, stream>>>(tmx, tmy, P);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

int lp_tma_splits(int xrows, int yrows, int k, int sms) {
  return pick_splits_tma((yrows + TB - 1) / TB, (xrows + TB - 1) / TB, sms, (size_t)(yrows > 0 ? yrows : 0) * k);
}

size_t lp_tma_ws_bytes(int xrows, int yrows, int dim, int k, int sms) {
  if (k < 1 || k > LP_TMA_MAX_K || dim < 1) return 0;   // larger lists leave no room for the ring: register-staged kernel
  const size_t lists = (size_t)lp_tma_splits(xrows, yrows, k, sms) * (size_t)(yrows > 0 ? yrows : 0) * k * sizeof(unsigned long long);
  return al256(lists) + al256((size_t)dim * pitch_of(xrows) * 4) + al256((size_t)dim * pitch_of(yrows) * 4) +
         al256((size_t)(yrows > 0 ? yrows : 1) * 4) + 256;
}

// didx / ddist must already hold the caller's zeros (lp_topk does that); *partial_out / *splits_out tell the caller
// what to hand to the merge kernel
cudaError_t lp_topk_tma(const float* dx, const float* dy, int xrows, int yrows, int dim, int k, double p, void* ws,
                        size_t ws_bytes, int sms, unsigned long long** partial_out, int* splits_out, cudaStream_t stream) {
  if (ws_bytes < lp_tma_ws_bytes(xrows, yrows, dim, k, sms)) return cudaErrorInvalidValue;
  const int splits = lp_tma_splits(xrows, yrows, k, sms);
  const size_t xp = pitch_of(xrows), yp = pitch_of(yrows);
  unsigned char* w = static_cast<unsigned char*>(ws);
  unsigned long long* partial = reinterpret_cast<unsigned long long*>(w);
  w += al256((size_t)splits * yrows * k * sizeof(unsigned long long));
  float* xT = reinterpret_cast<float*>(w);
  w += al256((size_t)dim * xp * 4);
  float* yT = reinterpret_cast<float*>(w);
  w += al256((size_t)dim * yp * 4);
  uint32_t* gthr = reinterpret_cast<uint32_t*>(w);
  cudaError_t e = cudaMemsetAsync(gthr, 0xFF, (size_t)yrows * 4, stream);
  if (e != cudaSuccess) return e;
  const dim3 tb(32, 8);
  transpose_f32_kernel<<<dim3((xrows + 31) / 32, (dim + 31) / 32), tb, 0, stream>>>(dx, xrows, dim, xT, xp);
  transpose_f32_kernel<<<dim3((yrows + 31) / 32, (dim + 31) / 32), tb, 0, stream>>>(dy, yrows, dim, yT, yp);
  count_launch(2);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  CUtensorMap tmx, tmy;
  if (!make_f32_panel_map(&tmx, xT, xrows, dim, xp, TB, KC) || !make_f32_panel_map(&tmy, yT, yrows, dim, yp, TB, KC))
    return cudaErrorInvalidValue;
  TmaParams P;
  P.xrows = xrows; P.yrows = yrows; P.dim = dim; P.k = k; P.p = p;
  P.nxt = (xrows + TB - 1) / TB; P.splits = splits; P.partial = partial; P.gthr = gthr;
  const size_t fixed = fixed_smem(k) + 2 * 8 * 16 + 1024, stage = (size_t)STAGE_FLOATS * 4;
  int nst = (int)(((size_t)200 * 1024 - fixed) / stage);
  nst = nst > 10 ? 10 : nst;
  if (nst < 2) return cudaErrorInvalidValue;
  P.nst = nst;
  const size_t smem = (size_t)nst * stage + fixed_smem(k) + (size_t)2 * nst * 8;
  const dim3 grid((yrows + TB - 1) / TB, splits);
  if (p == 1.0) e = launch<P_ONE>(tmx, tmy, P, grid, smem, stream);        // BruteForceNn.h:70-78: exact comparisons on m_p
  else if (p == 2.0) e = launch<P_TWO>(tmx, tmy, P, grid, smem, stream);
  else if (p == 0.5) e = launch<P_HALF>(tmx, tmy, P, grid, smem, stream);
  else e = launch<P_POW>(tmx, tmy, P, grid, smem, stream);
  *partial_out = partial;
  *splits_out = splits;
  return e;
}

}  // namespace spv
