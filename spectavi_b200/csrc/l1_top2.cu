// l1_top2.cu -- K1: all-pairs L1 distance + per-query top-2 over uint8 rows,
// hand-written for sm_100a.  Replaces BruteForceNnL1K2::find_neighbours
// (reference src/BruteForceNnL1K2.h:84-145) and its SSE2 sad_16 (:43-48).
//
// Bound: the CUDA-core ALU pipe.  One pair of dim-byte rows costs dim/4
// VABSDIFF4.U8.ACC (packed 4 x u8 |a-b| accumulate, 64 thread-ops/clk/SM
// measured, pipe_peak.cu); everything else is arranged so that it does NOT
// issue on that pipe:
//   * operands: TMA (UTMALDG) stages 64-row train tiles and the 256-row query
//     tile into shared memory as [16-byte k-chunk][row][16 B], so that an
//     8 x 8 register tile per thread is fed by 16 conflict-free LDS.128 per
//     256 VABSDIFF4 (LSU pipe, 6 % of the spare issue slots);
//   * threshold test: each accumulator starts at -(2nd best distance of its
//     query row), so "candidate" == "accumulator ended negative"; the 64 sign
//     bits of a thread tile are folded with IDP.4A (fma pipe), one ISETP and
//     one VOTE per tile;
//   * the top-2 update itself is a rare slow path (about 2 ln(xrows) hits per
//     query row) that keeps one (best, second) key pair per query row in shared
//     memory, updated with 64-bit atomicMin.
// The reference's early-exit prune (:118-121) is not reproduced: it never
// changes the result (strict '>' there, strict '<' at :129,135), and divergent
// early exit loses on a GPU.
//
// Work decomposition: units of (256 query rows) x (64 train rows); every CTA
// owns a contiguous range of units (query-tile major), so the grid is one
// persistent CTA per SM balanced to +-1 unit.  A CTA folds its per-thread
// results into the global key buffer with two 64-bit atomicMin per key, so
// train-side splits (across CTAs, launches, or GPUs) need no extra pass.
#include "spv_common.h"

#include <cuda.h>
#include <cstdlib>
#include <atomic>
#include <mutex>

namespace spv {
namespace {

constexpr int BY = 256;        // query rows per CTA tile (8 warps x 32)
constexpr int BX = 64;         // train rows per stage   (8 lanes x 8)
constexpr int TY = 8, TX = 8;  // per-thread register tile
constexpr int NCW = 8;         // compute warps
constexpr int NTHREADS = (NCW + 1) * 32;  // + 1 producer warp
constexpr uint32_t FULL = 0xFFFFFFFFu;
constexpr uint32_t THR_MAX = 0x7FFFFFFFu;  // INT_MAX, BruteForceNnL1K2.h:99-103

// ------------------------------------------------------------------ PTX glue
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// try_wait with a suspend-time hint: the hardware parks the warp until the phase flips or the hint (ns) runs
// out, so a waiting warp issues nothing.  Every issued instruction of a waiting warp -- even the one-lane
// ISETP / VIADD of a spin counter -- takes a 2-cycle slot of the half-rate ALU pipe the VABSDIFF4 stream is
// bound by: ncu r01 showed the producer's poll loop (its NANOSLEEP returned after ~12 ns) executing ~2 400
// warp-instructions per tile on its scheduler.  Parking it is worth 0.25 % (517.4 -> 518.7 G pairs/s).
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
// Bounded wait: a protocol bug must trap, never hang the device (2^20 time-outs of 10 us each).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t timeouts = 0;
  while (!mbar_try_wait(bar, parity, 10000u)) {
    if (++timeouts > (1u << 20)) {
      printf("l1_top2: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}
// Producer-side wait: the slot it waits for is microseconds away and the ring is 16 tiles deep
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity) {
  uint32_t timeouts = 0;
  while (!mbar_try_wait(bar, parity, 100000u)) {
    if (++timeouts > (1u << 18)) {
      printf("l1_top2: producer mbarrier wait timed out (block %d)\n", blockIdx.x);
      __trap();
    }
  }
}
// whole tile in one op: the tensor map (make_map) views the row-major matrix as
// (16 B, rows, dim/16 chunks) with strides (1, dim, 16), so a box of
// (16, box_rows, dim/16) lands in shared memory as [chunk][row][16 B]
__device__ __forceinline__ void tma_load_tile(uint32_t dst, const CUtensorMap* map, int row0, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(map), "r"(0), "r"(row0), "r"(0), "r"(bar)
      : "memory");
}
// packed 4 x u8 sum of absolute differences, accumulated: SASS VABSDIFF4.U8.ACC
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
// s += (signed top byte of v): negative iff some folded accumulator was negative
// (non-negative accumulators are < 2^24 for dim <= 65536).  SASS IDP.4A, fma pipe.
__device__ __forceinline__ int fold_sign(uint32_t v, int s) {
  int r;
  asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(0x01000000u), "r"(s));
  return r;
}

struct Params {
  int xrows, yrows, nk16;        // nk16 = dim / 16
  uint32_t xoffset;              // global index of local train row 0
  int nxt, nyt;                  // tiles along train / query
  long long units;               // nxt * nyt
  unsigned long long* keys;      // rank 0 at keys[row], rank 1 at keys[kstride + row]
  long long kstride;             // rows of the whole key buffer (>= yrows: a launch may cover a slice of the query rows)
  int nst;                       // train-tile stages in the shared-memory ring
};

// shared memory carve-up (dynamic):  [ytile][x stages][state][barriers]
__host__ __device__ inline size_t smem_ytile_bytes(int nk16) { return (size_t)nk16 * BY * 16; }
__host__ __device__ inline size_t smem_xstage_bytes(int nk16) { return (size_t)nk16 * BX * 16; }
constexpr size_t SMEM_STATE_BYTES = (size_t)2 * BY * 8;   // best[BY], second[BY] keys of the CTA's query tile

// global fold: keys[0][row] / keys[1][row] <- top2(old pair U {k}).  The loser of
// the first atomicMin is offered to rank 1; the 2nd smallest key overall is
// always the loser of some exchange, and no loser is smaller than it.  A key that
// is already the best (the same train row folded twice: a shard or xoffset range
// offered again, or a pair seeded from this buffer and published unchanged) is
// not offered to rank 1, so the two ranks never hold the same train row.
__device__ __forceinline__ void fold_key(unsigned long long* keys, long long kstride, int row, unsigned long long k) {
  unsigned long long old = atomicMin(keys + row, k);
  if (old == k) return;
  unsigned long long loser = old > k ? old : k;
  if (loser != KEY_NONE) atomicMin(keys + kstride + row, loser);
}

// Threshold of a query row whose running 2nd-best key is `second`: a candidate takes the slow path iff
// dist < thr.  A stored key may carry a LARGER train index than a later candidate (keys seeded from the
// global buffer come from any train range), so a candidate that ties the 2nd-best distance must still be
// offered: thr = d2 + 1 (the whole-key atomicMin then decides).  No 2nd-best yet: INT_MAX admits every
// real distance (BruteForceNnL1K2.h:99-103).
__device__ __forceinline__ uint32_t thr_of(unsigned long long second) {
  const uint32_t d2 = (uint32_t)(second >> 32);
#ifdef SPV_K1_STRICT       // (A/B builds only) round 1's strict threshold: valid only without seeding
  return d2 < THR_MAX ? d2 : THR_MAX;
#endif
  return d2 < THR_MAX ? d2 + 1u : THR_MAX;
}

// slow path: offer one candidate to its query row's (best, second) keys in shared
// memory.  The 8 lanes that split a row's train columns share ONE pair per row, so
// the row threshold is the true running 2nd-best (a per-lane pair would make it
// about the 10th-best and the slow path 4x more frequent: ncu r01, 39 % of tiles).
__device__ __forceinline__ void offer_body(unsigned long long* best, uint32_t dist, uint32_t gidx) {
  const unsigned long long key = ((unsigned long long)dist << 32) | gidx;
  const unsigned long long old = atomicMin(best, key);
  if (old == key) return;   // the seeded best offered again (same train range folded twice)
  const unsigned long long loser = old > key ? old : key;
  if (loser != KEY_NONE) atomicMin(best + BY, loser);
}
__device__ __noinline__ void offer(unsigned long long* best, uint32_t dist, uint32_t gidx) { offer_body(best, dist, gidx); }

// Half a k-chunk (8 dims) of the 8 x 8 register tile.  Operands are loaded by kload() one half-chunk AHEAD
// of the ksad() that consumes them (two register sets of 16 x 8 B, the same 64 registers a whole chunk of
// 16 LDS.128 would take), so the 16 LDS.64 of the next half are in flight under the 128 VABSDIFF4 of this
// one instead of in front of them.  (9 warps per CTA cap the kernel at 168 registers -- three warps share
// one scheduler's 16 K registers -- so a second set of full 16-byte operands does not fit.)
__device__ __forceinline__ void kload(uint2 (&ra)[TY], uint2 (&rb)[TX], const uint2* __restrict__ ys,
                                      const uint2* __restrict__ xs) {
#pragma unroll
  for (int i = 0; i < TY; ++i) ra[i] = ys[i * 8];    // rows ly + 4 i   (uint2 units: 2 per 16-byte row slot)
#pragma unroll
  for (int j = 0; j < TX; ++j) rb[j] = xs[j * 16];   // rows lx + 8 j
}

template <bool FIRST>
__device__ __forceinline__ void ksad(uint32_t (&acc)[TY][TX], const uint32_t (&negthr)[TY], const uint2 (&ra)[TY],
                                     const uint2 (&rb)[TX]) {
#pragma unroll
  for (int i = 0; i < TY; ++i)
#pragma unroll
    for (int j = 0; j < TX; ++j) acc[i][j] = sad4(ra[i].x, rb[j].x, FIRST ? negthr[i] : acc[i][j]);
#pragma unroll
  for (int i = 0; i < TY; ++i)
#pragma unroll
    for (int j = 0; j < TX; ++j) acc[i][j] = sad4(ra[i].y, rb[j].y, acc[i][j]);
}

// the last half-chunk of a tile: each accumulator's sign is folded into its row's sum right behind its final
// VABSDIFF4, so the 64 IDP.4A ride in the SAD stream instead of forming a dependent tail after it
__device__ __forceinline__ void ksad_last(uint32_t (&acc)[TY][TX], int (&rs)[TY], const uint2 (&ra)[TY], const uint2 (&rb)[TX]) {
#pragma unroll
  for (int i = 0; i < TY; ++i)
#pragma unroll
    for (int j = 0; j < TX; ++j) acc[i][j] = sad4(ra[i].x, rb[j].x, acc[i][j]);
#pragma unroll
  for (int i = 0; i < TY; ++i)
#pragma unroll
    for (int j = 0; j < TX; ++j) {
      acc[i][j] = sad4(ra[i].y, rb[j].y, acc[i][j]);
      rs[i] = fold_sign(acc[i][j], rs[i]);
    }
}

// Seeds of one query tile: a warp's 32 rows' (best, second) keys from the global buffer into its shared state.
// Deliberately NOT inlined: with the loads inline the key pointer and stride stay live across the SAD loop of a
// kernel that sits at its 168-register cap, and the hot loop's schedule degrades (measured on the same box:
// 523.7 vs 528.6 G pairs/s); behind a call they are fetched from the constant bank when needed.
__device__ __noinline__ void seed_rows(const Params* __restrict__ p, unsigned long long* st, int srow) {
  unsigned long long s1 = KEY_NONE, s2 = KEY_NONE;
  if (srow < p->yrows) { s1 = __ldcg(p->keys + srow); s2 = __ldcg(p->keys + p->kstride + srow); }
  st[0] = s1;
  st[BY] = s2;
}

__global__ void __launch_bounds__(NTHREADS, 1)
l1_top2_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmy,
               const __grid_constant__ Params p) {
  extern __shared__ __align__(128) unsigned char smem[];
#ifndef SPV_K1_OLD_SLOW
  __shared__ uint4 slow_scr[NCW * 32 * 2];   // slow path: one 32-byte line per lane (8 accumulators of one query row)
#endif
  const int nk16 = p.nk16, NST = p.nst;
  unsigned char* ytile = smem;
  unsigned char* xst = ytile + smem_ytile_bytes(nk16);
  unsigned long long* state = reinterpret_cast<unsigned long long*>(xst + NST * smem_xstage_bytes(nk16));
  unsigned long long* bars = state + 2 * BY;
  // bars: full[NST], empty[NST], yfull, yempty
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * NST;
  const uint32_t bar_yfull = bar_empty + 8 * NST, bar_yempty = bar_yfull + 8;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, NCW); }
    mbar_init(bar_yfull, 1);
    mbar_init(bar_yempty, NCW);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // contiguous unit range of this CTA
  const long long u0 = p.units * blockIdx.x / gridDim.x;
  const long long u1 = p.units * (blockIdx.x + 1) / gridDim.x;
  if (u0 >= u1) return;
  int yt = (int)(u0 / p.nxt), xt = (int)(u0 % p.nxt);

  if (warp == NCW) {
    // ===================== producer: one thread issues all TMA loads
    if (lane == 0) {
      const uint32_t ytile_s = smem_u32(ytile), xst_s = smem_u32(xst);
      const uint32_t xbytes = (uint32_t)smem_xstage_bytes(nk16), ybytes = (uint32_t)smem_ytile_bytes(nk16);
      int stage = 0;
      uint32_t phase = 0, yphase = 0;
      bool newy = true, firsty = true;
      for (long long u = u0; u < u1; ++u) {
        if (newy) {
          if (!firsty) { mbar_wait_sleep(bar_yempty, yphase); yphase ^= 1; }
          firsty = false;
          mbar_expect_tx(bar_yfull, ybytes);
          tma_load_tile(ytile_s, &tmy, yt * BY, bar_yfull);
          newy = false;
        }
        mbar_wait_sleep(bar_empty + 8 * stage, phase ^ 1);
        mbar_expect_tx(bar_full + 8 * stage, xbytes);
        tma_load_tile(xst_s + stage * xbytes, &tmx, xt * BX, bar_full + 8 * stage);
        if (++stage == NST) { stage = 0; phase ^= 1; }
        if (++xt == p.nxt) { xt = 0; ++yt; newy = true; }
      }
    }
    return;
  }

  // ===================== compute warps
  const int ly = lane >> 3, lx = lane & 7;
  unsigned long long* st = state + warp * 32;                     // this warp's 32 rows: best at [r], second at [BY + r]
  const uint4* ys = reinterpret_cast<const uint4*>(ytile) + warp * 32 + ly;   // + c*BY + 4 i
  const uint4* xs0 = reinterpret_cast<const uint4*>(xst) + lx;                // + stage*nk16*BX + c*BX + 8 j
  const int xstage_u4 = nk16 * BX;

  uint32_t acc[TY][TX];
  uint32_t negthr[TY];
  int stage = 0;
  uint32_t phase = 0, yphase = 0;
  int cur_yt = -1;

  auto flush = [&](int ytile_idx) {   // lane r publishes row warp*32 + r
    __syncwarp();
    const int row = ytile_idx * BY + warp * 32 + lane;
    if (row < p.yrows) {
      const unsigned long long k1 = st[lane], k2 = st[BY + lane];
      if (k1 != KEY_NONE) fold_key(p.keys, p.kstride, row, k1);
      if (k2 != KEY_NONE) fold_key(p.keys, p.kstride, row, k2);
    }
    __syncwarp();
  };

  for (long long u = u0; u < u1; ++u) {
    if (yt != cur_yt) {
      if (cur_yt >= 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_yempty);   // this warp no longer reads the old query tile
        flush(cur_yt);
      }
      // Seed the rows' (best, second) from what the key buffer already holds -- earlier launches over other
      // train ranges (the host path streams the train set in chunks, host_ctx.cu) or other CTAs of this one --
      // so that a launch does not pay the threshold warm-up (~2 ln N slow-path hits per query row) again.
      // A stale read only means a weaker threshold; the pair is published again through fold_key, which
      // ignores a key that is already there.
#ifndef SPV_K1_NO_SEED   // (A/B builds only: tools/gpu_r2_c.sh)
      seed_rows(&p, st + lane, yt * BY + warp * 32 + lane);
#else
      st[lane] = KEY_NONE;
      st[BY + lane] = KEY_NONE;
#endif
      __syncwarp();
#pragma unroll
      for (int i = 0; i < TY; ++i) negthr[i] = 0u - thr_of(st[BY + ly + 4 * i]);
      mbar_wait(bar_yfull, yphase);
      yphase ^= 1;
      cur_yt = yt;
    }
    mbar_wait(bar_full + 8 * stage, phase);
    const uint4* xs = xs0 + stage * xstage_u4;
    int rs[TY];   // per query row of the thread tile: < 0 iff some accumulator of the row ended negative
#pragma unroll
    for (int i = 0; i < TY; ++i) rs[i] = 0;
    {
      // half-chunk h = 2 c + (0 | 1): the 8-byte halves of chunk c sit at byte offsets 0 and 8 of the 16-byte slots
      const uint2* y2 = reinterpret_cast<const uint2*>(ys);
      const uint2* x2 = reinterpret_cast<const uint2*>(xs);
      uint2 raA[TY], rbA[TX], raB[TY], rbB[TX];
      kload(raA, rbA, y2, x2);
      kload(raB, rbB, y2 + 1, x2 + 1);
      ksad<true>(acc, negthr, raA, rbA);
#pragma unroll 1   // measured: a compile-time trip count or any unrolling here loses 1-20 % (spills / I-cache)
      for (int c = 1; c < nk16; ++c) {     // entering: B holds the second half of chunk c - 1
        kload(raA, rbA, y2 + c * (2 * BY), x2 + c * (2 * BX));
        ksad<false>(acc, negthr, raB, rbB);
        kload(raB, rbB, y2 + c * (2 * BY) + 1, x2 + c * (2 * BX) + 1);
        ksad<false>(acc, negthr, raA, rbA);
      }
      ksad_last(acc, rs, raB, rbB);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * stage);   // operands are in registers: release the stage

    // candidate test: any accumulator negative  <=>  some dist < 2nd best of its row.  One sign sum per query
    // row of the thread tile (8 independent IDP.4A chains), so that the slow path only visits the rows that
    // tripped: a typical hit is one pair out of the warp's 2 048, and scanning all 64 accumulators of every
    // lane for it cost 3.4 % of the kernel (measured by disabling the slow path)
    const int anyneg = (rs[0] | rs[1] | rs[2] | rs[3]) | (rs[4] | rs[5] | rs[6] | rs[7]);
#ifndef SPV_K1_OLD_SLOW
    if (__any_sync(FULL, anyneg < 0)) {
      // Slow path (rare): exact (dist, idx) insertion, then tighten the row thresholds.  Written for FEW INSTRUCTION
      // CACHE LINES, not for few instructions: the code is cold every time it runs (the SAD loop owns the instruction
      // cache), and the straightforward form -- 64 unrolled "accumulator negative? offer it" sites, 15 KB of code of
      // which a visit touched ~27 lines -- cost 700-1500 cycles per visit (tools/castle_k1.py: 0.24 ms of a
      // 27 k x 27 k launch, 1.9 % at C3).  Here the tripped rows are walked by a loop: the 8 lanes that share a query
      // row agree on the row (mask OR-ed over the group), the lane's 8 accumulators of that row go to a per-lane
      // scratch line by predicated stores (the only place the register tile needs a static index), a compact loop
      // offers the negative ones, and only that row's threshold is re-read.
      uint32_t m = 0;
#pragma unroll
      for (int i = 0; i < TY; ++i) m |= ((uint32_t)rs[i] >> 31) << i;
      m |= __shfl_xor_sync(FULL, m, 1);
      m |= __shfl_xor_sync(FULL, m, 2);
      m |= __shfl_xor_sync(FULL, m, 4);
      const uint32_t g0 = (uint32_t)(xt * BX + lx);
      uint4* scr = slow_scr + (warp * 32 + lane) * 2;
#pragma unroll 1
      while (__any_sync(FULL, m != 0)) {
        const int i = m ? __ffs(m) - 1 : TY;      // TY: this group has no row left, it idles through the round
        m &= m - 1;
        uint32_t nt = 0;
#define SPV_ROW_OUT(k) case k: scr[0] = make_uint4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]); \
                               scr[1] = make_uint4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]); nt = negthr[k]; break;
        switch (i) { SPV_ROW_OUT(0) SPV_ROW_OUT(1) SPV_ROW_OUT(2) SPV_ROW_OUT(3) SPV_ROW_OUT(4) SPV_ROW_OUT(5) SPV_ROW_OUT(6) SPV_ROW_OUT(7) default: break; }
#undef SPV_ROW_OUT
        if (i < TY) {
          const uint32_t thr = 0u - nt;
          // the negative ones of the 8: a bit mask from the line just written, then one trip per candidate (the slow
          // path's instructions compete with the SAD stream for the ALU pipe: ncu, a cold 27 k x 27 k launch issues
          // 7.6 % more warp instructions than the same launch over seeded keys and runs 14 % longer)
          const uint32_t* a = reinterpret_cast<const uint32_t*>(scr);
          const uint4 a0 = scr[0], a1 = scr[1];
          uint32_t cm = (a0.x >> 31) | ((a0.y >> 31) << 1) | ((a0.z >> 31) << 2) | ((a0.w >> 31) << 3) |
                        ((a1.x >> 31) << 4) | ((a1.y >> 31) << 5) | ((a1.z >> 31) << 6) | ((a1.w >> 31) << 7);
#pragma unroll 1
          while (cm) {
            const int j = __ffs(cm) - 1;
            cm &= cm - 1;
            const uint32_t lidx = g0 + 8 * j;
            if (lidx < (uint32_t)p.xrows) offer_body(st + ly + 4 * i, a[j] + thr, p.xoffset + lidx);   // the one call site: inline
          }
        }
        __syncwarp();
        const uint32_t nn = i < TY ? 0u - thr_of(st[BY + ly + 4 * i]) : 0u;
#define SPV_THR_IN(k) case k: negthr[k] = nn; break;
        switch (i) { SPV_THR_IN(0) SPV_THR_IN(1) SPV_THR_IN(2) SPV_THR_IN(3) SPV_THR_IN(4) SPV_THR_IN(5) SPV_THR_IN(6) SPV_THR_IN(7) default: break; }
#undef SPV_THR_IN
      }
    }
#else
    if (__any_sync(FULL, anyneg < 0)) {
      // slow path (rare): exact (dist, idx) insertion, then tighten the row thresholds
      const uint32_t g0 = (uint32_t)(xt * BX + lx);
#pragma unroll
      for (int i = 0; i < TY; ++i) {
        if (rs[i] < 0) {
          const uint32_t thr = 0u - negthr[i];
#pragma unroll
          for (int j = 0; j < TX; ++j) {
            const uint32_t lidx = g0 + 8 * j;
            if ((int)acc[i][j] < 0 && lidx < (uint32_t)p.xrows) offer(st + ly + 4 * i, acc[i][j] + thr, p.xoffset + lidx);
          }
        }
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < TY; ++i) negthr[i] = 0u - thr_of(st[BY + ly + 4 * i]);
    }
#endif
    if (++stage == NST) { stage = 0; phase ^= 1; }
    if (++xt == p.nxt) { xt = 0; ++yt; }
  }
  flush(cur_yt);
}


// ---------------------------------------------------------------- wide rows
// dim > 512: a 256-row query tile no longer fits shared memory beside the train ring (the tiled
// kernel keeps whole rows resident).  The reference accepts any dim % 16 == 0
// (BruteForceNnL1K2.h:78-81), so wide rows take this plain tiled kernel instead: 64 x 64 row tiles,
// the columns streamed in 64-byte panels through shared memory (k-chunk-major, as K1 reads them),
// a 4 x 4 register tile per thread, the same key protocol (seeded thresholds, shared (best, second)
// per query row, fold_key).  Same results bit for bit; ALU-bound like K1 but at a lower rate (the
// operand loads are not hidden behind a producer warp) -- wide rows are not a graded configuration.
constexpr int WB = 64;            // rows per tile, both sides
constexpr int WCH = 4;            // 16-byte chunks per panel
constexpr int WTHREADS = 256;

struct WideParams {
  const uint8_t* x;
  const uint8_t* y;
  int xrows, yrows, dim;
  uint32_t xoffset;
  int nxt, splits;
  unsigned long long* keys;
  long long kstride;
};

__global__ void __launch_bounds__(WTHREADS) l1_top2_wide_kernel(const WideParams p) {
  __shared__ uint4 ys[WCH][WB], xs[WCH][WB];
  __shared__ unsigned long long st[2 * BY];   // offer() addresses the second at best + BY
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int y0 = blockIdx.x * WB;
  const int t0 = (int)((long long)p.nxt * blockIdx.y / p.splits), t1 = (int)((long long)p.nxt * (blockIdx.y + 1) / p.splits);
  if (tid < WB) {
    unsigned long long s1 = KEY_NONE, s2 = KEY_NONE;
    if (y0 + tid < p.yrows) { s1 = __ldcg(p.keys + y0 + tid); s2 = __ldcg(p.keys + p.kstride + y0 + tid); }
    st[tid] = s1;
    st[BY + tid] = s2;
  }
  __syncthreads();
  const int nk16 = p.dim / 16;
  for (int t = t0; t < t1; ++t) {
    const int x0 = t * WB;
    uint32_t acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0;
    for (int c0 = 0; c0 < nk16; c0 += WCH) {
      {   // stage: thread -> (chunk, row) of both tiles; rows past the end read as zeros (never offered)
        const int c = tid >> 6, r = tid & 63;
        uint4 vy = make_uint4(0, 0, 0, 0), vx = vy;
        if (c0 + c < nk16) {
          if (y0 + r < p.yrows) vy = __ldg(reinterpret_cast<const uint4*>(p.y + (size_t)(y0 + r) * p.dim) + c0 + c);
          if (x0 + r < p.xrows) vx = __ldg(reinterpret_cast<const uint4*>(p.x + (size_t)(x0 + r) * p.dim) + c0 + c);
        }
        ys[c][r] = vy;
        xs[c][r] = vx;
      }
      __syncthreads();
#pragma unroll
      for (int c = 0; c < WCH; ++c) {
        uint4 a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = ys[c][ty + 16 * i];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = xs[c][tx + 16 * j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            acc[i][j] = sad4(a[i].x, b[j].x, acc[i][j]); acc[i][j] = sad4(a[i].y, b[j].y, acc[i][j]);
            acc[i][j] = sad4(a[i].z, b[j].z, acc[i][j]); acc[i][j] = sad4(a[i].w, b[j].w, acc[i][j]);
          }
      }
      __syncthreads();
    }
    // selection: thresholds are read once per tile (stale = weaker, never wrong); offers are exact
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = ty + 16 * i;
      if (y0 + r >= p.yrows) continue;
      const uint32_t thr = thr_of(st[BY + r]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int lx = x0 + tx + 16 * j;
        if (lx < p.xrows && acc[i][j] < thr) offer(st + r, acc[i][j], p.xoffset + (uint32_t)lx);
      }
    }
    __syncthreads();
  }
  if (tid < WB && y0 + tid < p.yrows) {
    const unsigned long long k1 = st[tid], k2 = st[BY + tid];
    if (k1 != KEY_NONE) fold_key(p.keys, p.kstride, y0 + tid, k1);
    if (k2 != KEY_NONE) fold_key(p.keys, p.kstride, y0 + tid, k2);
  }
}

// ------------------------------------------------------- small helper kernels
__global__ void keys_merge_kernel(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src,
                                  int nsrc, int yrows) {
  // K5: lexicographic top-2 of dst's pair and nsrc more pairs per query row
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= yrows) return;
  unsigned long long a = dst[row], b = dst[yrows + row];
  for (int s = 0; s < nsrc; ++s) {
    const unsigned long long* q = src + (size_t)s * 2 * yrows;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const unsigned long long k = q[(size_t)r * yrows + row];
      if (k < a) { b = a; a = k; }
      else if (k < b && k != a) b = k;   // k == a: the same train row offered twice
    }
  }
  dst[row] = a;
  dst[yrows + row] = b;
}

template <typename D>
__global__ void keys_finalize_kernel(const unsigned long long* __restrict__ keys, int yrows, int kcols,
                                     uint64_t* __restrict__ idx, D* __restrict__ dist, long long kstride) {
  // BruteForceNnL1K2.h:99-103 sentinels: idx (size_t)-1, dist INT_MAX; the cascade
  // reports (float)INT_MAX (CascadingHashNn.h:244).  Columns >= 2 (k > 2 on the
  // cascade ABI, Spectavi.cpp:329-334) are left as the caller's zeros.
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= yrows) return;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    if (r >= kcols) break;
    const unsigned long long k = keys[(size_t)r * kstride + row];
    const bool none = (k == KEY_NONE);
    idx[(size_t)row * kcols + r] = none ? ~0ull : (k & 0xFFFFFFFFull);
    dist[(size_t)row * kcols + r] = none ? (D)2147483647 : (D)(int32_t)(k >> 32);
  }
}

// -------------------------------------------------------------- host helpers
std::atomic<unsigned long long> g_launches{0};
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// rows x dim uint8, row-major, seen as 3-D (16 B, rows, dim/16) with byte strides
// (1, dim, 16); box = (16, box_rows, dim/16) = one whole tile, delivered to shared
// memory k-chunk-major.  Measured (tools/ubench_tma_box.cu): one such op fills an
// 8 KB stage in 544 cycles with 148 SMs streaming, vs 850 as 8 separate 16 B boxes.
}  // namespace

bool make_rows_map(CUtensorMap* map, const uint8_t* base, int rows, int dim, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return false;
  cuuint64_t gdim[3] = {16u, (cuuint64_t)rows, (cuuint64_t)(dim / 16)};
  cuuint64_t gstride[2] = {(cuuint64_t)dim, 16u};
  cuuint32_t box[3] = {16u, (cuuint32_t)box_rows, (cuuint32_t)(dim / 16)};
  cuuint32_t estr[3] = {1u, 1u, 1u};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// rows x dim uint8, row-major, as a plain 2-D map with 128-byte-wide boxes and the 128-byte swizzle: every request is
// a whole 128-byte line (the permuted 16-byte-wide map above makes TMA fetch every line as eight 16-byte pieces: K2's
// L2 traffic was 3.4x its DRAM traffic), and the tile lands in the canonical UMMA K-major SWIZZLE_128B layout
bool make_rows_map_sw128(CUtensorMap* map, const uint8_t* base, int rows, int dim, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)dim};
  cuuint32_t box[2] = {128u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// column-major float32 matrix [dim][pitch] (pitch floats per column, rows valid), seen as 2-D (rows, dim); a box of
// (box_rows, box_cols) lands in shared memory as [col][row]; rows / columns outside the matrix read as zeros
bool make_f32_panel_map(CUtensorMap* map, const float* base, int rows, int dim, size_t pitch, int box_rows, int box_cols) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)(rows > 0 ? rows : 1), (cuuint64_t)dim};
  cuuint64_t gstride[1] = {(cuuint64_t)pitch * 4u};
  cuuint32_t box[2] = {(cuuint32_t)box_rows, (cuuint32_t)box_cols};
  cuuint32_t estr[2] = {1u, 1u};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

namespace {

cudaError_t launch_tiled(const CUtensorMap& tmx, const CUtensorMap& tmy, Params p, int sms, cudaStream_t stream) {
  // as many ring stages as fit beside the query tile: the ring is also the slack
  // between the 8 compute warps (a warp may run at most nst tiles ahead of the
  // slowest one), and slow-path hits make their pace uneven
  const size_t fixed = smem_ytile_bytes(p.nk16) + SMEM_STATE_BYTES + 1024, budget = 200 * 1024;
  int nst = (int)((budget - fixed) / smem_xstage_bytes(p.nk16));
  nst = nst > 16 ? 16 : nst;
  if (nst < 2) return cudaErrorInvalidValue;
  p.nst = nst;
  const size_t smem = smem_ytile_bytes(p.nk16) + nst * smem_xstage_bytes(p.nk16) + SMEM_STATE_BYTES + (2 * nst + 2) * 8;
  const int grid = (int)(p.units < sms ? p.units : sms);
  cudaError_t e = cudaFuncSetAttribute(l1_top2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // per device
  if (e != cudaSuccess) return e;
  l1_top2_kernel<<<grid, NTHREADS, smem, stream>>>(tmx, tmy, p);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

unsigned long long launches() { return g_launches.load(); }
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n); }

cudaError_t keys_reset(unsigned long long* keys, int yrows, cudaStream_t stream) {
  if (yrows <= 0) return cudaSuccess;
  return cudaMemsetAsync(keys, 0xFF, (size_t)2 * yrows * sizeof(unsigned long long), stream);
}

cudaError_t keys_merge(unsigned long long* dst, const unsigned long long* src, int nsrc, int yrows, cudaStream_t stream) {
  if (yrows <= 0 || nsrc <= 0) return cudaSuccess;
  keys_merge_kernel<<<(yrows + 255) / 256, 256, 0, stream>>>(dst, src, nsrc, yrows);
  count_launch();
  return cudaGetLastError();
}

cudaError_t keys_finalize_i32(const unsigned long long* keys, int yrows, int kcols, uint64_t* idx, int32_t* dist,
                              cudaStream_t stream, long long kstride) {
  if (yrows <= 0) return cudaSuccess;
  keys_finalize_kernel<int32_t><<<(yrows + 255) / 256, 256, 0, stream>>>(keys, yrows, kcols, idx, dist, kstride > 0 ? kstride : yrows);
  count_launch();
  return cudaGetLastError();
}

cudaError_t keys_finalize_f32(const unsigned long long* keys, int yrows, int kcols, uint64_t* idx, float* dist,
                              cudaStream_t stream) {
  if (yrows <= 0) return cudaSuccess;
  keys_finalize_kernel<float><<<(yrows + 255) / 256, 256, 0, stream>>>(keys, yrows, kcols, idx, dist, yrows);
  count_launch();
  return cudaGetLastError();
}

cudaError_t l1_top2_u8(const uint8_t* dx, const uint8_t* dy, int xrows, int yrows, int dim, uint32_t xoffset,
                       unsigned long long* keys, cudaStream_t stream, long long kstride) {
  if (xrows <= 0 || yrows <= 0) return cudaSuccess;   // nothing to fold in
  if (dim <= 0 || dim % 16 != 0) return cudaErrorInvalidValue;
  if (kstride < yrows) kstride = yrows;
  // keys carry 32 index bits: the last global index must stay below the "none" pattern
  if ((unsigned long long)xoffset + (unsigned long long)xrows > 0xFFFFFFFFull) return cudaErrorInvalidValue;
  int dev = 0, sms = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return e;
  if (dim > L1K2_TILED_MAX_DIM) {
    WideParams w;
    w.x = dx; w.y = dy; w.xrows = xrows; w.yrows = yrows; w.dim = dim; w.xoffset = xoffset;
    w.nxt = (xrows + WB - 1) / WB;
    const int nyt = (yrows + WB - 1) / WB;
    int splits = (4 * sms + nyt - 1) / nyt;          // about 4 CTAs per SM when the query side alone is short
    splits = splits < 1 ? 1 : (splits > w.nxt ? w.nxt : splits);
    w.splits = splits; w.keys = keys; w.kstride = kstride;
    l1_top2_wide_kernel<<<dim3(nyt, splits), WTHREADS, 0, stream>>>(w);
    count_launch();
    return cudaGetLastError();
  }
  CUtensorMap tmx, tmy;
  if (!make_rows_map(&tmx, dx, xrows, dim, BX) || !make_rows_map(&tmy, dy, yrows, dim, BY)) return cudaErrorInvalidValue;
  Params p;
  p.xrows = xrows; p.yrows = yrows; p.nk16 = dim / 16; p.xoffset = xoffset;
  p.nxt = (xrows + BX - 1) / BX; p.nyt = (yrows + BY - 1) / BY;
  p.units = (long long)p.nxt * p.nyt;
  p.keys = keys;
  p.kstride = kstride;
  return launch_tiled(tmx, tmy, p, sms, stream);
}

}  // namespace spv
