// cascade.cu -- the cascade-of-hashes matcher on sm_100a: replaces CascadingHashNn
// (reference src/CascadingHashNn.h) behind nn_cascading_hash (Spectavi.cpp:321-336).
//
//   K0 f32_to_u8          x.cast<uint8>() += 128            (CascadingHashNn.h:236-239)
//   K2 hash_project_kernel R = D * H on the tensor cores (tcgen05.mma, accumulators
//                         in TMEM), sign-packed to one int32 code per (row, table)
//                         (:102-135); for query rows also the g least-confident
//                         bit positions of each table (:150-169)
//   K3 bucket_*           per table: histogram of codes, exclusive scan, scatter of
//                         the train indices -> CSR buckets            (:187-196)
//   K4 probe_rerank_kernel per query: the 2^g probe codes of every table (:170-179),
//                         the union of their buckets (:208-227), exact L1 top-2 of
//                         those candidates on the uint8 rows           (:229-245)
//
// Exactness.  Rows are integers in [-128, 127] (exact in bf16).  The dictionary the
// drop-in entry point draws is quantised to multiples of 1/32 in [-127/32, 127/32]
// (exact in bf16), so every product is an integer multiple of 1/32 below 2^14 and
// every partial sum is exact in fp32 in ANY order: the tensor-core projection, a
// CPU float loop and Eigen's GEMM give identical signs, and codes are bit-exact
// against the oracle.  An arbitrary fp32 dictionary (device API, golden vectors
// produced by the reference's own generator) is split into three bf16 terms whose
// sum is the fp32 value exactly; the projection then differs from an fp32 CPU loop
// only by fp32 summation rounding, i.e. possibly in the sign of |R| < ~1e-3.
#include "spv_common.h"

#include <cuda_bf16.h>
#include <climits>
#include <cstdint>
#include <cstdlib>
#include <cstring>

namespace spv {
namespace {

// ------------------------------------------------------------------ PTX glue
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 24)) {   // a protocol bug must trap, never hang the device
      printf("cascade: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

// UMMA shared-memory matrix descriptor, K-major, no swizzle (canonical layout in
// 16-byte units ((8, n), 2) : ((1, SBO), LBO)): 8 rows x 16 B core matrices, SBO =
// byte stride between 8-row groups, LBO = byte stride between the two 8-element K
// halves of one K=16 step.  Bits: [0,14) addr>>4, [16,30) LBO>>4, [32,46) SBO>>4,
// [46,48) version = 1 (Blackwell), [61,64) layout type = 0 (no swizzle).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16: D fp32 (bits[4,6) = 1), A, B bf16 (bits[7,10) =
// bits[10,13) = 1), both K-major (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
__device__ __forceinline__ uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// ---------------------------------------------------------------- K0
// CascadingHashNn.h:236-239: cast<uint8_t>() then += 128 in uint8 arithmetic, i.e.
// (uint8)((int)v + 128) for the integer-valued rows the Python wrapper passes.
__global__ void f32_to_u8_kernel(const float4* __restrict__ in, uint32_t* __restrict__ out, size_t n4) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const float4 v = in[i];
  const uint32_t b0 = (uint32_t)((int)v.x + 128) & 0xFF, b1 = (uint32_t)((int)v.y + 128) & 0xFF;
  const uint32_t b2 = (uint32_t)((int)v.z + 128) & 0xFF, b3 = (uint32_t)((int)v.w + 128) & 0xFF;
  out[i] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
}

// ---------------------------------------------------------------- dictionary image
// dict [n][dim][m] fp32 (hash_dict(i, j), CascadingHashNn.h:93-97) -> the B operand
// as it sits in shared memory: for every group of tables, three bf16 terms
// (hi, mid, lo; hi + mid + lo == the fp32 value), each [dim/8 k-chunks][N][8 elems]
// with column N-index = table_in_group * MP + bit, zero for padding columns.
__global__ void dict_image_kernel(const float* __restrict__ dict, int n, int dim, int m, int MP, int tables_per_group,
                                  int N, __nv_bfloat16* __restrict__ img) {
  const int groups = (n + tables_per_group - 1) / tables_per_group;
  const size_t per_term = (size_t)(dim / 8) * N * 8;
  const size_t total = (size_t)groups * 3 * per_term;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const int g = (int)(e / (3 * per_term));
    const int term = (int)((e / per_term) % 3);
    const size_t r = e % per_term;
    const int kc = (int)(r / ((size_t)N * 8)), col = (int)((r / 8) % N), ke = (int)(r % 8);
    const int t = g * tables_per_group + col / MP, bit = col % MP, k = kc * 8 + ke;
    float v = 0.f;
    if (col / MP < tables_per_group && t < n && bit < m) v = dict[((size_t)t * dim + k) * m + bit];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const float r1 = v - __bfloat162float(hi);
    const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
    const __nv_bfloat16 lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
    img[e] = term == 0 ? hi : term == 1 ? mid : lo;
  }
}

// ---------------------------------------------------------------- K2
constexpr int HP_ROWS = 128;     // rows per tile = UMMA M
constexpr int HP_KPANEL = 128;   // columns of A staged per pass (16 k-chunks of 8)
constexpr int HP_THREADS = 128;  // 4 warps: warp w owns TMEM lanes 32 w .. 32 w + 31

struct HpParams {
  const uint8_t* rows_u8;        // [rows][dim], value + 128
  const __nv_bfloat16* img;      // dictionary image (dict_image_kernel)
  int rows, dim, n, m, MP, tpg, N, groups, terms;
  int g;                         // > 0: query mode, also emit the g least-confident bits
  int32_t* codes;                // [n][rows]
  unsigned long long* flips;     // [n][rows], 5 bits per position (query mode)
  int tmem_cols;
};

template <int MP>
__device__ __forceinline__ void hp_epilogue_table(uint32_t taddr, int m, int g, int32_t* code_out,
                                                  unsigned long long* flip_out) {
  float v[MP];
#pragma unroll
  for (int c = 0; c < MP; c += 8) {
    float t8[8];
    tmem_ld8(taddr + c, t8);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[c + i] = t8[i];
  }
  // convert_to_hashcode (CascadingHashNn.h:102-111): bit i set iff R(i) >= 0
  uint32_t code = 0;
#pragma unroll
  for (int b = 0; b < MP; ++b)
    if (b < m && v[b] >= 0.f) code |= 1u << b;
  *code_out = (int32_t)code;
  if (g > 0) {
    // generate_y_candidate_hashcodes (:153-159): the g smallest (|R|, bit) pairs
    uint32_t taken = 0;
    unsigned long long flips = 0;
    const int nb = g < m ? g : m;
    for (int q = 0; q < nb; ++q) {
      float best = __int_as_float(0x7f800000);
      int bi = 0;
      bool found = false;
#pragma unroll
      for (int b = 0; b < MP; ++b) {
        const float a = fabsf(v[b]);
        if (b < m && !((taken >> b) & 1u) && (!found || a < best)) { best = a; bi = b; found = true; }
      }
      taken |= 1u << bi;
      flips |= (unsigned long long)bi << (5 * q);
    }
    *flip_out = flips;
  }
}

__global__ void __launch_bounds__(HP_THREADS) hash_project_kernel(const HpParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int dim = P.dim, nkc = dim / 8, N = P.N;
  const int bterm_bytes = nkc * N * 16;                      // one bf16 term of one group
  unsigned char* sB = smem;                                  // [terms][nkc][N][16 B]
  unsigned char* sA = sB + (size_t)P.terms * bterm_bytes;    // [16 k-chunks][128 rows][16 B]
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(sA + (HP_KPANEL / 8) * HP_ROWS * 16);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(smem_u32(bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(P.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  const uint32_t idesc = umma_idesc_bf16(HP_ROWS, N);
  const int ntiles = (P.rows + HP_ROWS - 1) / HP_ROWS;
  const int npanels = (dim + HP_KPANEL - 1) / HP_KPANEL;
  uint32_t parity = 0;

  for (int grp = 0; grp < P.groups; ++grp) {
    // B operand of this group of tables: a straight copy of the prepared image
    {
      const uint4* src = reinterpret_cast<const uint4*>(P.img) + (size_t)grp * 3 * (bterm_bytes / 16);
      uint4* dst = reinterpret_cast<uint4*>(sB);
      for (int term = 0; term < P.terms; ++term)
        for (int e = tid; e < bterm_bytes / 16; e += HP_THREADS)
          dst[(size_t)term * (bterm_bytes / 16) + e] = src[(size_t)term * (bterm_bytes / 16) + e];
    }
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int row0 = tile * HP_ROWS;
      for (int pn = 0; pn < npanels; ++pn) {
        const int c0 = pn * HP_KPANEL;                        // first column of the panel
        const int pcols = min(HP_KPANEL, dim - c0);           // multiple of 16
        const int p16 = pcols / 16;                           // 16-byte input chunks per row
        // stage A: uint8 (value + 128) -> bf16 value, k-chunk-major
        for (int e = tid; e < HP_ROWS * p16; e += HP_THREADS) {
          const int r = e / p16, j = e % p16;
          uint4 raw = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);   // zero rows
          if (row0 + r < P.rows) raw = *reinterpret_cast<const uint4*>(P.rows_u8 + (size_t)(row0 + r) * dim + c0 + j * 16);
          const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
          uint32_t o[8];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float f0 = (float)((int)(w[q] & 0xFF) - 128), f1 = (float)((int)((w[q] >> 8) & 0xFF) - 128);
            const float f2 = (float)((int)((w[q] >> 16) & 0xFF) - 128), f3 = (float)((int)(w[q] >> 24) - 128);
            __nv_bfloat162 lo2 = __floats2bfloat162_rn(f0, f1), hi2 = __floats2bfloat162_rn(f2, f3);
            o[2 * q] = *reinterpret_cast<uint32_t*>(&lo2);
            o[2 * q + 1] = *reinterpret_cast<uint32_t*>(&hi2);
          }
          *reinterpret_cast<uint4*>(sA + ((size_t)(2 * j) * HP_ROWS + r) * 16) = make_uint4(o[0], o[1], o[2], o[3]);
          *reinterpret_cast<uint4*>(sA + ((size_t)(2 * j + 1) * HP_ROWS + r) * 16) = make_uint4(o[4], o[5], o[6], o[7]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
          for (int term = 0; term < P.terms; ++term)
            for (int s = 0; s < pcols / 16; ++s) {              // one UMMA per K = 16
              const uint64_t ad = umma_desc(a0 + (2 * s) * (HP_ROWS * 16), HP_ROWS * 16, 128);
              const uint64_t bd = umma_desc(b0 + term * bterm_bytes + (c0 / 8 + 2 * s) * (N * 16), N * 16, 128);
              umma_bf16(tmem, ad, bd, idesc, (pn | term | s) != 0);
            }
          umma_commit(smem_u32(bar));   // implies tcgen05.fence::before_thread_sync
        }
        mbar_wait(smem_u32(bar), parity);   // MMAs done: A may be overwritten, D may be read
        parity ^= 1;
      }
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      // epilogue: thread = row (TMEM lane), columns [t_in_group * MP, + MP)
      const int row = row0 + tid;
      for (int tg = 0; tg < P.tpg; ++tg) {
        const int t = grp * P.tpg + tg;
        if (t >= P.n) break;
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + tg * P.MP;
        int32_t code;
        unsigned long long fl = 0;
        if (P.MP == 8) hp_epilogue_table<8>(taddr, P.m, P.g, &code, &fl);
        else if (P.MP == 16) hp_epilogue_table<16>(taddr, P.m, P.g, &code, &fl);
        else if (P.MP == 24) hp_epilogue_table<24>(taddr, P.m, P.g, &code, &fl);
        else hp_epilogue_table<32>(taddr, P.m, P.g, &code, &fl);
        if (row < P.rows) {
          P.codes[(size_t)t * P.rows + row] = code;
          if (P.g > 0) P.flips[(size_t)t * P.rows + row] = fl;
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();   // everyone is done with sB before the next group's image lands
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(P.tmem_cols) : "memory");
}

// ---------------------------------------------------------------- K2, int8 path
// The dictionary the drop-in entry point draws is q / 32 with q an int8 (cascade_dictionary,
// quantize = true), and the rows are uint8 (value + 128), so the projection is an exact
// integer GEMM:  32 R(row, col) = sum_k u8(row, k) * q(k, col) - 128 * sum_k q(k, col).
// tcgen05.mma kind::i8 takes the uint8 rows AS THEY LIE IN HBM (A unsigned, B signed, int32
// accumulators in TMEM), so nothing is converted: one TMA op per 128-row tile drops the rows
// into shared memory already in the UMMA K-major core-matrix order (make_rows_map), one elected
// thread issues dim/32 MMAs, and the epilogue subtracts the per-column constant.  Warp roles:
// warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue (TMEM lane quarter = warp % 4, two warps per
// quarter taking alternate tables);
// a ring of nst row tiles and two TMEM accumulators keep load, MMA and epilogue overlapped.
// Bound: HBM (dim bytes in, 4n (+8n) bytes out per row).  SASS: UTMALDG, UTCIMMA, LDTM.
constexpr int HQ_EPI_WARPS = 8;                       // two per TMEM lane quarter, tables split between them
constexpr int HQ_THREADS = (2 + HQ_EPI_WARPS) * 32;

struct HqParams {
  const int8_t* img;             // [groups][dim/16][N][16] int8 hyperplane coefficients * 32
  const int32_t* colsum;         // [groups][N]  sum_k q(k, col)
  int rows, dim, n, m, MP, tpg, N, groups, g, nst, tmem_cols;
  int swz;                       // 1: row tiles come as 128-byte-wide SWIZZLE_128B panels (dim % 128 == 0)
  int32_t* codes;                // [n][rows]
  unsigned long long* flips;     // [n][rows]
};

__global__ void dict_image_i8_kernel(const float* __restrict__ dict, int n, int dim, int m, int MP, int tpg, int N,
                                     int groups, int8_t* __restrict__ img, int32_t* __restrict__ colsum) {
  const size_t per_group = (size_t)dim * N, total = (size_t)groups * per_group;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const int g = (int)(e / per_group);
    const size_t r = e % per_group;
    const int kc = (int)(r / ((size_t)N * 16)), col = (int)((r / 16) % N), ke = (int)(r % 16);
    const int t = g * tpg + col / MP, bit = col % MP, k = kc * 16 + ke;
    float v = 0.f;
    if (col / MP < tpg && t < n && bit < m) v = dict[((size_t)t * dim + k) * m + bit];
    img[e] = (int8_t)(int)rintf(v * 32.f);
  }
  // column sums: one thread per (group, column); dim <= 512 terms
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < groups * N; c += gridDim.x * blockDim.x) {
    const int g = c / N, col = c % N;
    const int t = g * tpg + col / MP, bit = col % MP;
    int s = 0;
    if (col / MP < tpg && t < n && bit < m)
      for (int k = 0; k < dim; ++k) s += (int)rintf(dict[((size_t)t * dim + k) * m + bit] * 32.f);
    colsum[c] = s;
  }
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_rows(uint32_t dst, const CUtensorMap* map, int row0, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(map), "r"(0), "r"(row0), "r"(0), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_load_panel_2d(uint32_t dst, const CUtensorMap* map, int col0, int row0, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(col0), "r"(row0), "r"(bar)
      : "memory");
}
// UMMA descriptor of a K-major SWIZZLE_128B operand: rows are 128 B apart, 8-row groups (the swizzle atom) 1024 B
// (SBO); LBO is unused when one instruction's K extent (32 B here) lies inside the 128-byte atom; layout type 2
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8: D int32 (c_format = 2 at bits [4,6)), A uint8 (a_format = 0 at [7,10)), B int8 (b_format = 1 at
// [10,13)), both K-major, N >> 3 at [17,23), M >> 4 at [24,29); K = 32 per instruction
__device__ __forceinline__ uint32_t umma_idesc_u8s8(int M, int N) {
  return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// 8 accumulator columns of this thread's TMEM lane, NO wait: the caller issues all loads of a table back to back
// and waits once (one TMEM round trip per table instead of one per 8 columns -- ncu r01 showed the epilogue warps
// parked on these loads, long_scoreboard 6.1 per issue)
__device__ __forceinline__ void tmem_ld8_i32_nowait(uint32_t taddr, int* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}

// one table of one row: MP accumulator columns -> code (+ least-confident bits); v = 32 R exactly
template <int MP>
__device__ __forceinline__ void hq_epilogue_table(uint32_t taddr, const int32_t* __restrict__ scol, int m, int nb,
                                                  int32_t* code_out, unsigned long long* flip_out) {
  int v[MP];
#pragma unroll
  for (int c = 0; c < MP; c += 8) tmem_ld8_i32_nowait(taddr + c, v + c);
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  // the loaded registers are only valid behind the wait: make every later use depend on a statement that is
  // ordered after it (volatile asm statements keep their order)
#pragma unroll
  for (int c = 0; c < MP; c += 8)
    asm volatile("" : "+r"(v[c]), "+r"(v[c + 1]), "+r"(v[c + 2]), "+r"(v[c + 3]), "+r"(v[c + 4]), "+r"(v[c + 5]), "+r"(v[c + 6]), "+r"(v[c + 7]));
#pragma unroll
  for (int c = 0; c < MP; ++c) v[c] = v[c] - 128 * scol[c];
  uint32_t code = 0;   // convert_to_hashcode (CascadingHashNn.h:102-111): bit i set iff R(i) >= 0
#pragma unroll
  for (int b = MP - 1; b >= 0; --b) code = code * 2u + (~(uint32_t)v[b] >> 31);
  code &= (m >= 32) ? 0xFFFFFFFFu : ((1u << m) - 1u);
  *code_out = (int32_t)code;
  if (nb > 0) {
    // generate_y_candidate_hashcodes (:153-159): the nb smallest (|R|, bit) pairs; |v| < 2^24, so
    // key = |v| * 32 + bit orders them and keys are distinct (the bit index is part of the key)
    uint32_t key[MP];
#pragma unroll
    for (int b = 0; b < MP; ++b) key[b] = b < m ? ((uint32_t)(v[b] < 0 ? -v[b] : v[b]) * 32u + (uint32_t)b) : 0xFFFFFFFFu;
    unsigned long long flips = 0;
    if (nb <= 2) {
      // the default probe depth (g = 2): both minima in ONE pass, 3 min / max per key
      // (VIMNMX issues at 5.6 T/s on this GPU against 18 T/s for other ALU ops -- pipe_peak.cu -- and this loop is what
      // keeps the query side ALU-bound; spelling it as setp + selp does not help: ptxas folds those back into VIMNMX)
      // (also measured: the minimum by a tree of 3-input minima, knock-out by compare + select, second tree -- 24 VIMNMX3 +
      // 48 full-rate ops instead of 72 VIMNMX for MP = 24 -- runs the query side at 97 us instead of 47: two live arrays of MP keys)
      uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
#pragma unroll
      for (int b = 0; b < MP; ++b) {
        m2 = min(m2, max(m1, key[b]));
        m1 = min(m1, key[b]);
      }
      flips = (unsigned long long)(m1 & 31u);
      if (nb == 2) flips |= (unsigned long long)(m2 & 31u) << 5;
    } else {
      uint32_t last = 0;   // take the keys in increasing order
      for (int q = 0; q < nb; ++q) {
        uint32_t best = 0xFFFFFFFFu;
#pragma unroll
        for (int b = 0; b < MP; ++b) {
          const uint32_t k = (q == 0 || key[b] > last) ? key[b] : 0xFFFFFFFFu;
          best = k < best ? k : best;
        }
        last = best;
        flips |= (unsigned long long)(best & 31u) << (5 * q);
      }
    }
    *flip_out = flips;
  }
}

__global__ void __launch_bounds__(HQ_THREADS) hash_project_i8_kernel(const __grid_constant__ CUtensorMap tmA, const HqParams P) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // the swizzle atom wants 1024-byte alignment
  const int dim = P.dim, N = P.N, nst = P.nst;
  const uint32_t stage_bytes = (uint32_t)HP_ROWS * dim;      // [dim/16][128 rows][16 B], or dim/128 swizzled panels of [128 rows][128 B]
  unsigned char* sA = smem;
  unsigned char* sB = sA + (size_t)nst * stage_bytes;         // [dim/16][N][16 B]
  int32_t* sCol = reinterpret_cast<int32_t*>(sB + (size_t)dim * N);
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(sCol + N + (N & 1));
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * nst, bar_tfull = bar_empty + 8 * nst,
                 bar_tempty = bar_tfull + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * nst + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < nst; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, HQ_EPI_WARPS * 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(P.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  const uint32_t idesc = umma_idesc_u8s8(HP_ROWS, N);
  const int ntiles = (P.rows + HP_ROWS - 1) / HP_ROWS;
  const int my_tiles = blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  // The producer warp never touches sB / sCol, so it takes no part in the per-group barriers: its first row tiles
  // are in flight while the other nine warps are still fetching the B operand, and it runs ahead across groups,
  // held back only by the ring's empty barriers.
  constexpr int HQ_WORKERS = HQ_THREADS - 32;
  for (int grp = 0; grp < P.groups; ++grp) {
    if (warp != 0) {   // B operand and column constants of this group of tables
      const uint4* src = reinterpret_cast<const uint4*>(P.img + (size_t)grp * dim * N);
      uint4* dst = reinterpret_cast<uint4*>(sB);
      for (int e = tid - 32; e < dim * N / 16; e += HQ_WORKERS) dst[e] = src[e];
      for (int e = tid - 32; e < N; e += HQ_WORKERS) sCol[e] = P.colsum[grp * N + e];
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"n"(HQ_WORKERS) : "memory");
    }
    const uint32_t it0 = (uint32_t)grp * (uint32_t)my_tiles;   // the same iteration count in every role
    if (warp == 0) {
      if (lane == 0)
        for (int j = 0; j < my_tiles; ++j) {
          const uint32_t it = it0 + j, s = it % nst, ph = (it / nst) & 1u;
          mbar_wait(bar_empty + 8 * s, ph ^ 1u);
          mbar_expect_tx(bar_full + 8 * s, stage_bytes);
          const int row0 = (blockIdx.x + j * gridDim.x) * HP_ROWS;
          if (P.swz) {
            for (int pnl = 0; pnl < dim / 128; ++pnl)
              tma_load_panel_2d(smem_u32(sA) + s * stage_bytes + pnl * (HP_ROWS * 128), &tmA, pnl * 128, row0, bar_full + 8 * s);
          } else {
            tma_load_rows(smem_u32(sA) + s * stage_bytes, &tmA, row0, bar_full + 8 * s);
          }
        }
    } else if (warp == 1) {
      if (lane == 0)
        for (int j = 0; j < my_tiles; ++j) {
          const uint32_t it = it0 + j, s = it % nst, ph = (it / nst) & 1u, a = it & 1u, aph = (it >> 1) & 1u;
          mbar_wait(bar_tempty + 8 * a, aph ^ 1u);     // the epilogue has drained this accumulator
          mbar_wait(bar_full + 8 * s, ph);             // the row tile has landed
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a0 = smem_u32(sA) + s * stage_bytes, b0 = smem_u32(sB);
          for (int k = 0; k < dim / 32; ++k) {         // one UMMA per K = 32 bytes = two 16-byte k-chunks
            const uint64_t ad = P.swz ? umma_desc_sw128(a0 + (k >> 2) * (HP_ROWS * 128) + (k & 3) * 32)
                                      : umma_desc(a0 + (2 * k) * (HP_ROWS * 16), HP_ROWS * 16, 128);
            const uint64_t bd = umma_desc(b0 + (2 * k) * (N * 16), N * 16, 128);
            umma_i8(tmem + a * N, ad, bd, idesc, k != 0);
          }
          umma_commit(bar_empty + 8 * s);              // the stage may be refilled once these MMAs have read it
          umma_commit(bar_tfull + 8 * a);              // ... and the accumulator may be read
        }
    } else {
      const int quarter = warp & 3;                    // TMEM lanes 32 * (warp % 4) ..
      const int half = (warp - 2) >> 2;                // which of the two warps of this quarter
      for (int j = 0; j < my_tiles; ++j) {
        const uint32_t it = it0 + j, a = it & 1u, aph = (it >> 1) & 1u;
        mbar_wait(bar_tfull + 8 * a, aph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = (blockIdx.x + j * gridDim.x) * HP_ROWS + quarter * 32 + lane;
        for (int tg = half; tg < P.tpg; tg += HQ_EPI_WARPS / 4) {
          const int t = grp * P.tpg + tg;
          if (t >= P.n) break;
          const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + a * N + tg * P.MP;
          int32_t code;
          unsigned long long fl = 0;
          if (P.MP == 8) hq_epilogue_table<8>(taddr, sCol + tg * 8, P.m, P.g, &code, &fl);
          else if (P.MP == 16) hq_epilogue_table<16>(taddr, sCol + tg * 16, P.m, P.g, &code, &fl);
          else if (P.MP == 24) hq_epilogue_table<24>(taddr, sCol + tg * 24, P.m, P.g, &code, &fl);
          else hq_epilogue_table<32>(taddr, sCol + tg * 32, P.m, P.g, &code, &fl);
          if (row < P.rows) {
            P.codes[(size_t)t * P.rows + row] = code;
            if (P.g > 0) P.flips[(size_t)t * P.rows + row] = fl;
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(bar_tempty + 8 * a);
      }
    }
    // every tile of the group is through its epilogue, hence through its MMAs: sB is free
    if (warp != 0) asm volatile("bar.sync 1, %0;" ::"n"(HQ_WORKERS) : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(P.tmem_cols) : "memory");
}

// ---------------------------------------------------------------- K3: CSR buckets
// Bucket key = code & (2^mb - 1), mb = min(m, 24) (full codes are re-checked in K4
// when m > 24).  All tables share one count/offset array of n * 2^mb entries and
// one id array of n * xrows entries, so a single scan serves every table.
__global__ void bucket_count_kernel(const int32_t* __restrict__ codes, size_t total, int xrows, uint32_t mask, int mb,
                                    uint32_t* __restrict__ cnt) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const size_t t = e / xrows;
  atomicAdd(&cnt[(t << mb) + ((uint32_t)codes[e] & mask)], 1u);
}

constexpr int SCAN_BLOCK = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total, uint32_t* ws) {
  // ws: SCAN_BLOCK / 32 + 1 words of shared memory
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) ws[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < SCAN_BLOCK / 32 ? ws[lane] : 0, winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, winc, d);
      if (lane >= d) winc += o;
    }
    if (lane < SCAN_BLOCK / 32) ws[lane] = winc - w;
    if (lane == 31) ws[SCAN_BLOCK / 32] = winc;
  }
  __syncthreads();
  *total = ws[SCAN_BLOCK / 32];
  const uint32_t r = ws[warp] + inc - v;
  __syncthreads();
  return r;
}

// phase 1: per-tile sums; phase 3 (apply = true): in-place exclusive scan + tile offset
__global__ void __launch_bounds__(SCAN_BLOCK) scan_tiles_kernel(uint32_t* __restrict__ data, size_t n,
                                                                uint32_t* __restrict__ tile_sums, bool apply) {
  __shared__ uint32_t ws[SCAN_BLOCK / 32 + 1];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) { v[i] = base + i < n ? data[base + i] : 0; s += v[i]; }
  uint32_t total;
  uint32_t ex = block_exclusive_scan(s, &total, ws);
  if (!apply) {
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
    return;
  }
  ex += tile_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    if (base + i < n) data[base + i] = ex;
    ex += v[i];
  }
}

// phase 2: exclusive scan of the tile sums by one block
__global__ void __launch_bounds__(SCAN_BLOCK) scan_sums_kernel(uint32_t* __restrict__ sums, int n) {
  __shared__ uint32_t ws[SCAN_BLOCK / 32 + 1];
  uint32_t carry = 0;
  for (int b0 = 0; b0 < n; b0 += SCAN_BLOCK) {
    const int i = b0 + threadIdx.x;
    const uint32_t v = i < n ? sums[i] : 0;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, &total, ws);
    if (i < n) sums[i] = carry + ex;
    carry += total;
  }
}

// scatter: cursor starts as the bucket's begin offset and ends as its end offset,
// so afterwards  begin(b) = b ? cursor[b - 1] : 0,  end(b) = cursor[b]  over the
// concatenated (table, bucket) axis.  Order inside a bucket is arbitrary; K4's
// result does not depend on it (lexicographic (dist, idx) top-2).
__global__ void bucket_scatter_kernel(const int32_t* __restrict__ codes, size_t total, int xrows, uint32_t mask, int mb,
                                      uint32_t* __restrict__ cursor, int32_t* __restrict__ ids) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const size_t t = e / xrows;
  const uint32_t pos = atomicAdd(&cursor[(t << mb) + ((uint32_t)codes[e] & mask)], 1u);
  ids[pos] = (int32_t)(e - t * xrows);
}

// ---------------------------------------------------------------- K4
constexpr int PR_WARPS = 4;      // warps per CTA
constexpr int PR_QPW = 4;        // query rows per warp (one per 8-lane group)
constexpr int PR_PAD = INT_MIN;  // a list slot that holds no candidate: negative, and `& INT_MAX` is row 0
constexpr int PR_LCAP = 256;     // candidate ids staged per query before a distance pass (multiple of 8)

struct PrParams {
  const uint8_t* xu8;            // [xrows][dim]
  const uint8_t* yu8;            // [yrows][dim]
  const int32_t* xcodes;         // [n][xrows]   (full codes, consulted only when m > mb)
  const int32_t* ycodes;         // [n][yrows]
  const unsigned long long* yflips;   // [n][yrows]
  const uint32_t* cursor;        // [n << mb] bucket END offsets over the concatenated axis
  const int32_t* ids;            // [n * xrows]
  int xrows, yrows, dim, n, m, mb, nb;
  uint32_t xoffset;
  unsigned long long* keys;      // [2][yrows]
};

// the two smallest DISTINCT keys of {a1 < a2} U {b1 < b2}; equal keys are the same
// (dist, idx) reached through two tables / probes -- the reference's unordered_set
// (CascadingHashNn.h:223) removes exactly those
__device__ __forceinline__ void top2_merge(unsigned long long& a1, unsigned long long& a2, unsigned long long b1,
                                           unsigned long long b2) {
  const unsigned long long lo = a1 < b1 ? a1 : b1;
  unsigned long long hi = a1 < b1 ? b1 : a1;
  if (hi == lo) hi = KEY_NONE;
  unsigned long long s = a2 < b2 ? a2 : b2;
  s = hi < s ? hi : s;
  a1 = lo;
  a2 = s;
}

// One 8-lane group per query row, four queries per warp, control flow warp-uniform.
// NCH = 16-byte chunks of a row per lane = ceil(dim / 128).  Per group:
//  (1) lanes enumerate the probe codes 8 at a time and read their bucket ranges;
//  (2) the ranges are FLATTENED: candidate slot f of the chunk belongs to the probe found by a
//      3-step shuffle search in the group's inclusive prefix of range lengths, so all id loads
//      of a chunk are independent and issued together (a lane walking its own bucket would
//      serialise one memory latency per candidate);
//  (3) ids are staged in shared memory and drained 8 at a time: the 8 row gathers of a step
//      (8 lanes x 16 B = one 128-byte line each) are all issued before the first is consumed,
//      each lane folds 4 VABSDIFF4 per row, and a transposed butterfly (7 shuffles for 8 rows
//      instead of 24) leaves candidate l8's distance in lane l8, which makes ONE branch-free
//      top-2 update per lane and step.
// The kernel is bound by the latency of dependent random reads (codes -> bucket range -> ids ->
// rows), so what counts is queries in flight per SM: 4 per warp x 32 warps at 64 registers.
// NCH = 0: rows wider than 512 bytes -- the chunk count is a run-time value and the query slices are
// re-read (L1-resident) instead of living in registers.
template <int NCH>
__global__ void __launch_bounds__(PR_WARPS * 32, NCH == 1 ? 8 : 4) probe_rerank_kernel(const PrParams P) {
  constexpr int NYV = NCH > 0 ? NCH : 1;
  constexpr int NQ = PR_WARPS * PR_QPW;                 // queries per CTA
  static_assert(NQ == 16, "the rank computation below takes two slots per lane of an 8-lane group");
  __shared__ __align__(16) int32_t lists[PR_WARPS][PR_QPW][PR_LCAP];
  __shared__ int Tq[NQ], order[NQ], ovf;
  __shared__ __align__(16) int32_t padv[8];             // eight PR_PADs: what a group with no ids left reads in a drain step
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = lane >> 3, l8 = lane & 7;
  const long long first = ((long long)blockIdx.x * PR_WARPS + warp) * PR_QPW;
  const int P_total = P.n << P.nb;   // probe codes of a query: n tables x 2^nb settings
  // With at most 8 probe codes (one chunk) the whole candidate list of a query is known before the first
  // gather, and the CTA's 16 lists can be dealt out again by length (below); CTA-uniform.
  const bool balanced = P_total <= 8;
  if (threadIdx.x == 0) ovf = 0;
  if (threadIdx.x < 8) padv[threadIdx.x] = PR_PAD;
  __syncthreads();
  if (!balanced && first >= P.yrows) return;   // whole warp
  bool live = first + q < P.yrows;
  int r = live ? (int)(first + q) : P.yrows - 1;   // a dead group shadows the last row and offers nothing
  const unsigned FULL = 0xFFFFFFFFu;
  const int nk16 = P.dim / 16;
  const uint32_t mask = (P.mb >= 32) ? 0xFFFFFFFFu : ((1u << P.mb) - 1u);
  const bool verify = P.m > P.mb;
  int32_t* list = lists[warp][q];

  // this lane's 16-byte slices of the query row: chunks l8, l8 + 8, ...
  uint4 yv[NYV];
  const uint4* yrow = reinterpret_cast<const uint4*>(P.yu8 + (size_t)r * P.dim);
  const int nch = NCH > 0 ? NCH : (nk16 + 7) / 8;
#pragma unroll
  for (int c = 0; c < NYV; ++c) yv[c] = (l8 + 8 * c < nk16) ? __ldg(yrow + l8 + 8 * c) : make_uint4(0, 0, 0, 0);

  unsigned long long k1 = KEY_NONE, k2 = KEY_NONE;   // this lane's running top-2 (distinct keys)
  const bool b2 = (l8 & 4) != 0, b1 = (l8 & 2) != 0, b0 = (l8 & 1) != 0;
  int listn = 0;                                     // ids staged for this group's query

  auto drain = [&]() {
    if ((listn & 7) && l8 >= (listn & 7)) list[(listn & ~7) + l8] = PR_PAD;   // pad the last step
    __syncwarp();
    const int nmax = __reduce_max_sync(FULL, listn);
    // how many steps ahead rows are asked into L2 (below).  Measured on C5, K4 alone: none 1.381 ms, 1 step 1.332,
    // 2 steps 1.358, 4 steps 1.397, 8 steps 1.445 -- beyond one step the requests only queue up behind each other
    constexpr int PF_DIST = 1;
    for (int c0 = 0; c0 < nmax; c0 += 8) {
      // a group whose list is exhausted reads eight pads instead: one address select, no predicated loads
      const int32_t* src = c0 < listn ? list + c0 : padv;
      const int4 ia = *reinterpret_cast<const int4*>(src), ib = *reinterpret_cast<const int4*>(src + 4);
      const int own = src[l8];            // the candidate whose distance ends up in this lane
      // A slot without a candidate (PR_PAD) gathers row 0 instead of being predicated off: its distance is never
      // looked at (`own < 0` makes the key KEY_NONE below), row 0 stays in L1, and the step loses 8 predicates and the
      // 32 register moves that gave the skipped loads a value (SASS: 145 -> ~125 instructions per step).  The mask is
      // a LOP3; max(id, 0) would be a VIMNMX, which issues at a third of the rate (pipe_peak.cu).
      const uint32_t id[8] = {(uint32_t)ia.x & 0x7FFFFFFFu, (uint32_t)ia.y & 0x7FFFFFFFu, (uint32_t)ia.z & 0x7FFFFFFFu,
                              (uint32_t)ia.w & 0x7FFFFFFFu, (uint32_t)ib.x & 0x7FFFFFFFu, (uint32_t)ib.y & 0x7FFFFFFFu,
                              (uint32_t)ib.z & 0x7FFFFFFFu, (uint32_t)ib.w & 0x7FFFFFFFu};
      uint32_t d[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) d[u] = 0;
      // The NEXT step's rows are asked into L2 now, one row per lane (ncu r02: the kernel waits on the gathers --
      // long_scoreboard 9 per issue, 44 % of issue slots used -- and a prefetch holds no registers)
      if (c0 + 8 * PF_DIST + l8 < listn) {
        const int nid = list[c0 + 8 * PF_DIST + l8];
        if (nid >= 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.xu8 + (unsigned long long)(uint32_t)nid * (uint32_t)P.dim));
      }
#pragma unroll
      for (int c = 0; c < nch; ++c) {
        if (l8 + 8 * c < nk16) {          // this lane holds a 16-byte slice of chunk c (all lanes when dim % 128 == 0)
          const uint4 yc = NCH > 0 ? yv[NCH > 0 ? c : 0] : __ldg(yrow + l8 + 8 * c);
          const uint4* xs = reinterpret_cast<const uint4*>(P.xu8) + l8 + 8 * c;
          uint4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u)     // 32 x 32 -> 64-bit product: one IMAD.WIDE
            v[u] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(xs) + (unsigned long long)id[u] * (uint32_t)P.dim));
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            d[u] = sad4(v[u].x, yc.x, d[u]); d[u] = sad4(v[u].y, yc.y, d[u]);
            d[u] = sad4(v[u].z, yc.z, d[u]); d[u] = sad4(v[u].w, yc.w, d[u]);
          }
        }
      }
      // transposed butterfly over the 8 lanes of the group: lane l8 ends with sum_lanes d[l8]
      uint32_t e[4], f[2];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t send = b2 ? d[u] : d[u + 4], keep = b2 ? d[u + 4] : d[u];
        e[u] = keep + __shfl_xor_sync(FULL, send, 4);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const uint32_t send = b1 ? e[u] : e[u + 2], keep = b1 ? e[u + 2] : e[u];
        f[u] = keep + __shfl_xor_sync(FULL, send, 2);
      }
      const uint32_t send = b0 ? f[0] : f[1], keep = b0 ? f[1] : f[0];
      const uint32_t dist = keep + __shfl_xor_sync(FULL, send, 1);
      // branch-free insert of one key into the distinct top-2
      const unsigned long long key =
          own >= 0 ? (((unsigned long long)dist << 32) | (uint32_t)(P.xoffset + (uint32_t)own)) : KEY_NONE;
      const unsigned long long lo = key < k1 ? key : k1, hi = key < k1 ? k1 : key;
      k2 = (hi == lo) ? k2 : (hi < k2 ? hi : k2);
      k1 = lo;
    }
    __syncwarp();
    listn = 0;
  };

  for (int p0 = 0; p0 < P_total; p0 += 8) {
    const int p = p0 + l8;
    uint32_t start = 0, len = 0, code = 0;
    int t = 0;
    if (live && p < P_total) {
      t = p >> P.nb;
      const uint32_t combo = (uint32_t)p & ((1u << P.nb) - 1u);
      code = (uint32_t)__ldg(P.ycodes + (size_t)t * P.yrows + r);
      const unsigned long long fl = __ldg(P.yflips + (size_t)t * P.yrows + r);
      for (int b = 0; b < P.nb; ++b) {   // generate_y_candidate_hashcodes :170-179
        const uint32_t bit = (uint32_t)(fl >> (5 * b)) & 31u;
        code = (code & ~(1u << bit)) | (((combo >> b) & 1u) << bit);
      }
      const size_t bk = ((size_t)t << P.mb) + (code & mask);
      const uint32_t end = __ldg(P.cursor + bk);
      start = bk ? __ldg(P.cursor + bk - 1) : 0u;
      len = end - start;
    }
    // inclusive prefix of len over the 8 lanes of the group
    uint32_t inc = len;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) {
      const uint32_t o = __shfl_up_sync(FULL, inc, d, 8);
      if (l8 >= d) inc += o;
    }
    const uint32_t T = __shfl_sync(FULL, inc, 7, 8), excl = inc - len;
    uint32_t done = 0;
    bool redeal = false;
    if (balanced) {
      if (l8 == 0) Tq[warp * PR_QPW + q] = (int)T;
      if (T > (uint32_t)PR_LCAP) ovf = 1;
      __syncthreads();
      redeal = ovf == 0;   // every list of the CTA fits: fill them all, then deal them out by length
    }
    if (!__any_sync(FULL, (uint32_t)listn + T > (uint32_t)PR_LCAP)) {
      // common case, the chunk fits every group's list: copy bucket by bucket, the 8 lanes of a group striding
      // one bucket.  The first 8 ids of all 8 buckets are loaded before any is stored (8 independent loads in
      // flight); longer buckets (mean occupancy is ~6-8 by the auto-m rule) finish in a short loop.
      uint32_t b_start[8], b_len[8], b_excl[8];
      int idv[8];
#pragma unroll
      for (int pp = 0; pp < 8; ++pp) {
        b_start[pp] = __shfl_sync(FULL, start, pp, 8);
        b_len[pp] = __shfl_sync(FULL, len, pp, 8);
        b_excl[pp] = __shfl_sync(FULL, excl, pp, 8);
        idv[pp] = ((uint32_t)l8 < b_len[pp]) ? __ldg(P.ids + b_start[pp] + l8) : -1;
      }
#pragma unroll
      for (int pp = 0; pp < 8; ++pp) {
        uint32_t s_code = 0;
        int s_t = 0;
        if (verify) { s_code = __shfl_sync(FULL, code, pp, 8); s_t = __shfl_sync(FULL, t, pp, 8); }   // warp-uniform branch
        if ((uint32_t)l8 < b_len[pp]) {
          int id = idv[pp];
          if (verify && (uint32_t)__ldg(P.xcodes + (size_t)s_t * P.xrows + id) != s_code) id = PR_PAD;
          list[listn + b_excl[pp] + l8] = id;
        }
      }
      if (__reduce_max_sync(FULL, len) > 8u) {
#pragma unroll
        for (int pp = 0; pp < 8; ++pp) {
          const uint32_t lmax = __reduce_max_sync(FULL, b_len[pp]);
          uint32_t s_code = 0;
          int s_t = 0;
          if (verify) { s_code = __shfl_sync(FULL, code, pp, 8); s_t = __shfl_sync(FULL, t, pp, 8); }
          for (uint32_t k0 = 8; k0 < lmax; k0 += 8) {
            const uint32_t k = k0 + l8;
            if (k < b_len[pp]) {
              int id = __ldg(P.ids + b_start[pp] + k);
              if (verify && (uint32_t)__ldg(P.xcodes + (size_t)s_t * P.xrows + id) != s_code) id = PR_PAD;
              list[listn + b_excl[pp] + k] = id;
            }
          }
        }
      }
      listn += (int)T;
      done = T;
    }
    if (redeal) {
      // Load balance: a warp drains in lock step, so its four groups all pay for the longest list (measured:
      // 16.3 steps per warp where 10.6 would carry the candidates).  Sort the CTA's 16 lists by length and give
      // warp w the ranks 4w .. 4w+3, so lists of similar length share a warp.
      __syncthreads();                                  // all 16 lists are complete
      const int s = warp * PR_QPW + q, Ts = Tq[s];
      const int ja = l8, jb = l8 + 8;
      int rank = (Tq[ja] > Ts || (Tq[ja] == Ts && ja < s)) + (Tq[jb] > Ts || (Tq[jb] == Ts && jb < s));
      rank += __shfl_xor_sync(FULL, rank, 1);
      rank += __shfl_xor_sync(FULL, rank, 2);
      rank += __shfl_xor_sync(FULL, rank, 4);
      if (l8 == 0) order[rank] = s;
      __syncthreads();
      const int mine = order[s];                        // the query this group drains
      list = &lists[0][0][0] + mine * PR_LCAP;
      listn = Tq[mine];
      const long long row = (long long)blockIdx.x * NQ + mine;
      live = row < P.yrows;
      r = live ? (int)row : P.yrows - 1;
      yrow = reinterpret_cast<const uint4*>(P.yu8 + (size_t)r * P.dim);
#pragma unroll
      for (int c = 0; c < NYV; ++c) yv[c] = (l8 + 8 * c < nk16) ? __ldg(yrow + l8 + 8 * c) : make_uint4(0, 0, 0, 0);
    }
    while (__any_sync(FULL, done < T)) {
      if (__any_sync(FULL, done < T && listn == PR_LCAP)) drain();   // some group is full: all groups drain
      const uint32_t room = (uint32_t)(PR_LCAP - listn), take = (T - done) < room ? (T - done) : room;
      const uint32_t takemax = __reduce_max_sync(FULL, take);
      for (uint32_t f0 = 0; f0 < takemax; f0 += 8) {
        const uint32_t f = done + f0 + l8;
        // owner probe of flat slot f: the number of group lanes whose inclusive prefix is <= f
        int pos = 0;
#pragma unroll
        for (int step = 4; step >= 1; step >>= 1) {
          const uint32_t v = __shfl_sync(FULL, inc, pos + step - 1, 8);
          pos += (v <= f) ? step : 0;
        }
        pos = pos > 7 ? 7 : pos;
        const uint32_t s_start = __shfl_sync(FULL, start, pos, 8), s_excl = __shfl_sync(FULL, excl, pos, 8);
        uint32_t s_code = 0;
        int s_t = 0;
        if (verify) { s_code = __shfl_sync(FULL, code, pos, 8); s_t = __shfl_sync(FULL, t, pos, 8); }
        if (f0 + l8 < take) {
          int id = __ldg(P.ids + s_start + (f - s_excl));
          if (verify && (uint32_t)__ldg(P.xcodes + (size_t)s_t * P.xrows + id) != s_code) id = PR_PAD;
          list[listn + f0 + l8] = id;
        }
      }
      listn += (int)take;
      done += take;
    }
  }
  if (__any_sync(FULL, listn != 0)) drain();

  // every lane of the group holds the top-2 of a disjoint-or-duplicate candidate subset
#pragma unroll
  for (int d = 1; d <= 4; d <<= 1) {
    const unsigned long long o1 = __shfl_xor_sync(FULL, k1, d), o2 = __shfl_xor_sync(FULL, k2, d);
    top2_merge(k1, k2, o1, o2);
  }
  if (l8 == 0 && live) {
    // fold into what other train shards / launches left there: the same two-atomicMin protocol as K1's
    // fold_key (the loser of the rank-0 exchange is offered to rank 1; a key that is already the best is
    // not offered again), so folds into one key buffer may run concurrently on any streams
    const unsigned long long ks[2] = {k1, k2};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const unsigned long long k = ks[i];
      if (k == KEY_NONE) continue;
      const unsigned long long old = atomicMin(P.keys + r, k);
      if (old == k) continue;
      const unsigned long long loser = old > k ? old : k;
      if (loser != KEY_NONE) atomicMin(P.keys + (size_t)P.yrows + r, loser);
    }
  }
}

}  // namespace

// ---------------------------------------------------------------- host side
namespace {

struct Plan {
  int MP, tpg, N, groups, tmem_cols, mb, nb, terms;
  size_t smem_hp;
  // workspace offsets (bytes)
  size_t o_img, o_xcodes, o_ycodes, o_yflips, o_cursor, o_ids, o_sums, total;
  size_t ncursor;
  int nsum_tiles;
};

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

bool make_plan(int xrows, int yrows, int dim, int m, int n, int g, int terms, Plan* pl) {
  if (dim < 16 || dim % 16 != 0 || m < 1 || m > 31 || n < 1 || g < 0 || xrows < 0 || yrows < 0) return false;
  // bucket counts, scan, cursor and id positions are uint32 / int32 over the concatenated n * xrows axis
  if ((unsigned long long)n * (unsigned long long)(xrows > 0 ? xrows : 1) >= (1ull << 31)) return false;
  pl->terms = terms;
  pl->nb = g < m ? g : m;                       // :153-159 keeps at most m pairs
  if (pl->nb > 12 || ((long long)n << pl->nb) > (1ll << 24)) return false;
  pl->mb = m < 24 ? m : 24;
  pl->MP = (m + 7) / 8 * 8;
  const int nkc = dim / 8;
  const size_t a_bytes = (size_t)(HP_KPANEL / 8) * HP_ROWS * 16, budget = 190 * 1024;
  int nmax = (int)((budget - a_bytes - 64) / ((size_t)terms * nkc * 16)) / 16 * 16;   // shrinks with dim: wide rows take fewer tables per pass
  if (nmax > 256) nmax = 256;
  int tpg = nmax / pl->MP;
  if (tpg > n) tpg = n;
  if (tpg > 4 && n > 4) tpg = 4;                // keep the B image small enough for several CTAs per SM
  if (tpg < 1) return false;
  pl->tpg = tpg;
  pl->N = (tpg * pl->MP + 15) / 16 * 16;
  pl->groups = (n + tpg - 1) / tpg;
  int cols = 32;
  while (cols < pl->N) cols <<= 1;
  pl->tmem_cols = cols;
  pl->smem_hp = (size_t)terms * nkc * pl->N * 16 + a_bytes + 64;
  pl->ncursor = (size_t)n << pl->mb;
  pl->nsum_tiles = (int)((pl->ncursor + SCAN_TILE - 1) / SCAN_TILE);
  size_t o = 0;
  pl->o_img = o;     o = align256(o + (size_t)pl->groups * 3 * nkc * pl->N * 8 * sizeof(__nv_bfloat16));
  pl->o_xcodes = o;  o = align256(o + (size_t)n * (xrows > 0 ? xrows : 1) * 4);
  pl->o_ycodes = o;  o = align256(o + (size_t)n * (yrows > 0 ? yrows : 1) * 4);
  pl->o_yflips = o;  o = align256(o + (size_t)n * (yrows > 0 ? yrows : 1) * 8);
  pl->o_cursor = o;  o = align256(o + pl->ncursor * 4);
  pl->o_ids = o;     o = align256(o + (size_t)n * (xrows > 0 ? xrows : 1) * 4);
  pl->o_sums = o;    o = align256(o + (size_t)pl->nsum_tiles * 4 + 4);
  pl->total = o;
  return true;
}

// shared memory and ring depth of the int8 kernel; nst = 0: does not fit, use the bf16 kernel
size_t i8_smem(const Plan& pl, int dim, int* nst_out) {
  const size_t stage = (size_t)HP_ROWS * dim, fixed = (size_t)dim * pl.N + 4 * (size_t)(pl.N + 1) + 8 * 24 + 64 + 1024;
  int nst = 0;
  if (fixed + 2 * stage <= 200 * 1024) {
    nst = (int)((100 * 1024 > fixed ? 100 * 1024 - fixed : 0) / stage);
    nst = nst < 2 ? 2 : (nst > 8 ? 8 : nst);
  }
  *nst_out = nst;
  return fixed + (size_t)nst * stage;
}

// dict_terms 0 (int8-exact dictionary): the int8 kernel when its tiles fit, else the bf16 kernel (1 term is exact too)
bool plan_for(int xrows, int yrows, int dim, int m, int n, int g, int dict_terms, Plan* pl) {
  if (!make_plan(xrows, yrows, dim, m, n, g, dict_terms == 0 ? 1 : dict_terms, pl)) return false;
  if (dict_terms == 0) {
    int nst = 0;
    i8_smem(*pl, dim, &nst);
    if (nst > 0 && dim % 32 == 0) pl->terms = 0;
  }
  return true;
}

cudaError_t run_project_i8(const Plan& pl, const uint8_t* rows_u8, int rows, int dim, int n, int m, int g, const void* img,
                           int32_t* codes, unsigned long long* flips, int sms, cudaStream_t stream) {
  int nst = 0;
  const size_t smem = i8_smem(pl, dim, &nst);
  HqParams hq;
  hq.img = static_cast<const int8_t*>(img);
  hq.colsum = reinterpret_cast<const int32_t*>(static_cast<const unsigned char*>(img) + align256((size_t)pl.groups * dim * pl.N));
  hq.rows = rows; hq.dim = dim; hq.n = n; hq.m = m; hq.MP = pl.MP; hq.tpg = pl.tpg; hq.N = pl.N; hq.groups = pl.groups;
  hq.g = g; hq.nst = nst; hq.codes = codes; hq.flips = flips;
  int cols = 32;
  while (cols < 2 * pl.N) cols <<= 1;
  hq.tmem_cols = cols;
  CUtensorMap tmA;
  const char* k2 = getenv("SPECTAVI_K2");
  hq.swz = (dim % 128 == 0 && !(k2 && strcmp(k2, "legacy") == 0)) ? 1 : 0;
  if (hq.swz ? !make_rows_map_sw128(&tmA, rows_u8, rows, dim, HP_ROWS) : !make_rows_map(&tmA, rows_u8, rows, dim, HP_ROWS))
    return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(hash_project_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int per_sm = (smem <= 110 * 1024 && cols <= 256) ? 2 : 1;
  const int ntiles = (rows + HP_ROWS - 1) / HP_ROWS;
  const int grid = ntiles < sms * per_sm ? ntiles : sms * per_sm;
  hash_project_i8_kernel<<<grid, HQ_THREADS, smem, stream>>>(tmA, hq);
  count_launch();
  return cudaGetLastError();
}

cudaError_t run_project(const Plan& pl, const uint8_t* rows_u8, int rows, int dim, int n, int m, int g,
                        const __nv_bfloat16* img, int32_t* codes, unsigned long long* flips, int sms, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  if (pl.terms == 0) return run_project_i8(pl, rows_u8, rows, dim, n, m, g, img, codes, flips, sms, stream);
  HpParams hp;
  hp.rows_u8 = rows_u8; hp.img = img; hp.rows = rows; hp.dim = dim; hp.n = n; hp.m = m; hp.MP = pl.MP; hp.tpg = pl.tpg;
  hp.N = pl.N; hp.groups = pl.groups; hp.terms = pl.terms; hp.g = g; hp.codes = codes; hp.flips = flips;
  hp.tmem_cols = pl.tmem_cols;
  cudaError_t e = cudaFuncSetAttribute(hash_project_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_hp);
  if (e != cudaSuccess) return e;
  int per_sm = (int)((200 * 1024) / (pl.smem_hp + 1024));
  if (per_sm > 512 / pl.tmem_cols) per_sm = 512 / pl.tmem_cols;   // TMEM columns are a per-SM resource
  if (per_sm > 4) per_sm = 4;
  if (per_sm < 1) per_sm = 1;
  const int ntiles = (rows + HP_ROWS - 1) / HP_ROWS;
  const int grid = ntiles < sms * per_sm ? ntiles : sms * per_sm;
  hash_project_kernel<<<grid, HP_THREADS, pl.smem_hp, stream>>>(hp);
  count_launch();
  return cudaGetLastError();
}

void build_image(const Plan& pl, const float* ddict, int n, int dim, int m, __nv_bfloat16* img, cudaStream_t stream) {
  if (pl.terms == 0) {
    int8_t* q = reinterpret_cast<int8_t*>(img);
    int32_t* cs = reinterpret_cast<int32_t*>(reinterpret_cast<unsigned char*>(img) + align256((size_t)pl.groups * dim * pl.N));
    dict_image_i8_kernel<<<64, 256, 0, stream>>>(ddict, n, dim, m, pl.MP, pl.tpg, pl.N, pl.groups, q, cs);
  } else {
    dict_image_kernel<<<64, 256, 0, stream>>>(ddict, n, dim, m, pl.MP, pl.tpg, pl.N, img);
  }
  count_launch();
}

}  // namespace

size_t cascade_ws_bytes(int xrows, int yrows, int dim, int m, int n, int g) {
  // the layout depends on the number of bf16 terms (tables per pass -> padded image size); a caller that
  // allocates exactly this many bytes must be safe for every dict_terms, so: the largest of the plans
  Plan p1, p3;
  const bool ok1 = make_plan(xrows, yrows, dim, m, n, g, 1, &p1), ok3 = make_plan(xrows, yrows, dim, m, n, g, 3, &p3);
  if (!ok1) return 0;
  return ok3 && p3.total > p1.total ? p3.total : p1.total;
}

cudaError_t f32_to_u8(const float* din, uint8_t* dout, size_t count, cudaStream_t stream) {
  if (count == 0) return cudaSuccess;
  if (count % 4) return cudaErrorInvalidValue;   // rows are multiples of 16 columns
  const size_t n4 = count / 4;
  f32_to_u8_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const float4*>(din),
                                                                      reinterpret_cast<uint32_t*>(dout), n4);
  count_launch();
  return cudaGetLastError();
}

// K2 alone (tests, profiling): codes [n][rows] and, when g > 0, flips [n][rows]
cudaError_t hash_codes(const uint8_t* drows, int rows, int dim, int m, int n, int g, const float* ddict, int dict_terms,
                       int32_t* codes, unsigned long long* flips, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  Plan pl;
  if (dict_terms != 0 && dict_terms != 1 && dict_terms != 3) return cudaErrorInvalidValue;
  if (!plan_for(rows, rows, dim, m, n, g, dict_terms, &pl) || ws_bytes < pl.total) return cudaErrorInvalidValue;
  int dev = 0, sms = 0;
  cudaError_t e;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  __nv_bfloat16* img = reinterpret_cast<__nv_bfloat16*>(static_cast<unsigned char*>(ws) + pl.o_img);
  build_image(pl, ddict, n, dim, m, img, stream);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  return run_project(pl, drows, rows, dim, n, m, g > 0 ? pl.nb : 0, img, codes, flips, sms, stream);
}

// Optional stage timing (SPECTAVI_STAGE_TIMING=1): CUDA events on the launching stream at the stage boundaries of the
// calling thread's last cascade_match; read back with cascade_stage_ms().  Stages: dictionary image, K2 train side,
// K3 (count + scan + scatter), K2 query side, K4.
namespace {
constexpr int NSTAGE = 5;
thread_local cudaEvent_t t_ev[NSTAGE + 1];
thread_local bool t_ev_made = false, t_ev_valid = false;
bool stage_timing_on() {
  static const bool on = getenv("SPECTAVI_STAGE_TIMING") != nullptr;
  return on;
}
void stage_mark(int i, cudaStream_t stream) {
  if (!stage_timing_on()) return;
  if (!t_ev_made) {
    for (int e = 0; e <= NSTAGE; ++e) cudaEventCreate(&t_ev[e]);
    t_ev_made = true;
  }
  cudaEventRecord(t_ev[i], stream);
  if (i == NSTAGE) t_ev_valid = true;
}
}  // namespace

int cascade_stage_ms(double* out, int n) {
  if (!t_ev_valid || n < NSTAGE) return 0;
  if (cudaEventSynchronize(t_ev[NSTAGE]) != cudaSuccess) return 0;
  for (int i = 0; i < NSTAGE; ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, t_ev[i], t_ev[i + 1]) != cudaSuccess) return 0;
    out[i] = ms;
  }
  return NSTAGE;
}

cudaError_t cascade_match(const uint8_t* dxu8, const uint8_t* dyu8, int xrows, int yrows, int dim, int m, int n, int g,
                          const float* ddict, int dict_terms, uint32_t xoffset, unsigned long long* keys, void* ws,
                          size_t ws_bytes, cudaStream_t stream) {
  if (yrows <= 0 || xrows <= 0) return cudaSuccess;   // nothing to fold in
  Plan pl;
  if (dict_terms != 0 && dict_terms != 1 && dict_terms != 3) return cudaErrorInvalidValue;
  if (!plan_for(xrows, yrows, dim, m, n, g, dict_terms, &pl) || ws_bytes < pl.total) return cudaErrorInvalidValue;
  if ((unsigned long long)xoffset + (unsigned long long)xrows > 0xFFFFFFFFull) return cudaErrorInvalidValue;   // 32 index bits per key
  int dev = 0, sms = 0;
  cudaError_t e;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  unsigned char* w = static_cast<unsigned char*>(ws);
  __nv_bfloat16* img = reinterpret_cast<__nv_bfloat16*>(w + pl.o_img);
  int32_t* xcodes = reinterpret_cast<int32_t*>(w + pl.o_xcodes);
  int32_t* ycodes = reinterpret_cast<int32_t*>(w + pl.o_ycodes);
  unsigned long long* yflips = reinterpret_cast<unsigned long long*>(w + pl.o_yflips);
  uint32_t* cursor = reinterpret_cast<uint32_t*>(w + pl.o_cursor);
  int32_t* ids = reinterpret_cast<int32_t*>(w + pl.o_ids);
  uint32_t* sums = reinterpret_cast<uint32_t*>(w + pl.o_sums);

  stage_mark(0, stream);
  build_image(pl, ddict, n, dim, m, img, stream);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  stage_mark(1, stream);
  // train side: codes, then CSR buckets (initialize_hash_tables, :187-196)
  if ((e = run_project(pl, dxu8, xrows, dim, n, m, 0, img, xcodes, nullptr, sms, stream)) != cudaSuccess) return e;
  stage_mark(2, stream);
  if ((e = cudaMemsetAsync(cursor, 0, pl.ncursor * 4, stream)) != cudaSuccess) return e;
  const size_t total = (size_t)n * xrows;
  const uint32_t mask = (1u << pl.mb) - 1u;
  bucket_count_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(xcodes, total, xrows, mask, pl.mb, cursor);
  scan_tiles_kernel<<<pl.nsum_tiles, SCAN_BLOCK, 0, stream>>>(cursor, pl.ncursor, sums, false);
  scan_sums_kernel<<<1, SCAN_BLOCK, 0, stream>>>(sums, pl.nsum_tiles);
  scan_tiles_kernel<<<pl.nsum_tiles, SCAN_BLOCK, 0, stream>>>(cursor, pl.ncursor, sums, true);
  bucket_scatter_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(xcodes, total, xrows, mask, pl.mb, cursor, ids);
  count_launch(5);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  stage_mark(3, stream);
  // query side: codes + least-confident bits, then probe / gather / re-rank
  if ((e = run_project(pl, dyu8, yrows, dim, n, m, pl.nb > 0 ? pl.nb : 0, img, ycodes, yflips, sms, stream)) != cudaSuccess) return e;
  if (pl.nb == 0 && (e = cudaMemsetAsync(yflips, 0, (size_t)n * yrows * 8, stream)) != cudaSuccess) return e;
  stage_mark(4, stream);
  PrParams pr;
  pr.xu8 = dxu8; pr.yu8 = dyu8; pr.xcodes = xcodes; pr.ycodes = ycodes; pr.yflips = yflips; pr.cursor = cursor; pr.ids = ids;
  pr.xrows = xrows; pr.yrows = yrows; pr.dim = dim; pr.n = n; pr.m = m; pr.mb = pl.mb; pr.nb = pl.nb; pr.xoffset = xoffset;
  pr.keys = keys;
  const unsigned prgrid = (unsigned)((yrows + PR_WARPS * PR_QPW - 1) / (PR_WARPS * PR_QPW));
  switch ((dim + 127) / 128) {
    case 1: probe_rerank_kernel<1><<<prgrid, PR_WARPS * 32, 0, stream>>>(pr); break;
    case 2: probe_rerank_kernel<2><<<prgrid, PR_WARPS * 32, 0, stream>>>(pr); break;
    case 3: probe_rerank_kernel<3><<<prgrid, PR_WARPS * 32, 0, stream>>>(pr); break;
    case 4: probe_rerank_kernel<4><<<prgrid, PR_WARPS * 32, 0, stream>>>(pr); break;
    default: probe_rerank_kernel<0><<<prgrid, PR_WARPS * 32, 0, stream>>>(pr); break;
  }
  count_launch();
  stage_mark(5, stream);
  return cudaGetLastError();
}

}  // namespace spv
