// pipe_peak.cu -- issue-rate microbenchmarks for the instructions the L1 top-2
// kernel (l1_top2.cu) can be built from.  MEASUREMENT TOOL, not on the product
// path: bench.py calls spv_pipe_peak() to obtain the roofline denominator
// P_sad4 (packed-4xu8 SAD-accumulate equivalents per second) on
// the box the benchmark runs on.
//
// Every variant runs `iters` x UNROLL independent-chain instructions per
// thread in a fully resident grid (148 SMs x CTAS x 256 threads) and reports
//   * ops/clk/SM from clock64() deltas (frequency independent), and
//   * G ops/s from CUDA-event time (what the power/clock state really gave).
// An "op" is one thread-level instruction of the kind under test; for the SAD
// formulations the table also gives byte-lanes per clk per SM.
#include <cuda_runtime.h>
#include "../../include/spectavi_b200.h"
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "pipe_peak: %s failed: %s\n", #x, cudaGetErrorString(e_)); return -1; } } while (0)

namespace {

constexpr int NACC = 32;   // independent dependency chains per thread

enum Variant {
  V_SAD4_ACC = 0,      // VABSDIFF4.U8.ACC d, a, b, d
  V_SAD4_DP4A,         // VABSDIFF4.U8 + IDP.4A (ones)          (2 instr / 4 lanes)
  V_HALF2_ABS,         // HFMA2 (a-b) + HADD2 acc,|d|           (2 instr / 2 lanes)
  V_IADD3,
  V_LOP3,
  V_IMAD,
  V_FFMA,
  V_VIMNMX,
  V_VIMNMX3,
  V_DP4A,
  V_SAD4_ACC_IMAD,     // 1 IMAD per 4 SAD4 (does the fma pipe co-issue?)
  V_SAD4_ACC_IMAD11,   // 1 IMAD per 1 SAD4
  V_SAD4_TILE,         // 8x4 register tile, operands from shared memory (LDS.128)
  V_SAD_SCALAR,        // VABSDIFF.U32 (scalar sad)
  V_FADD,              // FADD acc, acc, a
  V_FSUB_ABSADD,       // FADD d, a, -b ; FADD acc, acc, |d|   (K1f's p = 1 pair-column, lp_topk.cu)
  V_FSUB_FMUL_FADD,    // FADD d ; FMUL d*d ; FADD acc         (K1f's p = 2 pair-column, no FMA contraction)
  V_FADD2,             // add.rn.f32x2 (SASS FADD2): two IEEE-rn float adds per instruction
  V_F2_SUB_ABSADD,     // FADD2 d, a, -b ; FADD2 acc, acc, |d|  (two p = 1 pair-columns per instruction pair)
  V_F2_SUB_MUL_ADD,    // FADD2 ; FMUL ; FMUL ; FADD2           (two p = 2 pair-columns per four instructions)
  V_DFMA,              // fma.rn.f64
  V_DMUL_DADD,         // mul.rn.f64 ; add.rn.f64               (K8's arithmetic: no contraction, ransac.cu)
  V_COUNT
};

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

template <int V>
__global__ void __launch_bounds__(256, 2) peak_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                   int iters, long long* __restrict__ cycles) {
  __shared__ uint4 sm[512];
  uint32_t acc[NACC];
  uint32_t a[8], b[8];
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = in[(t + i * 7) & 1023];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = in[(t * 3 + i) & 1023]; b[i] = in[(t * 5 + i + 11) & 1023]; }
  for (int i = t; i < 512; i += 256) sm[i] = make_uint4(in[i & 1023], in[(i + 1) & 1023], in[(i + 2) & 1023], in[(i + 3) & 1023]);
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (V == V_SAD4_ACC) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) acc[i] = sad4(a[i & 7], b[(i >> 3) & 7], acc[i]);
    } else if (V == V_SAD4_DP4A) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        uint32_t d;
        asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a[i & 7]), "r"(b[(i >> 3) & 7]), "r"(0u));
        acc[i] = __dp4a(d, 0x01010101u, acc[i]);
      }
    } else if (V == V_HALF2_ABS) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        __half2 x = *reinterpret_cast<__half2*>(&a[i & 7]);
        __half2 y = *reinterpret_cast<__half2*>(&b[(i >> 3) & 7]);
        __half2 c = *reinterpret_cast<__half2*>(&acc[i]);
        c = __hadd2(c, __habs2(__hsub2(x, y)));
        acc[i] = *reinterpret_cast<uint32_t*>(&c);
      }
    } else if (V == V_IADD3) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("add.u32 %0, %0, %1;" : "+r"(acc[i]) : "r"(a[i & 7]));
    } else if (V == V_LOP3) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(acc[i]) : "r"(a[i & 7]), "r"(b[(i >> 3) & 7]));
    } else if (V == V_IMAD) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(acc[i]) : "r"(a[i & 7]), "r"(b[(i >> 3) & 7]));
    } else if (V == V_FFMA) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(*reinterpret_cast<float*>(&acc[i])) : "f"(__uint_as_float(a[i & 7])), "f"(__uint_as_float(b[(i >> 3) & 7])));
    } else if (V == V_VIMNMX) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("min.u32 %0, %0, %1;" : "+r"(acc[i]) : "r"(a[(i + it) & 7]));
    } else if (V == V_VIMNMX3) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) acc[i] = min(min(acc[i], a[(i + it) & 7]), b[(i >> 3) & 7]);
    } else if (V == V_DP4A) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) acc[i] = __dp4a(a[i & 7], b[(i >> 3) & 7], acc[i]);
    } else if (V == V_SAD4_ACC_IMAD) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        acc[i] = sad4(a[i & 7], b[(i >> 3) & 7], acc[i]);
        if ((i & 3) == 3) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[(i >> 2) & 7]) : "r"(b[1]), "r"(b[2]));
      }
    } else if (V == V_SAD4_ACC_IMAD11) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        acc[i] = sad4(a[i & 7], b[(i >> 3) & 7], acc[i]);
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i & 7]) : "r"(b[1]), "r"(b[2]));
      }
    } else if (V == V_SAD4_TILE) {
      // 8 (rows) x 4 (cols) accumulators; per k4-step 8+4 LDS.128, 128 SAD4
      const uint4* pa = sm + ((t >> 3) & 31) * 8 + (it & 1);
      const uint4* pb = sm + 256 + (t & 7) * 4 + (it & 1) * 64;
      uint4 ra[8], rb[4];
#pragma unroll
      for (int i = 0; i < 8; ++i) ra[i] = pa[i * 2];
#pragma unroll
      for (int j = 0; j < 4; ++j) rb[j] = pb[j * 2];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t c = acc[i * 4 + j];
          c = sad4(ra[i].x, rb[j].x, c);
          c = sad4(ra[i].y, rb[j].y, c);
          c = sad4(ra[i].z, rb[j].z, c);
          c = sad4(ra[i].w, rb[j].w, c);
          acc[i * 4 + j] = c;
        }
    } else if (V == V_SAD_SCALAR) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) acc[i] = __usad(a[i & 7], b[(i >> 3) & 7], acc[i]);
    } else if (V == V_FADD) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(*reinterpret_cast<float*>(&acc[i])) : "f"(__uint_as_float(a[i & 7])));
    } else if (V == V_FSUB_ABSADD) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        float d;
        // the subtrahend is another chain's accumulator so that the subtraction cannot be hoisted out of the loop
        asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(__uint_as_float(a[i & 7])), "f"(__uint_as_float(acc[(i + 16) & (NACC - 1)])));
        asm volatile("{ .reg .f32 t; abs.f32 t, %1; add.rn.f32 %0, %0, t; }" : "+f"(*reinterpret_cast<float*>(&acc[i])) : "f"(d));
      }
    } else if (V == V_FSUB_FMUL_FADD) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        float d;
        asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(__uint_as_float(a[i & 7])), "f"(__uint_as_float(acc[(i + 16) & (NACC - 1)])));
        asm volatile("mul.rn.f32 %0, %0, %0;" : "+f"(d));
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(*reinterpret_cast<float*>(&acc[i])) : "f"(d));
      }
    } else if (V == V_FADD2) {
#pragma unroll
      for (int i = 0; i < NACC; i += 2) {
        unsigned long long c = ((unsigned long long)acc[i + 1] << 32) | acc[i], x = ((unsigned long long)a[(i + 1) & 7] << 32) | a[i & 7];
        asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(c) : "l"(x));
        acc[i] = (uint32_t)c; acc[i + 1] = (uint32_t)(c >> 32);
      }
    } else if (V == V_F2_SUB_ABSADD) {
#pragma unroll
      for (int i = 0; i < NACC; i += 2) {
        const int o = (i + 16) & (NACC - 1);
        float2 x = make_float2(__uint_as_float(a[i & 7]), __uint_as_float(a[(i + 1) & 7]));
        float2 y = make_float2(__uint_as_float(acc[o]), __uint_as_float(acc[o + 1]));   // another chain's accumulators: nothing hoists
        float2 c = make_float2(__uint_as_float(acc[i]), __uint_as_float(acc[i + 1]));
        float2 d = __fadd2_rn(x, make_float2(-y.x, -y.y));
        c = __fadd2_rn(c, make_float2(fabsf(d.x), fabsf(d.y)));
        acc[i] = __float_as_uint(c.x); acc[i + 1] = __float_as_uint(c.y);
      }
    } else if (V == V_F2_SUB_MUL_ADD) {
#pragma unroll
      for (int i = 0; i < NACC; i += 2) {
        const int o = (i + 16) & (NACC - 1);
        float2 x = make_float2(__uint_as_float(a[i & 7]), __uint_as_float(a[(i + 1) & 7]));
        float2 y = make_float2(__uint_as_float(acc[o]), __uint_as_float(acc[o + 1]));
        float2 c = make_float2(__uint_as_float(acc[i]), __uint_as_float(acc[i + 1]));
        // the form K1f uses: ptxas contracts a packed multiply + packed add into FFMA2 even with .rn, which would not be
        // bit-exact; the square is two scalar FMULs between the packed subtraction and the packed accumulation
        float2 d = __fadd2_rn(x, make_float2(-y.x, -y.y));
        asm volatile("mul.rn.f32 %0, %0, %0;" : "+f"(d.x));
        asm volatile("mul.rn.f32 %0, %0, %0;" : "+f"(d.y));
        c = __fadd2_rn(c, d);
        acc[i] = __float_as_uint(c.x); acc[i + 1] = __float_as_uint(c.y);
      }
    } else if (V == V_DFMA) {
#pragma unroll
      for (int i = 0; i < NACC; i += 2) {
        double c = __hiloint2double((int)acc[i + 1], (int)acc[i]);
        const double x = __hiloint2double((int)(a[i & 7] & 0x000FFFFFu) | 0x3FF00000, (int)a[(i + 1) & 7]);
        asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(c) : "d"(x));
        acc[i] = (uint32_t)__double2loint(c); acc[i + 1] = (uint32_t)__double2hiint(c);
      }
    } else if (V == V_DMUL_DADD) {
#pragma unroll
      for (int i = 0; i < NACC; i += 2) {
        double c = __hiloint2double((int)acc[i + 1], (int)acc[i]);
        const double x = __hiloint2double((int)(a[i & 7] & 0x000FFFFFu) | 0x3FF00000, (int)a[(i + 1) & 7]);
        asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(c) : "d"(x));
        asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(c) : "d"(x));
        acc[i] = (uint32_t)__double2loint(c); acc[i + 1] = (uint32_t)__double2hiint(c);
      }
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s ^= acc[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i];
  out[blockIdx.x * blockDim.x + t] = s;
  if (t == 0) cycles[blockIdx.x] = t1 - t0;
}

struct Desc { const char* name; double ops_per_iter; double lanes_per_op; };

const Desc kDesc[V_COUNT] = {
  {"sad4_acc(VABSDIFF4.U8.ACC)", NACC, 4},
  {"sad4+dp4a(VABSDIFF4+IDP.4A)", NACC, 4},      // op = one SAD4 equivalent (2 instr)
  {"half2 sub+|add|(HFMA2+HADD2)", NACC, 2},    // op = one half2 pair (2 instr)
  {"IADD3", NACC, 0}, {"LOP3", NACC, 0}, {"IMAD", NACC, 0}, {"FFMA", NACC, 0},
  {"VIMNMX", NACC, 0}, {"VIMNMX3", NACC, 0}, {"IDP.4A", NACC, 4},
  {"sad4_acc + 1/4 IMAD", NACC, 4}, {"sad4_acc + 1/1 IMAD", NACC, 4},
  {"sad4 8x4 tile, LDS.128 operands", 128, 4},
  {"VABSDIFF.U32 scalar sad", NACC, 1},
  {"FADD", NACC, 0},
  {"f32 sub + |.|add (2 FADD)", NACC, 1},          // op = one pair-column of K1f at p = 1
  {"f32 sub + mul + add (FADD FMUL FADD)", NACC, 1},
  {"FADD2 (add.rn.f32x2)", NACC / 2, 2},                        // op = one instruction = two float adds
  {"f32x2 sub + |.|add (2 FADD2 per 2)", NACC, 1},               // op = one p = 1 pair-column
  {"f32x2 sub + mul + add (FADD2 2xFMUL FADD2 per 2)", NACC, 1},  // op = one p = 2 pair-column
  {"DFMA", NACC / 2, 1},
  {"f64 mul + add (DMUL DADD)", NACC / 2, 1},                    // op = one multiply-add pair, not contracted
};

template <int V>
int run_variant(int sms, int ctas_per_sm, int iters, const uint32_t* din, uint32_t* dout, long long* dcyc,
                double* ops_per_clk_sm, double* gops) {
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, peak_kernel<V>, 256, 0));
  if (occ < ctas_per_sm) ctas_per_sm = occ;   // keep the whole grid resident: one wave
  const int grid = sms * ctas_per_sm;
  peak_kernel<V><<<grid, 256>>>(din, dout, 16, dcyc);  // warm-up
  CK(cudaGetLastError());
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaEventRecord(e0));
    peak_kernel<V><<<grid, 256>>>(din, dout, iters, dcyc);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  static long long hc[148 * 16];
  CK(cudaMemcpy(hc, dcyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost));
  double cyc = 0; for (int i = 0; i < grid; ++i) cyc += (double)hc[i]; cyc /= grid;
  const double ops_cta = (double)iters * kDesc[V].ops_per_iter * 256.0;
  *ops_per_clk_sm = ops_cta * ctas_per_sm / cyc;
  *gops = ops_cta * grid / (best * 1e-3) / 1e9;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return 0;
}

}  // namespace

// Runs every variant; fills out[v*2+0] = ops/clk/SM, out[v*2+1] = G ops/s.
// names (optional) receives V_COUNT pointers to static strings.
// Returns the number of variants, or -1 on a CUDA error.
extern "C" int spv_pipe_peak(int device, int iters, double* out, const char** names, double* lanes_per_op) {
  CK(cudaSetDevice(device));
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device));
  const int sms = prop.multiProcessorCount;
  if (sms > 148) { fprintf(stderr, "pipe_peak: more SMs than expected\n"); return -1; }
  const int cps = 2;  // 2 CTAs x 256 threads = 16 warps / SM, 32 independent chains each
  uint32_t* din; uint32_t* dout; long long* dcyc;
  CK(cudaMalloc(&din, 1024 * 4)); CK(cudaMalloc(&dout, (size_t)sms * cps * 256 * 4)); CK(cudaMalloc(&dcyc, sizeof(long long) * sms * cps));
  uint32_t h[1024]; for (int i = 0; i < 1024; ++i) h[i] = 0x9E3779B9u * (uint32_t)(i + 1);
  CK(cudaMemcpy(din, h, sizeof(h), cudaMemcpyHostToDevice));
  int rc = 0;
#define RUN(V) if (!rc) rc = run_variant<V>(sms, cps, iters, din, dout, dcyc, &out[V * 2], &out[V * 2 + 1]);
  RUN(V_SAD4_ACC) RUN(V_SAD4_DP4A) RUN(V_HALF2_ABS) RUN(V_IADD3) RUN(V_LOP3) RUN(V_IMAD) RUN(V_FFMA)
  RUN(V_VIMNMX) RUN(V_VIMNMX3) RUN(V_DP4A) RUN(V_SAD4_ACC_IMAD) RUN(V_SAD4_ACC_IMAD11) RUN(V_SAD4_TILE) RUN(V_SAD_SCALAR)
  RUN(V_FADD) RUN(V_FSUB_ABSADD) RUN(V_FSUB_FMUL_FADD) RUN(V_FADD2) RUN(V_F2_SUB_ABSADD) RUN(V_F2_SUB_MUL_ADD)
  RUN(V_DFMA) RUN(V_DMUL_DADD)
#undef RUN
  for (int v = 0; v < V_COUNT; ++v) { if (names) names[v] = kDesc[v].name; if (lanes_per_op) lanes_per_op[v] = kDesc[v].lanes_per_op; }
  cudaFree(din); cudaFree(dout); cudaFree(dcyc);
  return rc ? -1 : (int)V_COUNT;
}

#ifdef PIPE_PEAK_MAIN
int main(int argc, char** argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 20000;
  double out[V_COUNT * 2]; const char* names[V_COUNT]; double lanes[V_COUNT];
  int n = spv_pipe_peak(0, iters, out, names, lanes);
  if (n < 0) return 1;
  printf("%-36s %14s %12s %16s\n", "variant", "ops/clk/SM", "G ops/s", "byte-lanes/clk/SM");
  for (int v = 0; v < n; ++v)
    printf("%-36s %14.2f %12.1f %16.1f\n", names[v], out[v * 2], out[v * 2 + 1], out[v * 2] * lanes[v]);
  return 0;
}
#endif
