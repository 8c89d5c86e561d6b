// ransac.cu -- K8: RANSAC hypothesis scoring on sm_100a.  Replaces, for MANY candidate fundamental
// matrices at once, RansacFitter::process_fundamental_matrix (reference src/RansacFitter.h:42-98):
//   per hypothesis F:  SVD(F), the singular-value-ratio gate (:47-53), E = U diag(1,1,0) V^T (:54-56),
//                      the four cameras of Essential2Cameras (src/Camera.h:32-59);
//   per (hypothesis, camera, match):  DltTriangulator::solve (src/DltTriangulator.h:36-70: the 4 x 4 DLT
//                      system, its right singular vector of the smallest singular value, reprojection),
//                      reprojection_error() <= allowed (:72-79) and is_infront_both_cameras() (:81-91);
//   per hypothesis:    the camera with the most inliers under the reference's acceptance rule (:74-80).
// This is the step that consumes the matcher's output (example/ex01_essential_estimation.py:134-152: up
// to 10^7 tries x 4 cameras x all matches, each a 4 x 4 SVD) and it is embarrassingly parallel FP64 work.
//
// Arithmetic.  Every singular value decomposition is the two-sided Jacobi algorithm of Eigen's JacobiSVD
// for real square matrices (scaling by the largest |entry|, sweeps over p = 1..n-1, q = 0..p-1,
// real_2x2_jacobi_svd + makeJacobi, threshold 2 eps max|diag|, signs folded into U, selection sort
// descending), restated operation for operation as in oracle/eigen_shim_mvg/Eigen/Core, through which the
// reference's own headers are compiled into the CPU checker.  All other operations (the DLT rows, the
// products, hnormalized divisions, the two square roots of the reprojection error, the cheirality
// quotients) are performed in the reference's order.  The file is compiled with -fmad=false: the
// reference is built without -march (no FMA contraction), so every multiply and add rounds separately
// here too and the results are BIT-IDENTICAL to the checker's.  Bound: the FP64 pipe.
#include "spv_common.h"

#include <cfloat>
#include <cstdint>

namespace spv {
namespace {

struct Rot { double c, s; };

__device__ __forceinline__ Rot make_jacobi(double x, double y, double z) {
  Rot r;
  const double deno = 2.0 * fabs(y);
  if (deno < DBL_MIN) { r.c = 1.0; r.s = 0.0; return r; }
  const double tau = (x - z) / deno;
  const double w = sqrt(tau * tau + 1.0);
  const double t = tau > 0 ? 1.0 / (tau + w) : 1.0 / (tau - w);
  const double sign_t = t > 0 ? 1.0 : -1.0;
  const double n = 1.0 / sqrt(t * t + 1.0);
  // the reference divides y / |y|: +-1 exactly for every finite non-zero y (|y| >= DBL_MIN / 2 here), and a NaN y has
  // already made t NaN, so copysign gives the same product without a division
  r.s = -sign_t * copysign(1.0, y) * fabs(t) * n;
  r.c = n;
  return r;
}

// real_2x2_jacobi_svd: the left and right rotations that diagonalise the 2 x 2 block (m00 m01; m10 m11).
// NOT inlined on purpose: the divisions and square roots in here expand to ~300 SASS instructions, and one copy per
// (p, q) pair made the scoring kernel 3 800 instructions (61 KB) -- ncu r02 showed the warps waiting on instruction
// fetch (stall no_instruction 3.5 per issue).  One shared copy keeps the sweep inside the instruction cache; the
// arithmetic is the same instruction sequence either way.
__device__ __noinline__ void rotations_2x2(double m00, double m01, double m10, double m11, Rot& jl, Rot& jr) {
  Rot rot1;
  const double t = m00 + m11, d = m10 - m01;
  if (fabs(d) < DBL_MIN) { rot1.s = 0.0; rot1.c = 1.0; }
  else {
    const double uu = t / d;
    const double tmp = sqrt(1.0 + uu * uu);
    rot1.s = 1.0 / tmp;
    rot1.c = uu / tmp;
  }
  const double a00 = rot1.c * m00 + rot1.s * m10, a01 = rot1.c * m01 + rot1.s * m11;
  const double a11 = -rot1.s * m01 + rot1.c * m11;
  jr = make_jacobi(a00, a01, a11);
  const double c2 = jr.c, s2 = -jr.s;
  jl.c = rot1.c * c2 - rot1.s * s2;
  jl.s = rot1.c * s2 + rot1.s * c2;
}

// Two-sided Jacobi SVD of an N x N matrix held in registers (row-major a; destroyed).  WANT_U: also U.
// Outputs: sv descending, u / v with their columns permuted along.  All loops over p, q, k are unrolled
// at compile time so that a, u, v stay in registers.
template <int N, bool WANT_U>
__device__ __forceinline__ void jacobi_svd(double (&a)[N * N], double (&sv)[N], double (&u)[N * N], double (&v)[N * N]) {
  double scale = 0.0;
#pragma unroll
  for (int i = 0; i < N * N; ++i) scale = fmax(scale, fabs(a[i]));
  if (scale == 0.0) scale = 1.0;
#pragma unroll
  for (int i = 0; i < N * N; ++i) a[i] = a[i] / scale;
#pragma unroll
  for (int i = 0; i < N * N; ++i) {
    v[i] = (i / N == i % N) ? 1.0 : 0.0;
    if (WANT_U) u[i] = v[i];
  }
  const double precision = 2.0 * DBL_EPSILON;
  double max_diag = 0.0;
#pragma unroll
  for (int i = 0; i < N; ++i) max_diag = fmax(max_diag, fabs(a[i * N + i]));
  bool finished = false;
  int guard = 0;
  while (!finished && ++guard < 64) {   // converges in 4-7 sweeps; the guard only bounds a NaN input
    finished = true;
#pragma unroll
    for (int p = 1; p < N; ++p)
#pragma unroll
      for (int q = 0; q < p; ++q) {
        const double threshold = fmax(DBL_MIN, precision * max_diag);
        if (fabs(a[p * N + q]) > threshold || fabs(a[q * N + p]) > threshold) {
          finished = false;
          Rot jl, jr;   // real_2x2_jacobi_svd on the (p, q) block
          rotations_2x2(a[p * N + p], a[p * N + q], a[q * N + p], a[q * N + q], jl, jr);
          // work.applyOnTheLeft(p, q, j_left): rows p, q
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const double x = a[p * N + k], y = a[q * N + k];
            a[p * N + k] = jl.c * x + jl.s * y;
            a[q * N + k] = -jl.s * x + jl.c * y;
          }
          if (WANT_U) {   // U.applyOnTheRight(p, q, j_left.transpose()): columns p, q by (c, -s)^T = (c, s)
#pragma unroll
            for (int k = 0; k < N; ++k) {
              const double x = u[k * N + p], y = u[k * N + q];
              u[k * N + p] = jl.c * x - (-jl.s) * y;
              u[k * N + q] = (-jl.s) * x + jl.c * y;
            }
          }
          // work.applyOnTheRight(p, q, j_right), V.applyOnTheRight(p, q, j_right)
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const double x = a[k * N + p], y = a[k * N + q];
            a[k * N + p] = jr.c * x - jr.s * y;
            a[k * N + q] = jr.s * x + jr.c * y;
          }
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const double x = v[k * N + p], y = v[k * N + q];
            v[k * N + p] = jr.c * x - jr.s * y;
            v[k * N + q] = jr.s * x + jr.c * y;
          }
          max_diag = fmax(max_diag, fmax(fabs(a[p * N + p]), fabs(a[q * N + q])));
        }
      }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const double d = a[i * N + i];
    sv[i] = fabs(d);
    if (WANT_U && d < 0) {
#pragma unroll
      for (int k = 0; k < N; ++k) u[k * N + i] = -u[k * N + i];
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) sv[i] = sv[i] * scale;
  // selection sort, descending (first maximum wins), columns of U and V swapped along
#pragma unroll
  for (int i = 0; i < N - 1; ++i) {
    int best = i;
    double bestv = sv[i];
#pragma unroll
    for (int k = i + 1; k < N; ++k) if (sv[k] > bestv) { best = k; bestv = sv[k]; }
#pragma unroll
    for (int k = i + 1; k < N; ++k)
      if (best == k) {
        const double ts = sv[i]; sv[i] = sv[k]; sv[k] = ts;
#pragma unroll
        for (int r = 0; r < N; ++r) {
          const double tv = v[r * N + i]; v[r * N + i] = v[r * N + k]; v[r * N + k] = tv;
          if (WANT_U) { const double tu = u[r * N + i]; u[r * N + i] = u[r * N + k]; u[r * N + k] = tu; }
        }
      }
  }
}

// out = a * b, 3 x 3, out(i, j) = ((a(i,0) b(0,j) + a(i,1) b(1,j)) + a(i,2) b(2,j))   (the checker's product order)
__device__ __forceinline__ void mul3(const double (&a)[9], const double (&b)[9], double (&o)[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double acc = a[i * 3] * b[j];
      acc = acc + a[i * 3 + 1] * b[3 + j];
      acc = acc + a[i * 3 + 2] * b[6 + j];
      o[i * 3 + j] = acc;
    }
}
__device__ __forceinline__ void transpose3(const double (&a)[9], double (&o)[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) o[j * 3 + i] = a[i * 3 + j];
}

// per hypothesis: gate + the four candidate cameras.  cams[h][c][12] row-major 3 x 4, aux[h][c] = {sign det M, |m|^2}
// (DltTriangulator.h:28-34), valid[h] = 0 when the singular-value-ratio gate rejects F (RansacFitter.h:50-53).
// Gated-in hypotheses are appended to vlist (order irrelevant) so that the scoring kernel only ever sees real work:
// with seven-point draws from matches that hold outliers, most candidates fail the gate (bench: 157 of 4 001).
__global__ void ransac_setup_kernel(const double* __restrict__ F, int nF, double sv_ratio_allowed, double* __restrict__ cams,
                                    double* __restrict__ aux, int* __restrict__ valid, int* __restrict__ counts,
                                    int* __restrict__ vlist, int* __restrict__ nvalid) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= nF) return;
#pragma unroll
  for (int c = 0; c < 4; ++c) counts[h * 4 + c] = 0;
  double a[9], sv[3], u[9], v[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) a[i] = F[(size_t)9 * h + i];
  jacobi_svd<3, true>(a, sv, u, v);
  const double ratio = fabs(sv[0] - sv[1]) / (fabs(sv[0] + sv[1]) / 2.);
  if (ratio > sv_ratio_allowed) { valid[h] = 0; return; }   // also false for NaN, like the reference's comparison
  valid[h] = 1;
  vlist[atomicAdd(nvalid, 1)] = h;
  // E = U * diag(1, 1, 0) * V^T
  double ud[9], vt[9], E[9];
#pragma unroll
  for (int i = 0; i < 3; ++i) { ud[i * 3] = u[i * 3] * 1.0; ud[i * 3 + 1] = u[i * 3 + 1] * 1.0; ud[i * 3 + 2] = u[i * 3 + 2] * 0.0; }
  transpose3(v, vt);
  mul3(ud, vt, E);
  // Essential2Cameras (Camera.h:32-59)
  jacobi_svd<3, true>(E, sv, u, v);
  const double D[9] = {0., 1., 0., -1., 0., 0., 0., 0., 1.}, Dt[9] = {0., -1., 0., 1., 0., 0., 0., 0., 1.};
  double t3[3] = {u[2], u[5], u[8]};
  double tmp[9], Ra[9], Rb[9];
  transpose3(v, vt);
  mul3(u, D, tmp);
  mul3(tmp, vt, Ra);
  mul3(u, Dt, tmp);
  mul3(tmp, vt, Rb);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const double* R = c < 2 ? Ra : Rb;
    double* P = cams + ((size_t)h * 4 + c) * 12;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      P[i * 4] = R[i * 3]; P[i * 4 + 1] = R[i * 3 + 1]; P[i * 4 + 2] = R[i * 3 + 2];
      P[i * 4 + 3] = (c & 1) ? -t3[i] : t3[i];
    }
    const double det = R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
    aux[((size_t)h * 4 + c) * 2] = det < 0 ? -1. : 1.;
    aux[((size_t)h * 4 + c) * 2 + 1] = (R[2] * R[2] + R[5] * R[5]) + R[8] * R[8];   // column 2 of M, squared, summed in order
  }
}

constexpr int RS_THREADS = 128;
constexpr int RS_CHUNK = 4 * RS_THREADS;   // matches per work item

// Persistent blocks over work items (gated-in hypothesis, camera, chunk of 512 matches): every thread takes 4 matches
// of the chunk.  counts[h][c] (zeroed by the setup kernel) collects the chunk counts; masks[(h * 4 + c) * W + word]
// (bit = match & 31) are written whole words at a time, one warp per word.
// Blocks per SM: the kernel waits on FP64 latencies (ncu r02: stall wait 3.2 per issue at 16 warps / SM), so a fifth
// block (102 registers, 124 B of spills) is worth more than spill-free code: 1.18 ms -> 1.15 ms; six blocks spill 384 B.
constexpr int RS_BLOCKS = 5;
__global__ void __launch_bounds__(RS_THREADS, RS_BLOCKS) ransac_score_kernel(const double* __restrict__ cams, const double* __restrict__ aux,
                                                                  const int* __restrict__ vlist, const int* __restrict__ nvalid,
                                                                  const double* __restrict__ x0, const double* __restrict__ x1,
                                                                  int npt, double err_allowed, int* __restrict__ counts,
                                                                  uint32_t* __restrict__ masks, int W) {
  const int nchunk = (npt + RS_CHUNK - 1) / RS_CHUNK;
  const long long total = (long long)(*nvalid) * 4 * nchunk;
  for (long long item = blockIdx.x; item < total; item += gridDim.x) {
  const int chunk = (int)(item % nchunk);
  const int hc = vlist[item / nchunk / 4] * 4 + (int)((item / nchunk) % 4);
  double P[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) P[i] = cams[(size_t)hc * 12 + i];
  const double sign1 = aux[(size_t)hc * 2], norm1 = aux[(size_t)hc * 2 + 1];
  int mine = 0;
  const int cend = min(npt, (chunk + 1) * RS_CHUNK);
  for (int base = chunk * RS_CHUNK; base < cend; base += RS_THREADS) {
    const int ix = base + threadIdx.x;
    bool inlier = false;
    if (ix < npt) {
      // DltTriangulator::solve with P0 = [I | 0] (RansacFitter.h:57: Camera_t P0 is the identity camera)
      const double x = x0[3 * ix] / x0[3 * ix + 2], y = x0[3 * ix + 1] / x0[3 * ix + 2];      // hnormalized divides
      const double xp = x1[3 * ix] / x1[3 * ix + 2], yp = x1[3 * ix + 1] / x1[3 * ix + 2];
      double A[16], sv[4], uu[16], V[16];
      // A.row(0) = x * P0.row(2) - P0.row(0), P0 rows: (1 0 0 0), (0 1 0 0), (0 0 1 0)
      A[0] = x * 0.0 - 1.0; A[1] = x * 0.0 - 0.0; A[2] = x * 1.0 - 0.0; A[3] = x * 0.0 - 0.0;
      A[4] = y * 0.0 - 0.0; A[5] = y * 0.0 - 1.0; A[6] = y * 1.0 - 0.0; A[7] = y * 0.0 - 0.0;
#pragma unroll
      for (int j = 0; j < 4; ++j) { A[8 + j] = xp * P[8 + j] - P[j]; A[12 + j] = yp * P[8 + j] - P[4 + j]; }
      jacobi_svd<4, false>(A, sv, uu, V);
      const double X0 = V[3], X1 = V[7], X2 = V[11], X3 = V[15];   // V.col(3)
      // reprojections: rp0 = P0 X (k order, P0 entries are 0 / 1), rp1 = P1 X
      double r0[3], r1[3];
      r0[0] = ((1.0 * X0 + 0.0 * X1) + 0.0 * X2) + 0.0 * X3;
      r0[1] = ((0.0 * X0 + 1.0 * X1) + 0.0 * X2) + 0.0 * X3;
      r0[2] = ((0.0 * X0 + 0.0 * X1) + 1.0 * X2) + 0.0 * X3;
#pragma unroll
      for (int i = 0; i < 3; ++i) r1[i] = ((P[i * 4] * X0 + P[i * 4 + 1] * X1) + P[i * 4 + 2] * X2) + P[i * 4 + 3] * X3;
      const double e0x = r0[0] / r0[2] - x, e0y = r0[1] / r0[2] - y;
      const double e1x = r1[0] / r1[2] - xp, e1y = r1[1] / r1[2] - yp;
      const double err = sqrt(e0x * e0x + e0y * e0y) + sqrt(e1x * e1x + e1y * e1y);
      // distance2camera: sign det M / |m|^2 * rp(2) / X(3); camera 0: sign 1, |m|^2 = 1
      const double d0 = 1.0 / 1.0 * r0[2] / X3, d1 = sign1 / norm1 * r1[2] / X3;
      inlier = (err <= err_allowed) && (d0 > 0) && (d1 > 0);
    }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, inlier);
    if ((threadIdx.x & 31) == 0) {
      const int word = (base + threadIdx.x) >> 5;
      if (word < W) masks[(size_t)hc * W + word] = bal;
      mine += __popc(bal);
    }
  }
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(counts + hc, mine);
  }
}

// per hypothesis: the reference's camera choice (RansacFitter.h:59-80): cameras in order, a camera wins when
// (percent >= required || find_best) and percent > best so far (strict: the first of equals stays)
__global__ void ransac_pick_kernel(const int* __restrict__ valid, const int* __restrict__ counts, const double* __restrict__ cams,
                                   const uint32_t* __restrict__ masks, int W, int nF, int npt, double required, int find_best,
                                   int* __restrict__ ok, int* __restrict__ ninlier, double* __restrict__ out_cams,
                                   uint32_t* __restrict__ out_masks) {
  const int h = blockIdx.x;
  if (h >= nF) return;
  __shared__ int s_best;
  if (threadIdx.x == 0) {
    int best = -1;
    double best_pct = 0;
    if (valid[h]) {
      for (int c = 0; c < 4; ++c) {
        const double pct = counts[h * 4 + c] / (double)npt;
        if ((pct >= required || find_best) && pct > best_pct) { best_pct = pct; best = c; }
      }
    }
    s_best = best;
    ok[h] = best >= 0;
    ninlier[h] = best >= 0 ? counts[h * 4 + best] : 0;
  }
  __syncthreads();
  const int best = s_best;
  for (int i = threadIdx.x; i < 12; i += blockDim.x) out_cams[(size_t)h * 12 + i] = best >= 0 ? cams[((size_t)h * 4 + best) * 12 + i] : 0.0;
  if (out_masks)
    for (int w = threadIdx.x; w < W; w += blockDim.x) out_masks[(size_t)h * W + w] = best >= 0 ? masks[((size_t)h * 4 + best) * W + w] : 0u;
}

size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

size_t ransac_ws_bytes(int nF, int npt) {
  if (nF < 0 || npt < 0) return 0;
  const size_t W = ((size_t)npt + 31) / 32;
  return al256((size_t)nF * 4 * 12 * 8) + al256((size_t)nF * 4 * 2 * 8) + al256((size_t)nF * 4) + al256((size_t)nF * 4 * 4) +
         al256((size_t)nF * 4 + 16) + al256((size_t)nF * 4 * W * 4) + 256;
}

cudaError_t ransac_score(const double* dF, int nF, const double* dx0, const double* dx1, int npt, double required_percent_inliers,
                         double reprojection_error_allowed, bool find_best_even_in_failure, double singular_value_ratio_allowed,
                         int32_t* dok, int32_t* dcount, double* dcams, uint32_t* dmasks, void* ws, size_t ws_bytes,
                         cudaStream_t stream) {
  if (nF <= 0) return cudaSuccess;
  if (npt <= 0 || ws_bytes < ransac_ws_bytes(nF, npt)) return cudaErrorInvalidValue;
  const int W = (npt + 31) / 32;
  unsigned char* w = static_cast<unsigned char*>(ws);
  double* cams = reinterpret_cast<double*>(w);            w += al256((size_t)nF * 4 * 12 * 8);
  double* aux = reinterpret_cast<double*>(w);             w += al256((size_t)nF * 4 * 2 * 8);
  int* valid = reinterpret_cast<int*>(w);                 w += al256((size_t)nF * 4);
  int* counts = reinterpret_cast<int*>(w);                w += al256((size_t)nF * 4 * 4);
  int* vlist = reinterpret_cast<int*>(w);                 w += al256((size_t)nF * 4 + 16);
  int* nvalid = vlist + nF;
  uint32_t* masks = reinterpret_cast<uint32_t*>(w);
  cudaError_t e = cudaMemsetAsync(nvalid, 0, sizeof(int), stream);
  if (e != cudaSuccess) return e;
  int dev = 0, sms = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  ransac_setup_kernel<<<(nF + 63) / 64, 64, 0, stream>>>(dF, nF, singular_value_ratio_allowed, cams, aux, valid, counts, vlist,
                                                         nvalid);
  // RS_BLOCKS blocks of 128 threads per SM; the grid is that, capped by the work there can be at most
  const long long most = (long long)nF * 4 * ((npt + RS_CHUNK - 1) / RS_CHUNK);
  const int grid = (int)(most < (long long)sms * RS_BLOCKS ? most : (long long)sms * RS_BLOCKS);
  ransac_score_kernel<<<grid, RS_THREADS, 0, stream>>>(cams, aux, vlist, nvalid, dx0, dx1, npt, reprojection_error_allowed,
                                                       counts, masks, W);
  ransac_pick_kernel<<<nF, 32, 0, stream>>>(valid, counts, cams, masks, W, nF, npt, required_percent_inliers,
                                            find_best_even_in_failure ? 1 : 0, dok, dcount, dcams, dmasks);
  count_launch(3);
  return cudaGetLastError();
}

}  // namespace spv
