// spv_common.h -- shared declarations of the device-side matcher (internal).
//
// Result representation used by every kernel of the path: one 64-bit KEY per
// (query row, rank):   key = (uint64)dist << 32 | global_train_index.
// Unsigned comparison of keys IS the reference's selection rule: the scan of
// src/BruteForceNnL1K2.h:109-142 walks x ascending and updates on strict '<',
// so its answer is the two lexicographically smallest (dist, idx) pairs.  Any
// tiling, any merge order and any number of GPUs therefore give the same bits.
// "no neighbour" (BruteForceNnL1K2.h:99-103: dist INT_MAX, idx (size_t)-1) is
// the all-ones key, which loses every comparison.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace spv {

constexpr unsigned long long KEY_NONE = 0xFFFFFFFFFFFFFFFFull;

// keys layout on device: keys[rank * yrows + yrow], rank in {0,1}
struct KeyBuf {
  unsigned long long* keys;
  int yrows;
};

// ---- l1_top2.cu : K1, brute-force L1 k=2 over uint8 rows -------------------
// Accumulates the top-2 of y against x[0..xrows) (global index = xoffset + row)
// INTO keys (which must have been reset or hold earlier partial results).
// kstride: rows of the whole key buffer when dy / keys address a SLICE of the query rows (rank 1 of row r
// sits at keys[kstride + r]); 0 = yrows.  Rows wider than L1K2_TILED_MAX_DIM take the plain wide-row kernel.
// The thresholds of a launch are seeded from what keys already holds, so folding a train set chunk by chunk
// costs no more slow-path work than one launch; folding the same train rows twice is harmless.
constexpr int L1K2_TILED_MAX_DIM = 512;
cudaError_t l1_top2_u8(const uint8_t* dx, const uint8_t* dy, int xrows, int yrows, int dim,
                       uint32_t xoffset, unsigned long long* keys, cudaStream_t stream, long long kstride = 0);
cudaError_t keys_reset(unsigned long long* keys, int yrows, cudaStream_t stream);
// K5: keys_dst = top2(keys_dst U src[0] U ... U src[nsrc-1]); src s at src + s*2*yrows
cudaError_t keys_merge(unsigned long long* dst, const unsigned long long* src, int nsrc, int yrows,
                       cudaStream_t stream);
// keys -> reference output arrays (row-major yrows x 2): size_t idx, int32 dist
// (kstride as in l1_top2_u8: keys / idx / dist may address a slice of the query rows of a larger key buffer)
cudaError_t keys_finalize_i32(const unsigned long long* keys, int yrows, int kcols, uint64_t* idx,
                              int32_t* dist, cudaStream_t stream, long long kstride = 0);
// same but distances as float (CascadingHashNn.h:244 casts int -> float)
cudaError_t keys_finalize_f32(const unsigned long long* keys, int yrows, int kcols, uint64_t* idx,
                              float* dist, cudaStream_t stream);
// TMA view of a row-major rows x dim uint8 matrix as (16 B, rows, dim/16) with byte strides (1, dim, 16):
// a box of (16, box_rows, dim/16) lands in shared memory as [dim/16][box_rows][16 B], i.e. already in the
// k-chunk-major order K1 reads and in the UMMA K-major no-swizzle core-matrix order K2 feeds to tcgen05
bool make_rows_map(CUtensorMap* map, const uint8_t* base, int rows, int dim, int box_rows);
bool make_rows_map_sw128(CUtensorMap* map, const uint8_t* base, int rows, int dim, int box_rows);
bool make_f32_panel_map(CUtensorMap* map, const float* base, int rows, int dim, size_t pitch, int box_rows, int box_cols);
// ---- lp_topk.cu : K1f, brute-force sum |x-y|^p top-k over float32 / int32 rows
// (BruteForceNn.h:37-120).  partial: scratch of lp_partial_bytes(); didx/ddist:
// yrows x k outputs in the reference's layout (ascending, right-aligned).
constexpr int LP_MAX_K = 128;
size_t lp_partial_bytes(int xrows, int yrows, int k, int sms);     // scratch of the register-staged kernel (any dtype, any k)
// scratch that also admits the TMA-fed kernel (lp_topk_tma.cu: float32 rows, k <= LP_MAX_K): >= lp_partial_bytes
size_t lp_ws_bytes(int xrows, int yrows, int dim, int k, int sms);
// ws_bytes decides the kernel: with lp_ws_bytes() of scratch float32 rows take the TMA-fed kernel (unless the
// environment says SPECTAVI_K1F=v1), otherwise -- and for int32 rows, k > LP_MAX_K -- the register-staged one
cudaError_t lp_topk(const void* dx, const void* dy, bool is_int, int xrows, int yrows, int dim, int k, double p,
                    void* ws, size_t ws_bytes, uint64_t* didx, void* ddist, cudaStream_t stream);
// mu != 0 (BruteForceNn.h:80-86, the reference's approximate prune): a sequential-per-query emulation that returns
// the reference's result, not the exact top-k
size_t lp_mu_ws_bytes(int yrows, int k);
cudaError_t lp_topk_mu(const void* dx, const void* dy, bool is_int, int xrows, int yrows, int dim, int k, double p, float mu,
                       void* ws, size_t ws_bytes, uint64_t* didx, void* ddist, cudaStream_t stream);
// lp_topk_tma.cu
size_t lp_tma_ws_bytes(int xrows, int yrows, int dim, int k, int sms);
cudaError_t lp_topk_tma(const float* dx, const float* dy, int xrows, int yrows, int dim, int k, double p, void* ws,
                        size_t ws_bytes, int sms, unsigned long long** partial_out, int* splits_out, cudaStream_t stream);
// The filter overload (BruteForceNn.h:37-49), batched: tile t = {ypos, ycnt <= lp_filtered_tile_rows(), xpos, xcnt}
// searches query rows ysel[ypos .. ypos+ycnt) among train rows xsel[xpos .. xpos+xcnt).  A query row id may appear
// in ysel once.  didx / ddist (yrows x k) are in-out: only rows named by ysel are written, right-aligned.
int lp_filtered_tile_rows();
size_t lp_filtered_partial_bytes(int ntiles, int max_xcnt, int nysel, int k, int sms);
cudaError_t lp_topk_filtered(const float* dx, const float* dy, int dim, int k, double p, const int4* dtiles, int ntiles,
                             const uint32_t* dysel, int nysel, const uint32_t* dxsel, int max_xcnt,
                             unsigned long long* partial, uint64_t* didx, float* ddist, cudaStream_t stream);
// ---- cascade.cu : the cascade-of-hashes matcher (CascadingHashNn.h)
// rows are uint8 (value + 128); ddict: [n][dim][m] fp32 on the device; dict_terms 1
// when every entry is exact in bf16 (the library's own quantised dictionary), else 3.
// Folds the top-2 of the candidates of dy among dx[0..xrows) into keys.
size_t cascade_ws_bytes(int xrows, int yrows, int dim, int m, int n, int g);
cudaError_t f32_to_u8(const float* din, uint8_t* dout, size_t count, cudaStream_t stream);
cudaError_t cascade_match(const uint8_t* dxu8, const uint8_t* dyu8, int xrows, int yrows, int dim, int m, int n, int g,
                          const float* ddict, int dict_terms, uint32_t xoffset, unsigned long long* keys, void* ws,
                          size_t ws_bytes, cudaStream_t stream);
int cascade_stage_ms(double* out, int n);   // SPECTAVI_STAGE_TIMING=1: ms of {image, K2 train, K3, K2 query, K4} of the last call
cudaError_t hash_codes(const uint8_t* drows, int rows, int dim, int m, int n, int g, const float* ddict, int dict_terms,
                       int32_t* codes, unsigned long long* flips, void* ws, size_t ws_bytes, cudaStream_t stream);
// ---- ratio.cu : K6, ratio test + order-preserving compaction of the surviving matches
// (example/ex01_essential_estimation.py:102-105).  matches: [count][2] uint64 (train idx, query row),
// count: one uint32; mask (nullable): yrows bytes; ws: ratio_ws_bytes(yrows) scratch.
size_t ratio_ws_bytes(int yrows);
cudaError_t ratio_matches(const unsigned long long* keys, int yrows, double min_ratio, bool dist_f32, unsigned char* mask,
                          uint64_t* matches, unsigned* count, void* ws, size_t ws_bytes, cudaStream_t stream);
// the same with the scratch in a file-scope device array (no workspace, nothing allocated); yrows <= ratio_static_max_rows()
int ratio_static_max_rows();
cudaError_t ratio_matches_static(const unsigned long long* keys, int yrows, double min_ratio, bool dist_f32,
                                 unsigned char* mask, uint64_t* matches, unsigned* count, cudaStream_t stream);
// ---- quantize.cu : K7, normalize_to_ubyte_and_multiple_16_dim (spectavi/feature.py:384-407) on the device.
// dx: rows x dim float32; outputs rows x normalize_new_dim(dim): float32 (nullable) and / or uint8 = value + 128
// (nullable); ws: normalize_ws_bytes(dim) scratch.
size_t normalize_ws_bytes(int dim);
int normalize_new_dim(int dim);
cudaError_t normalize_rows(const float* dx, int rows, int dim, float* dout_f32, uint8_t* dout_u8, void* ws, size_t ws_bytes,
                           cudaStream_t stream);
// ---- ransac.cu : K8, RANSAC hypothesis scoring (RansacFitter::process_fundamental_matrix, src/RansacFitter.h:42-98,
// for nF candidate fundamental matrices at once).  dF: [nF][9] row-major; dx0 / dx1: [npt][3] homogeneous matches.
// Outputs per hypothesis: ok (the reference's return value), inlier count and camera (3 x 4 row-major) of the
// winning candidate, and (nullable) its inlier bitmask, [nF][ceil(npt / 32)] words, bit = match & 31.
size_t ransac_ws_bytes(int nF, int npt);
cudaError_t ransac_score(const double* dF, int nF, const double* dx0, const double* dx1, int npt, double required_percent_inliers,
                         double reprojection_error_allowed, bool find_best_even_in_failure, double singular_value_ratio_allowed,
                         int32_t* dok, int32_t* dcount, double* dcams, uint32_t* dmasks, void* ws, size_t ws_bytes,
                         cudaStream_t stream);
// launch count bookkeeping (bench.py's gpu_launches)
unsigned long long launches();
void count_launch(int n = 1);

}  // namespace spv
