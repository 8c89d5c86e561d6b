/* spectavi_b200.h -- C ABI of libspectavi.so (B200 build of spectavi's
 * descriptor-matching path).
 *
 * Part 1 is the DROP-IN boundary: the same symbols, argument order and output
 * protocol as the reference's extern "C" functions for this path, so that
 * spectavi/feature.py's ctypes bindings (reference spectavi/feature.py:208-243,
 * 346-358) work unchanged against this library.  All of them take HOST
 * pointers, return void, and -- like the reference, whose failures are C++
 * exceptions escaping extern "C" -- print one line to stderr and abort() on a
 * violated precondition or a CUDA/NCCL error.  There is no CPU fallback.
 *
 * Part 2 is the device-resident API the drop-in functions are built from; it
 * takes DEVICE pointers and a CUDA stream, returns 0 or a negative error code,
 * and is what bench.py times with inputs already in HBM and what the
 * one-process-per-GPU (torchrun) path shards the train set with.
 */
#ifndef SPECTAVI_B200_H
#define SPECTAVI_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

/* ------------------------------------------------------------------ NdArray
 * Byte-for-byte the reference's struct (ctypes_ndarray/src/NdArray.h:17-29;
 * Python mirror ctypes_ndarray/cndarray/ndarray.py:50-58).  Output memory is
 * allocated by, and owned by, the caller through m_alloc. */
typedef char *NumpyAllocator(size_t nbytes);
typedef void SetDtypeFunc(const char *newdtype);

struct NdArray {
  char *m_data;
  int m_ndim;
  uint64_t *m_shape;
  int m_sizeofdtype;
  NumpyAllocator *m_alloc;
  NumpyAllocator *m_realloc;
  SetDtypeFunc *m_set_dtype;
  char *m_dtype;
};
typedef struct NdArray NdArray;

/* ============================ Part 1: drop-in entry points (host pointers)
 * Host arrays are ordinary pageable memory; the library stages them itself.  Environment:
 *   SPECTAVI_GPUS=N|all       nn_bruteforcel1k2 / nn_cascading_hash shard the TRAIN rows over N GPUs (NCCL all-gather
 *                             of the per-GPU top-2 keys); nn_bruteforce / nn_bruteforcei shard the QUERY rows (from
 *                             8192 rows per GPU, no exchange).  Results are bit-identical to one GPU.
 *   SPECTAVI_UPLOAD_LANES=1..8  host threads staging a large pageable array (default 4)
 *   SPECTAVI_CVT_LANES=1..16  host threads converting nn_cascading_hash's float32 rows to uint8 while staging them
 *                             (default: 3/4 of the host's cores); SPECTAVI_CVT=device converts on the GPU (K0) instead
 *   SPECTAVI_SEED=<uint>      seed of the cascade's hyperplane dictionary */

/* replaces src/Spectavi.cpp:284-298 (BruteForceNnL1K2, src/BruteForceNnL1K2.h).
 * x: xrows x dim uint8 train rows, y: yrows x dim uint8 query rows, row-major,
 * dim % 16 == 0.  outidx: yrows x 2 size_t, outdist: yrows x 2 int32: the two
 * lexicographically smallest (dist, idx); missing = ((size_t)-1, INT_MAX).
 * nthreads (an OpenMP thread count in the reference) is accepted and ignored. */
void nn_bruteforcel1k2(const uint8_t *x, const uint8_t *y, int xrows, int yrows, int dim, int nthreads,
                       NdArray *outidx, NdArray *outdist);

/* replaces src/Spectavi.cpp:258-269 (BruteForceNn<float>, src/BruteForceNn.h:37-120).
 * k nearest under sum |d|^p (p == 1, 2, .5 special-cased, otherwise powf);
 * outidx yrows x k size_t, outdist yrows x k float32, ascending.  mu != 0 (the
 * reference's approximate, visiting-order-dependent prune, :80-86) is emulated in the
 * reference's order and returns the reference's result; mu == 0 takes the fast kernels. */
void nn_bruteforce(const float *x, const float *y, int xrows, int yrows, int dim, int k, float p, float mu,
                   NdArray *outidx, NdArray *outdist);

/* replaces src/Spectavi.cpp:271-282 (BruteForceNn<int>): int32 rows, per-term
 * value truncated to int as the reference's Scalar cast does; outdist int32. */
void nn_bruteforcei(const int *x, const int *y, int xrows, int yrows, int dim, int k, float p, float mu,
                    NdArray *outidx, NdArray *outdist);

/* replaces src/Spectavi.cpp:321-336 (CascadingHashNn, src/CascadingHashNn.h).
 * x, y float32 holding integers in [-128,127]; outputs yrows x k (k >= 2; only
 * columns 0,1 are written), outdist float32.  The hash dictionary is drawn
 * from a fixed seed (env SPECTAVI_SEED overrides) instead of random_device. */
void nn_cascading_hash(const float *x, const float *y, int xrows, int yrows, int dim, int k, int hash_bit_rate,
                       int num_hash_tables, int num_candidate_neighbours, NdArray *outidx, NdArray *outdist);

/* ---- Part 1b: the step after the matcher, same calling convention.
 * The reference applies the ratio test in numpy (example/ex01_essential_estimation.py:102-105:
 * ratio = d2 / float64(d1); pass = ratio >= min_ratio; matches = (idx0[pass], arange(yrows)[pass]))
 * on the yrows x 2 arrays the matcher returned.  These entry points run the matcher, the
 * ratio test and the compaction on the GPU and return only the surviving pairs:
 * outmatches count x 2 uint64, column 0 = train index, column 1 = query row, query order kept.
 * Identical to applying the numpy lines above to nn_bruteforcel1k2's / nn_cascading_hash's output. */
void spv_match_bruteforcel1k2(const uint8_t *x, const uint8_t *y, int xrows, int yrows, int dim, double min_ratio,
                              NdArray *outmatches);
void spv_match_cascading_hash(const float *x, const float *y, int xrows, int yrows, int dim, int hash_bit_rate,
                              int num_hash_tables, int num_candidate_neighbours, double min_ratio,
                              NdArray *outmatches);

/* The step before the matcher: spectavi/feature.py:384-407
 * normalize_to_ubyte_and_multiple_16_dim for float32 rows -- column de-mean (float32, rows in order, as
 * numpy sums them), scale by 128 / max|column|, round half to even, clip to [-128, 127], zero-pad the
 * columns to a multiple of 16.  out: rows x ceil16(dim) float32, bit-identical to the numpy function. */
void spv_normalize_to_ubyte(const float *x, int rows, int dim, NdArray *out);

/* ---- Part 1c: the second caller of the generic matcher.
 * spv_nn_bruteforce_filtered replaces BruteForceNn::find_neighbours(out_idx, out_dist, filter_x, filter_y, k)
 * (src/BruteForceNn.h:37-120), which the reference reaches only from KMedians (it has no C entry point of its
 * own): the k nearest rows of x, among those named by filter_x, for every row of y named by filter_y, under
 * sum |d|^p.  The filters are row-id sets (std::unordered_set<size_t> there; lists here, duplicates collapse);
 * an EMPTY filter means all rows (:45-46).  outidx (yrows x k uint64) / outdist (yrows x k float32) are the
 * caller's arrays and are IN-OUT as in the reference: only the rows in filter_y are written, and a row with
 * n < k candidates only gets its last n columns (:111-118).  Distances are bit-exact; candidates that tie in
 * distance are ordered by row id (the reference's order is its hash set's, i.e. unspecified).
 *
 * spv_kmedians_find_nearest_neighbours replaces KMedians::find_nearest_neighbours (src/KMedians.h:260-295), the
 * body of nn_kmedians after its two clusterings (src/Spectavi.cpp:305-319): x is `other` (train), y is `this`
 * (query).  For every query cluster g: its c nearest train medians under L1 (:268-273), then the filtered search
 * (p = 1) of the cluster's rows among the union of those train clusters (:275-293).  All clusters run in one
 * kernel launch.  The clusterings are inputs -- medians (nm x dim float32) and labels (row -> cluster id) --
 * because KMedians::run seeds its shuffle from std::random_device (KMedians.h:36-40) and is out of scope
 * (SURVEY.md 2 row 9).  Outputs are allocated like nn_kmedians' (Spectavi.cpp:310-313). */
void spv_nn_bruteforce_filtered(const float *x, const float *y, int xrows, int yrows, int dim, int k, float p,
                                float mu, const uint64_t *filter_x, int nfx, const uint64_t *filter_y, int nfy,
                                uint64_t *outidx, float *outdist);
void spv_kmedians_find_nearest_neighbours(const float *x, const float *y, int xrows, int yrows, int dim,
                                          const float *medians_x, const int32_t *labels_x, int nmx,
                                          const float *medians_y, const int32_t *labels_y, int nmy, int c, int k,
                                          NdArray *outidx, NdArray *outdist);

/* ---- Part 1d: the step AFTER the matcher: RANSAC hypothesis scoring.
 * The reference's ransac_fitter (src/Spectavi.cpp:70-88 -> RansacFitter::fit_essential, src/RansacFitter.h:154-277)
 * draws 7 matches per try from std::random_device, solves the seven-point system (out of scope here) and scores
 * every resulting fundamental matrix with process_fundamental_matrix (RansacFitter.h:42-98): SVD gate, the four
 * cameras of Essential2Cameras (src/Camera.h:32-59), a DLT triangulation (src/DltTriangulator.h:36-92) of EVERY
 * match against every camera, inliers = reprojection error <= allowed and in front of both cameras.  These entry
 * points take the candidate matrices (F: nF x 9 row-major, from any seven-point solver) and do the scoring on the
 * GPU, all hypotheses x 4 cameras x all matches in one launch.  x0, x1: npt x 3 homogeneous normalised points
 * (npt >= 10, as the reference's constructor demands).
 *
 * spv_ransac_process_fundamental_matrices: per hypothesis what process_fundamental_matrix returns -- ok (int32,
 * its bool), ninlier (int32), cameras (float64 nF x 12: the winning 3 x 4 camera, row-major) and, when masks is not
 * NULL, the winner's inliers as bit masks (uint32 nF x ceil(npt / 32), bit = match & 31).
 * spv_ransac_fit_hypotheses: ransac_fitter's outputs (Spectavi.cpp:81-87) for the hypothesis list taken in order
 * with fit_essential's bookkeeping: the first hypothesis whose inlier percent exceeds required_percent_inliers wins
 * (success = true); otherwise, with find_best_even_in_failure, the best one seen.  essential = the winning F,
 * camera 3 x 4, inlier_idx int32 (ascending).
 * Arithmetic: Eigen's JacobiSVD is restated (two-sided Jacobi); results equal the reference's code
 * compiled against that restatement bit for bit, and a LAPACK-based check up to matches within 1e-9 of the threshold. */
void spv_ransac_process_fundamental_matrices(const double *x0, const double *x1, int npt,
                                             double required_percent_inliers, double reprojection_error_allowed,
                                             int find_best_even_in_failure, double singular_value_ratio_allowed,
                                             const double *F, int nF, NdArray *ok, NdArray *ninlier, NdArray *cameras,
                                             NdArray *masks);
void spv_ransac_fit_hypotheses(const double *x0, const double *x1, int npt, double required_percent_inliers,
                               double reprojection_error_allowed, int find_best_even_in_failure,
                               double singular_value_ratio_allowed, const double *F, int nF, bool *success,
                               NdArray *essential, NdArray *camera, double *inlier_percent, NdArray *inlier_idx);

/* ============================ Part 2: device-resident API (device pointers)
 * Results live in a KEY buffer: uint64 keys[2][yrows], key = dist << 32 | idx,
 * all-ones = none; unsigned order of keys == the reference's (dist, idx) rule. */

/* keys <- all "none" */
int spv_keys_reset(uint64_t *dkeys, int yrows, void *stream);
/* K1: fold the L1 top-2 of dy against dx[0..xrows) into dkeys; reported index
 * = xoffset + local row (train shards of one query set fold into one buffer). */
int spv_l1k2_device(const uint8_t *dx, const uint8_t *dy, int xrows, int yrows, int dim, uint32_t xoffset,
                    uint64_t *dkeys, void *stream);
/* K5: dkeys <- top-2 of dkeys and nsrc more key buffers laid end to end
 * (what an all-gather of per-GPU key buffers produces). */
int spv_keys_merge(uint64_t *dkeys, const uint64_t *dsrc, int nsrc, int yrows, void *stream);
/* keys -> the reference's output arrays (device): idx yrows x kcols uint64,
 * dist yrows x kcols int32 / float32 */
int spv_keys_finalize_i32(const uint64_t *dkeys, int yrows, int kcols, uint64_t *didx, int32_t *ddist, void *stream);
int spv_keys_finalize_f32(const uint64_t *dkeys, int yrows, int kcols, uint64_t *didx, float *ddist, void *stream);

/* K6: ratio test + compaction on a key buffer.  dist_is_f32: divide as float32 / float64 (the
 * cascade matcher's output dtype, sentinel 2147483648.0f) instead of int32 / float64.  dmask
 * (nullable): yrows bytes, 1 = pass.  dmatches: room for yrows x 2 uint64; dcount: one uint32. */
size_t spv_ratio_ws_bytes(int yrows);
int spv_ratio_matches_device(const uint64_t *dkeys, int yrows, double min_ratio, int dist_is_f32, uint8_t *dmask,
                             uint64_t *dmatches, uint32_t *dcount, void *dws, size_t ws_bytes, void *stream);

/* K6 without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as spv_ratio_matches_device, but the per-block scratch is a file-scope device array
 * inside the library (65 540 bytes on every GPU that loads it, until exit) instead of a caller workspace, and the call
 * makes no cudaMalloc / cudaFree.  Limit: yrows <= spv_ratio_static_max_rows() = 4 194 304; beyond it the call returns
 * -1 with a message in spv_last_error() and launches nothing.  The array belongs to the CURRENT device and is shared
 * by all calls on it: calls that overlap on the GPU (different streams, different threads) race -- keep them on one
 * stream or order them yourself.  The first call on a device looks the array's address up (cudaGetSymbolAddress);
 * make one call before starting a graph capture. */
int spv_ratio_static_max_rows(void);
int spv_ratio_matches_device_static(const uint64_t *dkeys, int yrows, double min_ratio, int dist_is_f32, uint8_t *dmask,
                                    uint64_t *dmatches, uint32_t *dcount, void *stream);

/* K7: the quantiser on device rows.  dout_f32 and / or dout_u8 (= value + 128, what the matchers take;
 * either may be NULL): rows x spv_normalize_new_dim(dim). */
size_t spv_normalize_ws_bytes(int dim);
int spv_normalize_new_dim(int dim);
int spv_normalize_device(const float *dx, int rows, int dim, float *dout_f32, uint8_t *dout_u8, void *dws,
                         size_t ws_bytes, void *stream);

/* K1f on device rows: sum |x - y|^p top-k (BruteForceNn.h:37-120), float32 (is_int = 0) or int32 rows.
 * didx: yrows x k uint64, ddist: yrows x k float32 / int32, ascending, right-aligned and zero-filled in front
 * when xrows < k (what nn_bruteforce returns into numpy zeros).  dws: spv_nn_bruteforce_ws_bytes() of scratch. */
size_t spv_nn_bruteforce_ws_bytes(int xrows, int yrows, int k);
/* the same plus room for the TMA-fed float32 kernel's column-major copies of both matrices: with this much
 * scratch float32 rows (k <= 128) take the faster kernel; results are identical either way */
size_t spv_nn_bruteforce_ws_bytes_dim(int xrows, int yrows, int dim, int k);
int spv_nn_bruteforce_device(const void *dx, const void *dy, int is_int, int xrows, int yrows, int dim, int k,
                             double p, uint64_t *didx, void *ddist, void *dws, size_t ws_bytes, void *stream);

/* K0: (uint8)((int)v + 128), the cast of CascadingHashNn.h:236-239 (count % 4 == 0) */
int spv_f32_to_u8_device(const float *din, uint8_t *dout, size_t count, void *stream);
/* the same conversion on HOST arrays (any count): what nn_cascading_hash's staging threads apply to the caller's
 * float32 rows on their way into pinned memory, so that 1 byte per element crosses PCIe instead of 4; results equal
 * K0's bit for bit, including NaN, +-inf and out-of-range values (the device conversion saturates) */
void spv_f32_to_u8_host(const float *in, uint8_t *out, size_t count);
/* scratch bytes spv_cascade_device needs (0 = unsupported parameters) */
size_t spv_cascade_ws_bytes(int xrows, int yrows, int dim, int m, int n, int g);
/* K2-K4: cascade-of-hashes match of uint8 (value + 128) query rows dyu8 against
 * train rows dxu8 (replaces CascadingHashNn, src/CascadingHashNn.h:187-245), folded
 * into dkeys like spv_l1k2_device.  ddict: device float32 [n][dim][m] hyperplanes;
 * dict_terms: 0 if every entry is q / 32 with q an int8 in [-127, 127] (the dictionary of
 * spv_cascade_dictionary(quantize = 1): tcgen05 kind::i8 straight on the uint8 rows), 1 if every entry is
 * exact in bf16, else 3 (exact 3-term bf16 split). */
int spv_cascade_device(const uint8_t *dxu8, const uint8_t *dyu8, int xrows, int yrows, int dim, int m, int n, int g,
                       const float *ddict, int dict_terms, uint32_t xoffset, uint64_t *dkeys, void *dws,
                       size_t ws_bytes, void *stream);
/* With SPECTAVI_STAGE_TIMING=1 in the environment spv_cascade_device records CUDA events at its stage boundaries;
 * this returns the milliseconds of {dictionary image, K2 train side, K3 buckets, K2 query side, K4} of the calling
 * thread's last call (5 values; returns the count, 0 when timing is off).  Synchronises on the last event. */
int spv_cascade_stage_ms(double *out, int n);
/* K2 alone: hash codes [n][rows] int32 of uint8 (value + 128) rows (replaces
 * generate_hashcodes_for_data, CascadingHashNn.h:125-135) and, when g > 0, the
 * min(g, m) least-confident bit positions of every (table, row), 5 bits each
 * (generate_y_candidate_hashcodes, :153-169); scratch as spv_cascade_ws_bytes(rows, rows, ...) */
int spv_hash_codes_device(const uint8_t *drows, int rows, int dim, int m, int n, int g, const float *ddict,
                          int dict_terms, int32_t *dcodes, uint64_t *dflips, void *dws, size_t ws_bytes, void *stream);
/* the dictionary nn_cascading_hash draws (host): generate_hash_dict's loop order
 * (CascadingHashNn.h:86-100) over std::mt19937(seed); quantize: multiples of 1/32 */
void spv_cascade_dictionary(unsigned seed, int dim, int m, int n, int quantize, float *out);
unsigned spv_cascade_seed(void);

/* K8: RansacFitter::process_fundamental_matrix (src/RansacFitter.h:42-98) for nF fundamental matrices on device
 * pointers: dF [nF][9], dx0 / dx1 [npt][3] float64.  Outputs [nF]: dok, dcount (int32), dcams [nF][12] float64,
 * dmasks (nullable) [nF][ceil(npt / 32)] uint32.  dws: spv_ransac_ws_bytes(nF, npt) of scratch. */
size_t spv_ransac_ws_bytes(int nF, int npt);
int spv_ransac_score_device(const double *dF, int nF, const double *dx0, const double *dx1, int npt,
                            double required_percent_inliers, double reprojection_error_allowed,
                            int find_best_even_in_failure, double singular_value_ratio_allowed, int32_t *dok,
                            int32_t *dcount, double *dcams, uint32_t *dmasks, void *dws, size_t ws_bytes, void *stream);

/* number of kernels this library has launched since it was loaded */
uint64_t spv_launch_count(void);
/* last error text of the calling thread's most recent failed spv_* call */
const char *spv_last_error(void);
/* packed-SAD issue-rate microbenchmark (roofline denominator; pipe_peak.cu) */
int spv_pipe_peak(int device, int iters, double *out, const char **names, double *lanes_per_op);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* SPECTAVI_B200_H */
