// abi.cu -- the extern "C" boundary of libspectavi.so (include/spectavi_b200.h).
//
// Part 1 mirrors reference src/Spectavi.cpp:258-298,321-336 call for call:
// size both outputs, let the CALLER allocate them through the NdArray callback
// (ctypes_ndarray/src/NdArray.h:41-45,67-73) from the calling thread, run the
// matcher, write the results into the caller's memory.  What differs is where
// the matcher runs: inputs are copied to the GPU(s), the sm_100a kernels of
// l1_top2.cu / lp_topk.cu / cascade.cu do the work, the yrows x k results come
// back.  There is no CPU implementation behind these symbols.
#include "../../include/spectavi_b200.h"
#include "spv_common.h"
#include "host_ctx.h"

#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace spv {

thread_local char g_err[512] = "";

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -1;
}

[[noreturn]] void die(const char* fn, const char* fmt, ...) {
  // the reference's error convention on this path is an exception escaping
  // extern "C" (BruteForceNnL1K2.h:74-81) => terminate; keep it observable.
  char msg[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(msg, sizeof(msg), fmt, ap);
  va_end(ap);
  fprintf(stderr, "libspectavi(b200): %s: %s\n", fn, msg);
  fflush(stderr);
  abort();
}

void out_alloc(NdArray* arr, uint64_t d0, uint64_t d1) {
  // ndarray_set_size + ndarray_alloc (NdArray.h:41-45,67-73)
  arr->m_ndim = 2;
  arr->m_shape[0] = d0;
  arr->m_shape[1] = d1;
  arr->m_data = arr->m_alloc((size_t)arr->m_sizeofdtype * d0 * d1);
}

}  // namespace spv

using namespace spv;

#define CU_OR_FAIL(expr)                                                             \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess) return fail("%s: %s", #expr, cudaGetErrorString(e_));     \
  } while (0)

extern "C" {

// ----------------------------------------------------------------- Part 2
int spv_keys_reset(uint64_t* dkeys, int yrows, void* stream) {
  CU_OR_FAIL(keys_reset(reinterpret_cast<unsigned long long*>(dkeys), yrows, (cudaStream_t)stream));
  return 0;
}

int spv_l1k2_device(const uint8_t* dx, const uint8_t* dy, int xrows, int yrows, int dim, uint32_t xoffset,
                    uint64_t* dkeys, void* stream) {
  if (xrows < 0 || yrows < 0) return fail("negative row count");
  if (dim <= 0 || dim % 16 != 0) return fail("dim (%d) must be a positive multiple of 16", dim);
  CU_OR_FAIL(l1_top2_u8(dx, dy, xrows, yrows, dim, xoffset, reinterpret_cast<unsigned long long*>(dkeys),
                        (cudaStream_t)stream));
  return 0;
}

int spv_keys_merge(uint64_t* dkeys, const uint64_t* dsrc, int nsrc, int yrows, void* stream) {
  CU_OR_FAIL(keys_merge(reinterpret_cast<unsigned long long*>(dkeys), reinterpret_cast<const unsigned long long*>(dsrc),
                        nsrc, yrows, (cudaStream_t)stream));
  return 0;
}

int spv_keys_finalize_i32(const uint64_t* dkeys, int yrows, int kcols, uint64_t* didx, int32_t* ddist, void* stream) {
  if (kcols < 1) return fail("kcols must be >= 1");
  CU_OR_FAIL(keys_finalize_i32(reinterpret_cast<const unsigned long long*>(dkeys), yrows, kcols, didx, ddist,
                               (cudaStream_t)stream));
  return 0;
}

int spv_keys_finalize_f32(const uint64_t* dkeys, int yrows, int kcols, uint64_t* didx, float* ddist, void* stream) {
  if (kcols < 1) return fail("kcols must be >= 1");
  CU_OR_FAIL(keys_finalize_f32(reinterpret_cast<const unsigned long long*>(dkeys), yrows, kcols, didx, ddist,
                               (cudaStream_t)stream));
  return 0;
}

size_t spv_ratio_ws_bytes(int yrows) { return ratio_ws_bytes(yrows); }

int spv_ratio_matches_device(const uint64_t* dkeys, int yrows, double min_ratio, int dist_is_f32, uint8_t* dmask,
                             uint64_t* dmatches, uint32_t* dcount, void* dws, size_t ws_bytes, void* stream) {
  if (yrows < 0) return fail("negative row count");
  CU_OR_FAIL(ratio_matches(reinterpret_cast<const unsigned long long*>(dkeys), yrows, min_ratio, dist_is_f32 != 0, dmask,
                           dmatches, dcount, dws, ws_bytes, (cudaStream_t)stream));
  return 0;
}

int spv_ratio_static_max_rows(void) { return ratio_static_max_rows(); }

int spv_ratio_matches_device_static(const uint64_t* dkeys, int yrows, double min_ratio, int dist_is_f32, uint8_t* dmask,
                                    uint64_t* dmatches, uint32_t* dcount, void* stream) {
  if (yrows < 0) return fail("negative row count");
  if (yrows > ratio_static_max_rows())
    return fail("spv_ratio_matches_device_static: yrows = %d exceeds the built-in scratch (%d rows); use "
                "spv_ratio_matches_device with a workspace", yrows, ratio_static_max_rows());
  CU_OR_FAIL(ratio_matches_static(reinterpret_cast<const unsigned long long*>(dkeys), yrows, min_ratio, dist_is_f32 != 0,
                                  dmask, dmatches, dcount, (cudaStream_t)stream));
  return 0;
}

size_t spv_normalize_ws_bytes(int dim) { return normalize_ws_bytes(dim); }
int spv_normalize_new_dim(int dim) { return normalize_new_dim(dim); }

int spv_normalize_device(const float* dx, int rows, int dim, float* dout_f32, uint8_t* dout_u8, void* dws, size_t ws_bytes,
                         void* stream) {
  if (rows < 0 || dim <= 0) return fail("bad shape %d x %d", rows, dim);
  CU_OR_FAIL(normalize_rows(dx, rows, dim, dout_f32, dout_u8, dws, ws_bytes, (cudaStream_t)stream));
  return 0;
}

static int sm_count() {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  return sms;
}

size_t spv_nn_bruteforce_ws_bytes(int xrows, int yrows, int k) {
  const int sms = sm_count();
  if (sms <= 0 || xrows < 0 || yrows < 0 || k < 1) return 0;
  return lp_partial_bytes(xrows, yrows, k, sms) + 8;
}

size_t spv_nn_bruteforce_ws_bytes_dim(int xrows, int yrows, int dim, int k) {
  const int sms = sm_count();
  if (sms <= 0 || xrows < 0 || yrows < 0 || dim < 1 || k < 1) return 0;
  return lp_ws_bytes(xrows, yrows, dim, k, sms);
}

int spv_nn_bruteforce_device(const void* dx, const void* dy, int is_int, int xrows, int yrows, int dim, int k, double p,
                             uint64_t* didx, void* ddist, void* dws, size_t ws_bytes, void* stream) {
  if (xrows < 0 || yrows < 0 || dim <= 0) return fail("bad shape %d x %d x %d", xrows, yrows, dim);
  if (k < 1) return fail("k = %d must be >= 1", k);
  if (ws_bytes < spv_nn_bruteforce_ws_bytes(xrows, yrows, k)) return fail("workspace too small");
  CU_OR_FAIL(lp_topk(dx, dy, is_int != 0, xrows, yrows, dim, k, p, dws, ws_bytes, didx, ddist, (cudaStream_t)stream));
  return 0;
}

void spv_f32_to_u8_host(const float* in, uint8_t* out, size_t count) { f32_to_u8_host(in, out, count); }

int spv_f32_to_u8_device(const float* din, uint8_t* dout, size_t count, void* stream) {
  CU_OR_FAIL(f32_to_u8(din, dout, count, (cudaStream_t)stream));
  return 0;
}

size_t spv_cascade_ws_bytes(int xrows, int yrows, int dim, int m, int n, int g) {
  return cascade_ws_bytes(xrows, yrows, dim, m, n, g);
}

int spv_cascade_device(const uint8_t* dxu8, const uint8_t* dyu8, int xrows, int yrows, int dim, int m, int n, int g,
                       const float* ddict, int dict_terms, uint32_t xoffset, uint64_t* dkeys, void* dws, size_t ws_bytes,
                       void* stream) {
  if (xrows < 0 || yrows < 0) return fail("negative row count");
  CU_OR_FAIL(cascade_match(dxu8, dyu8, xrows, yrows, dim, m, n, g, ddict, dict_terms, xoffset,
                           reinterpret_cast<unsigned long long*>(dkeys), dws, ws_bytes, (cudaStream_t)stream));
  return 0;
}

int spv_hash_codes_device(const uint8_t* drows, int rows, int dim, int m, int n, int g, const float* ddict, int dict_terms,
                          int32_t* dcodes, uint64_t* dflips, void* dws, size_t ws_bytes, void* stream) {
  CU_OR_FAIL(hash_codes(drows, rows, dim, m, n, g, ddict, dict_terms, dcodes,
                        reinterpret_cast<unsigned long long*>(dflips), dws, ws_bytes, (cudaStream_t)stream));
  return 0;
}

size_t spv_ransac_ws_bytes(int nF, int npt) { return ransac_ws_bytes(nF, npt); }

int spv_ransac_score_device(const double* dF, int nF, const double* dx0, const double* dx1, int npt,
                            double required_percent_inliers, double reprojection_error_allowed, int find_best_even_in_failure,
                            double singular_value_ratio_allowed, int32_t* dok, int32_t* dcount, double* dcams,
                            uint32_t* dmasks, void* dws, size_t ws_bytes, void* stream) {
  if (nF < 0 || npt < 1) return fail("bad shape: %d hypotheses, %d matches", nF, npt);
  if (ws_bytes < ransac_ws_bytes(nF, npt)) return fail("workspace too small");
  CU_OR_FAIL(ransac_score(dF, nF, dx0, dx1, npt, required_percent_inliers, reprojection_error_allowed,
                          find_best_even_in_failure != 0, singular_value_ratio_allowed, dok, dcount, dcams, dmasks, dws,
                          ws_bytes, (cudaStream_t)stream));
  return 0;
}

int spv_cascade_stage_ms(double* out, int n) { return cascade_stage_ms(out, n); }

void spv_cascade_dictionary(unsigned seed, int dim, int m, int n, int quantize, float* out) {
  cascade_dictionary(seed, dim, m, n, quantize != 0, out);
}

unsigned spv_cascade_seed(void) { return cascade_seed(); }

uint64_t spv_launch_count(void) { return launches(); }
const char* spv_last_error(void) { return g_err; }

// ----------------------------------------------------------------- Part 1
void nn_bruteforcel1k2(const uint8_t* x, const uint8_t* y, int xrows, int yrows, int dim, int nthreads,
                       NdArray* outidx, NdArray* outdist) {
  (void)nthreads;  // OpenMP thread count in the reference (BruteForceNnL1K2.h:92); no meaning here
  const int K = 2;
  out_alloc(outidx, (uint64_t)yrows, K);    // Spectavi.cpp:288-291: both outputs first, on the calling thread
  out_alloc(outdist, (uint64_t)yrows, K);
  // BruteForceNnL1K2.h:71-82 throws on these
  if (dim % 16 != 0) die("nn_bruteforcel1k2", "data dimension (%d) must be a multiple of 16", dim);
  if (xrows < 0 || yrows < 0 || dim <= 0) die("nn_bruteforcel1k2", "bad shape %d x %d x %d", xrows, yrows, dim);
  if (yrows == 0) return;
  HostCall hc("nn_bruteforcel1k2");
  hc.l1k2(x, y, xrows, yrows, dim, reinterpret_cast<uint64_t*>(outidx->m_data),
          reinterpret_cast<int32_t*>(outdist->m_data));
}

// ---- Part 1b: matcher + ratio test + compaction (example/ex01_essential_estimation.py:102-105)
static MatchSink sink_for(NdArray* outmatches, double min_ratio) {
  return MatchSink{min_ratio, [outmatches](uint64_t count) {
                     out_alloc(outmatches, count, 2);   // sized after the device count is known
                     return reinterpret_cast<uint64_t*>(outmatches->m_data);
                   }};
}

void spv_match_bruteforcel1k2(const uint8_t* x, const uint8_t* y, int xrows, int yrows, int dim, double min_ratio,
                              NdArray* outmatches) {
  const char* fn = "spv_match_bruteforcel1k2";
  if (dim % 16 != 0) die(fn, "data dimension (%d) must be a multiple of 16", dim);
  if (xrows < 0 || yrows < 0 || dim <= 0) die(fn, "bad shape %d x %d x %d", xrows, yrows, dim);
  if (yrows == 0) { out_alloc(outmatches, 0, 2); return; }
  const MatchSink ms = sink_for(outmatches, min_ratio);
  HostCall hc(fn);
  hc.l1k2(x, y, xrows, yrows, dim, nullptr, nullptr, &ms);
}

void spv_match_cascading_hash(const float* x, const float* y, int xrows, int yrows, int dim, int hash_bit_rate,
                              int num_hash_tables, int num_candidate_neighbours, double min_ratio, NdArray* outmatches) {
  const char* fn = "spv_match_cascading_hash";
  if (xrows < 0 || yrows < 0 || dim <= 0) die(fn, "bad shape %d x %d x %d", xrows, yrows, dim);
  if (dim % 16 != 0) die(fn, "data dimension (%d) must be a multiple of 16", dim);
  if (yrows == 0) { out_alloc(outmatches, 0, 2); return; }
  const MatchSink ms = sink_for(outmatches, min_ratio);
  HostCall hc(fn);
  hc.cascade(x, y, xrows, yrows, dim, 2, hash_bit_rate, num_hash_tables, num_candidate_neighbours, nullptr, nullptr, &ms);
}

void spv_normalize_to_ubyte(const float* x, int rows, int dim, NdArray* out) {
  const char* fn = "spv_normalize_to_ubyte";
  if (rows < 0 || dim <= 0) die(fn, "bad shape %d x %d", rows, dim);
  out_alloc(out, (uint64_t)rows, (uint64_t)normalize_new_dim(dim));
  if (rows == 0) return;
  HostCall hc(fn);
  hc.normalize(x, rows, dim, reinterpret_cast<float*>(out->m_data));
}

static void lp_entry(const char* fn, const void* x, const void* y, bool is_int, int xrows, int yrows, int dim, int k,
                     float p, float mu, NdArray* outidx, NdArray* outdist) {
  if (xrows < 0 || yrows < 0 || dim <= 0) die(fn, "bad shape %d x %d x %d", xrows, yrows, dim);
  if (k < 1) die(fn, "k = %d must be >= 1", k);
  out_alloc(outidx, (uint64_t)yrows, (uint64_t)k);    // Spectavi.cpp:262-265 / :275-278
  out_alloc(outdist, (uint64_t)yrows, (uint64_t)k);
  if (yrows == 0) return;
  HostCall hc(fn);
  // BruteForceNn's ctor takes `double p` (BruteForceNn.h:26-28): the float widens exactly
  hc.lpk(x, y, is_int, xrows, yrows, dim, k, (double)p, reinterpret_cast<uint64_t*>(outidx->m_data), outdist->m_data, mu);
}

void nn_bruteforce(const float* x, const float* y, int xrows, int yrows, int dim, int k, float p, float mu,
                   NdArray* outidx, NdArray* outdist) {
  // mu != 0 is the reference's approximate prune (BruteForceNn.h:80-86): emulated in its visiting order (lp_topk_mu)
  lp_entry("nn_bruteforce", x, y, false, xrows, yrows, dim, k, p, mu, outidx, outdist);
}
void nn_bruteforcei(const int* x, const int* y, int xrows, int yrows, int dim, int k, float p, float mu,
                    NdArray* outidx, NdArray* outdist) {
  lp_entry("nn_bruteforcei", x, y, true, xrows, yrows, dim, k, p, mu, outidx, outdist);
}
// ---- Part 1c: the filter overload of the generic matcher and the caller that drives it
void spv_nn_bruteforce_filtered(const float* x, const float* y, int xrows, int yrows, int dim, int k, float p, float mu,
                                const uint64_t* filter_x, int nfx, const uint64_t* filter_y, int nfy, uint64_t* outidx,
                                float* outdist) {
  const char* fn = "spv_nn_bruteforce_filtered";
  (void)mu;
  if (xrows < 0 || yrows < 0 || dim <= 0 || nfx < 0 || nfy < 0) die(fn, "bad shape %d x %d x %d", xrows, yrows, dim);
  if (k < 1 || k > LP_MAX_K) die(fn, "k = %d outside [1, %d]", k, LP_MAX_K);
  if (yrows == 0 || xrows == 0) return;
  HostCall hc(fn);
  hc.lpk_filtered(x, y, xrows, yrows, dim, k, (double)p, filter_x, nfx, filter_y, nfy, outidx, outdist);
}

void spv_kmedians_find_nearest_neighbours(const float* x, const float* y, int xrows, int yrows, int dim,
                                          const float* medians_x, const int32_t* labels_x, int nmx, const float* medians_y,
                                          const int32_t* labels_y, int nmy, int c, int k, NdArray* outidx, NdArray* outdist) {
  const char* fn = "spv_kmedians_find_nearest_neighbours";
  if (xrows < 0 || yrows < 0 || dim <= 0 || nmx < 1 || nmy < 1) die(fn, "bad shape %d x %d x %d, %d / %d medians", xrows, yrows, dim, nmx, nmy);
  if (k < 1 || k > LP_MAX_K) die(fn, "k = %d outside [1, %d]", k, LP_MAX_K);
  // median_idx is nmy x c and only min(c, nmx) columns are ever written there (KMedians.h:271-273)
  if (c < 1 || c > nmx || c > LP_MAX_K) die(fn, "c = %d outside [1, min(%d medians, %d)]", c, nmx, LP_MAX_K);
  out_alloc(outidx, (uint64_t)yrows, (uint64_t)k);    // Spectavi.cpp:310-313
  out_alloc(outdist, (uint64_t)yrows, (uint64_t)k);
  if (yrows == 0 || xrows == 0) return;
  HostCall hc(fn);
  hc.kmedians_nn(x, y, xrows, yrows, dim, medians_x, labels_x, nmx, medians_y, labels_y, nmy, c, k,
                 reinterpret_cast<uint64_t*>(outidx->m_data), reinterpret_cast<float*>(outdist->m_data));
}

// ---- Part 1d: the step after the matcher: RANSAC hypothesis scoring
static void ransac_check(const char* fn, int npt, int nF) {
  // RansacFitter's constructor throws on these (RansacFitter.h:139-148)
  if (npt < 10) die(fn, "Supplied less than 10 point matches, unsupported.");
  if (nF < 0) die(fn, "negative hypothesis count");
}

void spv_ransac_process_fundamental_matrices(const double* x0, const double* x1, int npt, double required_percent_inliers,
                                             double reprojection_error_allowed, int find_best_even_in_failure,
                                             double singular_value_ratio_allowed, const double* F, int nF, NdArray* ok,
                                             NdArray* ninlier, NdArray* cameras, NdArray* masks) {
  const char* fn = "spv_ransac_process_fundamental_matrices";
  ransac_check(fn, npt, nF);
  out_alloc(ok, (uint64_t)nF, 1);
  out_alloc(ninlier, (uint64_t)nF, 1);
  out_alloc(cameras, (uint64_t)nF, 12);
  if (masks) out_alloc(masks, (uint64_t)nF, (uint64_t)((npt + 31) / 32));
  if (nF == 0) return;
  HostCall hc(fn);
  hc.ransac(x0, x1, npt, required_percent_inliers, reprojection_error_allowed, find_best_even_in_failure != 0,
            singular_value_ratio_allowed, F, nF, false, reinterpret_cast<int32_t*>(ok->m_data),
            reinterpret_cast<int32_t*>(ninlier->m_data), reinterpret_cast<double*>(cameras->m_data),
            masks ? reinterpret_cast<uint32_t*>(masks->m_data) : nullptr);
}

void spv_ransac_fit_hypotheses(const double* x0, const double* x1, int npt, double required_percent_inliers,
                               double reprojection_error_allowed, int find_best_even_in_failure,
                               double singular_value_ratio_allowed, const double* F, int nF, bool* success,
                               NdArray* essential, NdArray* camera, double* inlier_percent, NdArray* inlier_idx) {
  const char* fn = "spv_ransac_fit_hypotheses";
  ransac_check(fn, npt, nF);
  const int W = (npt + 31) / 32;
  std::vector<int32_t> ok((size_t)nF), cnt((size_t)nF);
  std::vector<double> cams((size_t)nF * 12);
  std::vector<uint32_t> masks((size_t)nF * W);
  int scored = 0;
  if (nF > 0) {
    HostCall hc(fn);
    scored = hc.ransac(x0, x1, npt, required_percent_inliers, reprojection_error_allowed, find_best_even_in_failure != 0,
                       singular_value_ratio_allowed, F, nF, true, ok.data(), cnt.data(), cams.data(), masks.data());
  }
  // fit_essential's bookkeeping (RansacFitter.h:166-199), the tries taken in order: a scored hypothesis replaces the
  // best so far when (percent > required || find_best) and percent > best percent; the first one above `required`
  // ends the search.
  bool succ = false;
  double best_pct = 0;
  int best = -1;
  for (int i = 0; i < scored && !succ; ++i) {
    if (!ok[i]) continue;
    const double pct = (double)cnt[i] / (double)npt;
    if ((pct > required_percent_inliers || find_best_even_in_failure) && best_pct < pct) {
      succ = pct > required_percent_inliers;
      best_pct = pct;
      best = i;
    }
  }
  *success = succ;
  *inlier_percent = best_pct;
  // ndarray_copy_matrix of the members (Spectavi.cpp:82-87): never-assigned members are 0 x 0, the default camera is I(3,4)
  if (best >= 0) {
    out_alloc(essential, 3, 3);
    memcpy(essential->m_data, F + (size_t)best * 9, 9 * 8);
    out_alloc(camera, 3, 4);
    memcpy(camera->m_data, cams.data() + (size_t)best * 12, 12 * 8);
    out_alloc(inlier_idx, (uint64_t)cnt[best], 1);
    int32_t* o = reinterpret_cast<int32_t*>(inlier_idx->m_data);
    int k = 0;
    for (int ix = 0; ix < npt; ++ix)
      if ((masks[(size_t)best * W + (ix >> 5)] >> (ix & 31)) & 1u) o[k++] = ix;
  } else {
    out_alloc(essential, 0, 0);
    out_alloc(camera, 3, 4);
    double* c = reinterpret_cast<double*>(camera->m_data);
    for (int i = 0; i < 12; ++i) c[i] = (i % 5 == 0) ? 1.0 : 0.0;
    out_alloc(inlier_idx, 0, 0);
  }
}

void nn_cascading_hash(const float* x, const float* y, int xrows, int yrows, int dim, int k, int hash_bit_rate,
                       int num_hash_tables, int num_candidate_neighbours, NdArray* outidx, NdArray* outdist) {
  const char* fn = "nn_cascading_hash";
  if (xrows < 0 || yrows < 0 || dim <= 0) die(fn, "bad shape %d x %d x %d", xrows, yrows, dim);
  // the reference sizes its outputs yrows x k but assigns a yrows x 2 matrix to them
  // (CascadingHashNn.h:244): only k == 2 is well defined there; k > 2 leaves columns >= 2 zero here
  if (k < 2) die(fn, "k = %d: the matcher always produces 2 neighbours (CascadingHashNn.h:229-245)", k);
  out_alloc(outidx, (uint64_t)yrows, (uint64_t)k);     // Spectavi.cpp:329-332
  out_alloc(outdist, (uint64_t)yrows, (uint64_t)k);
  if (dim % 16 != 0) die(fn, "data dimension (%d) must be a multiple of 16", dim);   // BruteForceNnL1K2.h:78-81
  if (yrows == 0) return;
  HostCall hc(fn);
  hc.cascade(x, y, xrows, yrows, dim, k, hash_bit_rate, num_hash_tables, num_candidate_neighbours,
             reinterpret_cast<uint64_t*>(outidx->m_data), reinterpret_cast<float*>(outdist->m_data));
}

}  // extern "C"
