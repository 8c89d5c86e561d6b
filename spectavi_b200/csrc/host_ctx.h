// host_ctx.h -- host-side runtime behind the drop-in entry points: cached
// per-GPU streams and device buffers, host<->device staging, and the
// single-process multi-GPU path (train rows sharded over SPECTAVI_GPUS devices,
// query rows broadcast, per-GPU top-2 all-gathered with NCCL and merged by K5).
//
// The reference keeps no state between calls (every call rebuilds everything,
// Spectavi.cpp:326-328) and its functions are re-entrant; the cache here holds
// only scratch memory and communicators, never results.  Single-GPU calls each
// take their own context (streams + buffers) from a pool, so concurrent callers
// run concurrently; the multi-GPU layout serialises on its NCCL communicators.
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>
#include <mutex>
#include <vector>

namespace spv {

int fail(const char* fmt, ...);
[[noreturn]] void die(const char* fn, const char* fmt, ...);

// The dictionary of the drop-in entry point (the role of generate_hash_dict,
// CascadingHashNn.h:86-100): n matrices dim x m of N(0,1) draws from std::mt19937
// in the reference's loop order, seeded with `seed` instead of std::random_device,
// and -- when quantize -- rounded to multiples of 1/32 in [-127/32, 127/32] so the
// projection is exact in every arithmetic (cascade.cu header).  out: [n][dim][m].
void cascade_dictionary(unsigned seed, int dim, int m, int n, bool quantize, float* out);
unsigned cascade_seed();   // env SPECTAVI_SEED, else a fixed default

// Asks for the ratio-tested, compacted match list (K6) instead of the yrows x k arrays: alloc(count)
// must return host memory for count x 2 uint64 (called once, from the calling thread).
struct MatchSink {
  double min_ratio;
  std::function<uint64_t*(uint64_t)> alloc;
};

struct Ctx;

class HostCall {
 public:
  explicit HostCall(const char* fn);
  ~HostCall();
  // brute-force L1 k=2, host in / host out (reference layout: yrows x 2)
  // (with a MatchSink: idx/dist are ignored and the surviving (train idx, query row) pairs are delivered)
  void l1k2(const uint8_t* x, const uint8_t* y, int xrows, int yrows, int dim, uint64_t* idx, int32_t* dist,
            const MatchSink* ms = nullptr);
  // brute-force sum |x-y|^p top-k over float32 / int32 rows (yrows x k out)
  // mu != 0 (for int32 rows: (int)mu != 0) selects the emulation of the reference's approximate prune
  void lpk(const void* x, const void* y, bool is_int, int xrows, int yrows, int dim, int k, double p, uint64_t* idx,
           void* dist, float mu = 0.f);
  // The filter overload of the generic matcher (BruteForceNn.h:37-49).  filter_x / filter_y: row-id sets, empty = all
  // rows.  idx / dist (yrows x k) are in-out: only rows in filter_y are written.
  void lpk_filtered(const float* x, const float* y, int xrows, int yrows, int dim, int k, double p,
                    const uint64_t* filter_x, int nfx, const uint64_t* filter_y, int nfy, uint64_t* idx, float* dist);
  // KMedians::find_nearest_neighbours (KMedians.h:260-295) given both clusterings: x = `other`, y = `this`.
  void kmedians_nn(const float* x, const float* y, int xrows, int yrows, int dim, const float* medians_x,
                   const int32_t* labels_x, int nmx, const float* medians_y, const int32_t* labels_y, int nmy, int c, int k,
                   uint64_t* idx, float* dist);
  // cascade-of-hashes matcher, float32 rows holding integers in [-128, 127] (yrows x kcols out)
  void cascade(const float* x, const float* y, int xrows, int yrows, int dim, int kcols, int m, int n, int g,
               uint64_t* idx, float* dist, const MatchSink* ms = nullptr);

  // K8: RansacFitter::process_fundamental_matrix for nF fundamental matrices (host in / host out).  ok / ninlier
  // [nF], cams [nF][12], masks (nullable) [nF][ceil(npt / 32)] words.  stop_at_success: hypotheses are scored in
  // batches and scoring stops behind the first batch that holds a hypothesis with percent > required, as
  // fit_essential's tries do (RansacFitter.h:189-191); entries of later batches are then left ok = 0.
  // Returns the number of hypotheses actually scored.
  int ransac(const double* x0, const double* x1, int npt, double required, double err_allowed, bool find_best,
             double sv_ratio, const double* F, int nF, bool stop_at_success, int32_t* ok, int32_t* ninlier, double* cams,
             uint32_t* masks);

  // the quantiser (feature.py:384-407): rows x dim float32 in, rows x ceil16(dim) float32 out
  void normalize(const float* x, int rows, int dim, float* out);

 private:
  struct FilterGroup { std::vector<uint32_t> ys, xs; };   // empty list = all rows (BruteForceNn.h:45-46)
  void filt_begin(const float* x, const float* y, int xrows, int yrows, int dim, int k, const uint64_t* idx, const float* dist);
  void filt_run(const std::vector<FilterGroup>& groups, int xrows, int yrows, int dim, int k, double p);
  void filt_end(int yrows, int k, uint64_t* idx, float* dist);
  void deliver_matches(const MatchSink& ms, int yrows, bool dist_f32);   // K6 on GPU 0's merged keys
  const char* fn_;
  Ctx* ctx_ = nullptr;                  // streams, staging lanes, scratch buffers of this call
  std::unique_lock<std::mutex> multi_lock_;    // held for the call when SPECTAVI_GPUS > 1
  int prev_dev_ = 0;                           // the caller's current device, restored on return
};

// K0 on the host (what nn_cascading_hash's staging threads run): K0's result bit for bit, any count
void f32_to_u8_host(const float* in, uint8_t* out, size_t count);

}  // namespace spv
