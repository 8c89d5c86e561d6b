// host_ctx.cu -- see host_ctx.h.
#include "host_ctx.h"
#include <algorithm>
#include "spv_common.h"

#include <dlfcn.h>
#include <nccl.h>   // types only; the library is dlopen()ed when SPECTAVI_GPUS > 1

#include <cmath>
#include <cstdlib>
#include <random>
#include <cstring>
#include <vector>

namespace spv {
namespace {

std::mutex g_mu;

struct Buf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = n + n / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
};

struct Gpu {
  int dev = -1;
  cudaStream_t stream = nullptr;
  Buf x, y, keys, gather, oidx, odist, xf, yf, dict, ws, rws, sel;
};

struct Nccl {
  void* lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::vector<ncclComm_t> comms;
};

struct Runtime {
  std::vector<Gpu> gpus;
  Nccl nccl;
  bool ready = false;
};
Runtime g_rt;

#define CU(fn, expr)                                                     \
  do {                                                                   \
    cudaError_t e_ = (expr);                                             \
    if (e_ != cudaSuccess) die(fn, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)
#define NC(fn, expr)                                                     \
  do {                                                                   \
    ncclResult_t r_ = (expr);                                            \
    if (r_ != ncclSuccess) die(fn, "%s: %s", #expr, g_rt.nccl.GetErrorString ? g_rt.nccl.GetErrorString(r_) : "nccl error"); \
  } while (0)

void load_nccl(const char* fn, int ngpu) {
  Nccl& n = g_rt.nccl;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    n.lib = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
    if (n.lib) break;
  }
  if (!n.lib) die(fn, "SPECTAVI_GPUS=%d needs NCCL, dlopen(libnccl.so.2) failed: %s", ngpu, dlerror());
#define SYM(field, name)                                                  \
  n.field = reinterpret_cast<decltype(n.field)>(dlsym(n.lib, name));      \
  if (!n.field) die(fn, "NCCL symbol %s missing", name);
  SYM(CommInitAll, "ncclCommInitAll") SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd")
  SYM(AllGather, "ncclAllGather") SYM(Broadcast, "ncclBroadcast") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
  std::vector<int> devs(ngpu);
  for (int g = 0; g < ngpu; ++g) devs[g] = g_rt.gpus[g].dev;
  n.comms.resize(ngpu);
  NC(fn, n.CommInitAll(n.comms.data(), ngpu, devs.data()));
}

void init_runtime(const char* fn) {
  if (g_rt.ready) return;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1)
    die(fn, "no CUDA device (%s); this library has no CPU implementation", cudaGetErrorString(e));
  int want = 1;
  if (const char* s = getenv("SPECTAVI_GPUS")) {
    want = (strcmp(s, "all") == 0) ? ndev : atoi(s);
    if (want < 1 || want > ndev) die(fn, "SPECTAVI_GPUS=%s but %d device(s) visible", s, ndev);
  }
  int first = 0;
  cudaGetDevice(&first);   // respect the caller's current device for the 1-GPU case
  g_rt.gpus.resize(want);
  for (int g = 0; g < want; ++g) {
    Gpu& G = g_rt.gpus[g];
    G.dev = (want == 1) ? first : g;
    CU(fn, cudaSetDevice(G.dev));
    CU(fn, cudaStreamCreateWithFlags(&G.stream, cudaStreamNonBlocking));
  }
  if (want > 1) load_nccl(fn, want);
  CU(fn, cudaSetDevice(g_rt.gpus[0].dev));
  g_rt.ready = true;
}

}  // namespace

HostCall::HostCall(const char* fn) : fn_(fn), lock_(g_mu) { init_runtime(fn); }
HostCall::~HostCall() {}

// K6 on the merged key buffer of GPU 0: ratio test + compaction on the device, then only
// count x 16 bytes come back (instead of yrows x 2 x 12)
void HostCall::deliver_matches(const MatchSink& ms, int yrows, bool dist_f32) {
  Gpu& R = g_rt.gpus[0];
  CU(fn_, cudaSetDevice(R.dev));
  const size_t wsb = ratio_ws_bytes(yrows);
  CU(fn_, R.rws.ensure(wsb + 16));
  CU(fn_, R.oidx.ensure((size_t)yrows * 2 * sizeof(uint64_t)));
  unsigned* dcount = reinterpret_cast<unsigned*>(static_cast<char*>(R.rws.p) + ((wsb + 7) / 8) * 8);
  CU(fn_, ratio_matches((const unsigned long long*)R.keys.p, yrows, ms.min_ratio, dist_f32, nullptr, (uint64_t*)R.oidx.p,
                        dcount, R.rws.p, wsb, R.stream));
  unsigned count = 0;
  CU(fn_, cudaMemcpyAsync(&count, dcount, sizeof(count), cudaMemcpyDeviceToHost, R.stream));
  for (size_t g = 0; g < g_rt.gpus.size(); ++g) {
    CU(fn_, cudaSetDevice(g_rt.gpus[g].dev));
    CU(fn_, cudaStreamSynchronize(g_rt.gpus[g].stream));
  }
  CU(fn_, cudaSetDevice(R.dev));
  uint64_t* out = ms.alloc(count);
  if (count) {
    CU(fn_, cudaMemcpyAsync(out, R.oidx.p, (size_t)count * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, R.stream));
    CU(fn_, cudaStreamSynchronize(R.stream));
  }
}

void HostCall::l1k2(const uint8_t* x, const uint8_t* y, int xrows, int yrows, int dim, uint64_t* idx, int32_t* dist,
                    const MatchSink* ms) {
  const int G = (int)g_rt.gpus.size();
  const size_t ybytes = (size_t)yrows * dim, kbytes = (size_t)2 * yrows * sizeof(unsigned long long);
  const int per = (xrows + G - 1) / G;   // train rows per GPU
  int prev = 0;
  cudaGetDevice(&prev);
  // stage inputs: query rows to GPU 0 (broadcast below), train shard g to GPU g
  for (int g = 0; g < G; ++g) {
    Gpu& D = g_rt.gpus[g];
    CU(fn_, cudaSetDevice(D.dev));
    const int x0 = g * per < xrows ? g * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    CU(fn_, D.y.ensure(ybytes));
    CU(fn_, D.keys.ensure(kbytes));
    if (xn > 0) {
      CU(fn_, D.x.ensure((size_t)xn * dim));
      CU(fn_, cudaMemcpyAsync(D.x.p, x + (size_t)x0 * dim, (size_t)xn * dim, cudaMemcpyHostToDevice, D.stream));
    }
    if (g == 0) CU(fn_, cudaMemcpyAsync(D.y.p, y, ybytes, cudaMemcpyHostToDevice, D.stream));
    CU(fn_, keys_reset((unsigned long long*)D.keys.p, yrows, D.stream));
  }
  if (G > 1) {
    NC(fn_, g_rt.nccl.GroupStart());
    for (int g = 0; g < G; ++g)
      NC(fn_, g_rt.nccl.Broadcast(g_rt.gpus[0].y.p, g_rt.gpus[g].y.p, ybytes, ncclUint8, 0, g_rt.nccl.comms[g],
                                  g_rt.gpus[g].stream));
    NC(fn_, g_rt.nccl.GroupEnd());
  }
  // K1 on every shard
  for (int g = 0; g < G; ++g) {
    Gpu& D = g_rt.gpus[g];
    const int x0 = g * per < xrows ? g * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    CU(fn_, cudaSetDevice(D.dev));
    CU(fn_, l1_top2_u8((const uint8_t*)D.x.p, (const uint8_t*)D.y.p, xn, yrows, dim, (uint32_t)x0,
                       (unsigned long long*)D.keys.p, D.stream));
  }
  Gpu& R = g_rt.gpus[0];
  if (G > 1) {
    // all-gather the per-GPU key buffers, K5 merge on GPU 0
    for (int g = 0; g < G; ++g) {
      CU(fn_, cudaSetDevice(g_rt.gpus[g].dev));
      CU(fn_, g_rt.gpus[g].gather.ensure(kbytes * G));
    }
    NC(fn_, g_rt.nccl.GroupStart());
    for (int g = 0; g < G; ++g)
      NC(fn_, g_rt.nccl.AllGather(g_rt.gpus[g].keys.p, g_rt.gpus[g].gather.p, (size_t)2 * yrows, ncclUint64,
                                  g_rt.nccl.comms[g], g_rt.gpus[g].stream));
    NC(fn_, g_rt.nccl.GroupEnd());
    CU(fn_, cudaSetDevice(R.dev));
    CU(fn_, keys_merge((unsigned long long*)R.keys.p, (const unsigned long long*)R.gather.p, G, yrows, R.stream));
  }
  if (ms) {
    deliver_matches(*ms, yrows, false);
    cudaSetDevice(prev);
    return;
  }
  CU(fn_, cudaSetDevice(R.dev));
  CU(fn_, R.oidx.ensure((size_t)yrows * 2 * sizeof(uint64_t)));
  CU(fn_, R.odist.ensure((size_t)yrows * 2 * sizeof(int32_t)));
  CU(fn_, keys_finalize_i32((const unsigned long long*)R.keys.p, yrows, 2, (uint64_t*)R.oidx.p, (int32_t*)R.odist.p,
                            R.stream));
  CU(fn_, cudaMemcpyAsync(idx, R.oidx.p, (size_t)yrows * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, R.stream));
  CU(fn_, cudaMemcpyAsync(dist, R.odist.p, (size_t)yrows * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, R.stream));
  for (int g = 0; g < G; ++g) {
    CU(fn_, cudaSetDevice(g_rt.gpus[g].dev));
    CU(fn_, cudaStreamSynchronize(g_rt.gpus[g].stream));
  }
  cudaSetDevice(prev);
}

void cascade_dictionary(unsigned seed, int dim, int m, int n, bool quantize, float* out) {
  std::mt19937 gen(seed);
  std::normal_distribution<float> normal(0.f, 1.f);
  for (int t = 0; t < n; ++t)
    for (int i = 0; i < dim; ++i)
      for (int j = 0; j < m; ++j) {
        float v = normal(gen);
        if (quantize) {
          float q = nearbyintf(v * 32.f);
          q = q > 127.f ? 127.f : (q < -127.f ? -127.f : q);
          v = q / 32.f;
        }
        out[((size_t)t * dim + i) * m + j] = v;
      }
}

unsigned cascade_seed() {
  if (const char* s = getenv("SPECTAVI_SEED")) return (unsigned)strtoul(s, nullptr, 0);
  return 0x5EED5EEDu;
}

void HostCall::cascade(const float* x, const float* y, int xrows, int yrows, int dim, int kcols, int m, int n, int g,
                       uint64_t* idx, float* dist, const MatchSink* ms) {
  const int G = (int)g_rt.gpus.size();
  const size_t ybytes = (size_t)yrows * dim, kbytes = (size_t)2 * yrows * sizeof(unsigned long long);
  const int per = (xrows + G - 1) / G;   // train rows per GPU, as for brute force
  std::vector<float> hdict((size_t)n * dim * m);
  cascade_dictionary(cascade_seed(), dim, m, n, true, hdict.data());
  int prev = 0;
  cudaGetDevice(&prev);
  for (int gi = 0; gi < G; ++gi) {
    Gpu& D = g_rt.gpus[gi];
    CU(fn_, cudaSetDevice(D.dev));
    const int x0 = gi * per < xrows ? gi * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    const size_t wsb = cascade_ws_bytes(xn, yrows, dim, m, n, g);
    if (wsb == 0) die(fn_, "unsupported parameters dim=%d m=%d n=%d g=%d (need dim %% 16 == 0, dim <= 512, 1 <= m <= 31, g <= 12)", dim, m, n, g);
    CU(fn_, D.y.ensure(ybytes));
    CU(fn_, D.keys.ensure(kbytes));
    CU(fn_, D.dict.ensure(hdict.size() * 4));
    CU(fn_, D.ws.ensure(wsb));
    CU(fn_, cudaMemcpyAsync(D.dict.p, hdict.data(), hdict.size() * 4, cudaMemcpyHostToDevice, D.stream));
    if (xn > 0) {
      CU(fn_, D.xf.ensure((size_t)xn * dim * 4));
      CU(fn_, D.x.ensure((size_t)xn * dim));
      CU(fn_, cudaMemcpyAsync(D.xf.p, x + (size_t)x0 * dim, (size_t)xn * dim * 4, cudaMemcpyHostToDevice, D.stream));
      CU(fn_, f32_to_u8((const float*)D.xf.p, (uint8_t*)D.x.p, (size_t)xn * dim, D.stream));   // CascadingHashNn.h:236-239
    }
    if (gi == 0) {
      CU(fn_, D.yf.ensure(ybytes * 4));
      CU(fn_, cudaMemcpyAsync(D.yf.p, y, ybytes * 4, cudaMemcpyHostToDevice, D.stream));
      CU(fn_, f32_to_u8((const float*)D.yf.p, (uint8_t*)D.y.p, ybytes, D.stream));
    }
    CU(fn_, keys_reset((unsigned long long*)D.keys.p, yrows, D.stream));
  }
  if (G > 1) {
    NC(fn_, g_rt.nccl.GroupStart());
    for (int gi = 0; gi < G; ++gi)
      NC(fn_, g_rt.nccl.Broadcast(g_rt.gpus[0].y.p, g_rt.gpus[gi].y.p, ybytes, ncclUint8, 0, g_rt.nccl.comms[gi],
                                  g_rt.gpus[gi].stream));
    NC(fn_, g_rt.nccl.GroupEnd());
  }
  for (int gi = 0; gi < G; ++gi) {
    Gpu& D = g_rt.gpus[gi];
    const int x0 = gi * per < xrows ? gi * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    CU(fn_, cudaSetDevice(D.dev));
    CU(fn_, cascade_match((const uint8_t*)D.x.p, (const uint8_t*)D.y.p, xn, yrows, dim, m, n, g, (const float*)D.dict.p, 0,
                          (uint32_t)x0, (unsigned long long*)D.keys.p, D.ws.p, D.ws.cap, D.stream));
  }
  Gpu& R = g_rt.gpus[0];
  if (G > 1) {
    for (int gi = 0; gi < G; ++gi) {
      CU(fn_, cudaSetDevice(g_rt.gpus[gi].dev));
      CU(fn_, g_rt.gpus[gi].gather.ensure(kbytes * G));
    }
    NC(fn_, g_rt.nccl.GroupStart());
    for (int gi = 0; gi < G; ++gi)
      NC(fn_, g_rt.nccl.AllGather(g_rt.gpus[gi].keys.p, g_rt.gpus[gi].gather.p, (size_t)2 * yrows, ncclUint64,
                                  g_rt.nccl.comms[gi], g_rt.gpus[gi].stream));
    NC(fn_, g_rt.nccl.GroupEnd());
    CU(fn_, cudaSetDevice(R.dev));
    CU(fn_, keys_merge((unsigned long long*)R.keys.p, (const unsigned long long*)R.gather.p, G, yrows, R.stream));
  }
  if (ms) {
    deliver_matches(*ms, yrows, true);
    cudaSetDevice(prev);
    return;
  }
  CU(fn_, cudaSetDevice(R.dev));
  const size_t ib = (size_t)yrows * kcols * sizeof(uint64_t), db = (size_t)yrows * kcols * sizeof(float);
  CU(fn_, R.oidx.ensure(ib));
  CU(fn_, R.odist.ensure(db));
  CU(fn_, cudaMemsetAsync(R.oidx.p, 0, ib, R.stream));    // columns >= 2 keep the caller's zeros (Spectavi.cpp:329-334)
  CU(fn_, cudaMemsetAsync(R.odist.p, 0, db, R.stream));
  CU(fn_, keys_finalize_f32((const unsigned long long*)R.keys.p, yrows, kcols, (uint64_t*)R.oidx.p, (float*)R.odist.p, R.stream));
  CU(fn_, cudaMemcpyAsync(idx, R.oidx.p, ib, cudaMemcpyDeviceToHost, R.stream));
  CU(fn_, cudaMemcpyAsync(dist, R.odist.p, db, cudaMemcpyDeviceToHost, R.stream));
  for (int gi = 0; gi < G; ++gi) {
    CU(fn_, cudaSetDevice(g_rt.gpus[gi].dev));
    CU(fn_, cudaStreamSynchronize(g_rt.gpus[gi].stream));
  }
  cudaSetDevice(prev);
}

void HostCall::normalize(const float* x, int rows, int dim, float* out) {
  Gpu& D = g_rt.gpus[0];
  int prev = 0;
  cudaGetDevice(&prev);
  CU(fn_, cudaSetDevice(D.dev));
  const size_t ib = (size_t)rows * dim * 4, ob = (size_t)rows * normalize_new_dim(dim) * 4;
  CU(fn_, D.xf.ensure(ib));
  CU(fn_, D.yf.ensure(ob));
  CU(fn_, D.rws.ensure(normalize_ws_bytes(dim)));
  CU(fn_, cudaMemcpyAsync(D.xf.p, x, ib, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, normalize_rows((const float*)D.xf.p, rows, dim, (float*)D.yf.p, nullptr, D.rws.p, D.rws.cap, D.stream));
  CU(fn_, cudaMemcpyAsync(out, D.yf.p, ob, cudaMemcpyDeviceToHost, D.stream));
  CU(fn_, cudaStreamSynchronize(D.stream));
  cudaSetDevice(prev);
}

void HostCall::lpk(const void* x, const void* y, bool is_int, int xrows, int yrows, int dim, int k, double p,
                   uint64_t* idx, void* dist) {
  // single GPU: the graded shapes of this entry point are small (test/test_feature.py:67-81)
  Gpu& D = g_rt.gpus[0];
  int prev = 0, sms = 0;
  cudaGetDevice(&prev);
  CU(fn_, cudaSetDevice(D.dev));
  CU(fn_, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, D.dev));
  const size_t xb = (size_t)xrows * dim * 4, yb = (size_t)yrows * dim * 4;
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  CU(fn_, D.x.ensure(xb ? xb : 4));
  CU(fn_, D.y.ensure(yb ? yb : 4));
  CU(fn_, D.gather.ensure(lp_partial_bytes(xrows, yrows, k, sms) + 8));
  CU(fn_, D.oidx.ensure(ib));
  CU(fn_, D.odist.ensure(db));
  if (xb) CU(fn_, cudaMemcpyAsync(D.x.p, x, xb, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, cudaMemcpyAsync(D.y.p, y, yb, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, lp_topk(D.x.p, D.y.p, is_int, xrows, yrows, dim, k, p, (unsigned long long*)D.gather.p, (uint64_t*)D.oidx.p,
                  D.odist.p, D.stream));
  CU(fn_, cudaMemcpyAsync(idx, D.oidx.p, ib, cudaMemcpyDeviceToHost, D.stream));
  CU(fn_, cudaMemcpyAsync(dist, D.odist.p, db, cudaMemcpyDeviceToHost, D.stream));
  CU(fn_, cudaStreamSynchronize(D.stream));
  cudaSetDevice(prev);
}

// ---- the filter overload and its caller --------------------------------
// One GPU, like lpk.  filt_begin uploads the rows and the caller's in-out arrays once; filt_run runs
// any number of groups in ONE launch (their query lists must be disjoint); filt_end downloads.
void HostCall::filt_begin(const float* x, const float* y, int xrows, int yrows, int dim, int k, const uint64_t* idx,
                          const float* dist) {
  Gpu& D = g_rt.gpus[0];
  CU(fn_, cudaSetDevice(D.dev));
  const size_t xb = (size_t)xrows * dim * 4, yb = (size_t)yrows * dim * 4;
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  CU(fn_, D.x.ensure(xb ? xb : 4));
  CU(fn_, D.y.ensure(yb ? yb : 4));
  CU(fn_, D.oidx.ensure(ib ? ib : 8));
  CU(fn_, D.odist.ensure(db ? db : 4));
  if (xb) CU(fn_, cudaMemcpyAsync(D.x.p, x, xb, cudaMemcpyHostToDevice, D.stream));
  if (yb) CU(fn_, cudaMemcpyAsync(D.y.p, y, yb, cudaMemcpyHostToDevice, D.stream));
  if (ib) CU(fn_, cudaMemcpyAsync(D.oidx.p, idx, ib, cudaMemcpyHostToDevice, D.stream));
  if (db) CU(fn_, cudaMemcpyAsync(D.odist.p, dist, db, cudaMemcpyHostToDevice, D.stream));
}

void HostCall::filt_run(const std::vector<FilterGroup>& groups, int xrows, int yrows, int dim, int k, double p) {
  Gpu& D = g_rt.gpus[0];
  int sms = 0;
  CU(fn_, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, D.dev));
  const int TR = lp_filtered_tile_rows();
  std::vector<uint32_t> ysel, xsel;
  std::vector<int4> tiles;
  int max_xcnt = 0;
  for (const FilterGroup& g : groups) {
    const int ypos = (int)ysel.size(), xpos = (int)xsel.size();
    if (g.ys.empty()) for (int i = 0; i < yrows; ++i) ysel.push_back((uint32_t)i);
    else ysel.insert(ysel.end(), g.ys.begin(), g.ys.end());
    if (g.xs.empty()) for (int i = 0; i < xrows; ++i) xsel.push_back((uint32_t)i);
    else xsel.insert(xsel.end(), g.xs.begin(), g.xs.end());
    const int ycnt = (int)ysel.size() - ypos, xcnt = (int)xsel.size() - xpos;
    if (xcnt > max_xcnt) max_xcnt = xcnt;
    for (int o = 0; o < ycnt; o += TR) tiles.push_back(make_int4(ypos + o, ycnt - o < TR ? ycnt - o : TR, xpos, xcnt));
  }
  if (tiles.empty() || max_xcnt == 0) return;
  if (ysel.size() > 0x7FFFFFFFull || xsel.size() > 0x7FFFFFFFull) die(fn_, "filter lists too long");
  const size_t yb = ysel.size() * 4, xb = xsel.size() * 4, tb = tiles.size() * sizeof(int4);
  const size_t yoff = 0, xoff = (yb + 15) / 16 * 16, toff = xoff + (xb + 15) / 16 * 16;
  CU(fn_, D.sel.ensure(toff + tb));
  CU(fn_, D.gather.ensure(lp_filtered_partial_bytes((int)tiles.size(), max_xcnt, (int)ysel.size(), k, sms) + 8));
  char* base = static_cast<char*>(D.sel.p);
  CU(fn_, cudaMemcpyAsync(base + yoff, ysel.data(), yb, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, cudaMemcpyAsync(base + xoff, xsel.data(), xb, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, cudaMemcpyAsync(base + toff, tiles.data(), tb, cudaMemcpyHostToDevice, D.stream));
  CU(fn_, lp_topk_filtered((const float*)D.x.p, (const float*)D.y.p, dim, k, p, (const int4*)(base + toff), (int)tiles.size(),
                           (const uint32_t*)(base + yoff), (int)ysel.size(), (const uint32_t*)(base + xoff), max_xcnt,
                           (unsigned long long*)D.gather.p, (uint64_t*)D.oidx.p, (float*)D.odist.p, D.stream));
  CU(fn_, cudaStreamSynchronize(D.stream));   // the host vectors above are about to go away
}

void HostCall::filt_end(int yrows, int k, uint64_t* idx, float* dist) {
  Gpu& D = g_rt.gpus[0];
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  if (ib) CU(fn_, cudaMemcpyAsync(idx, D.oidx.p, ib, cudaMemcpyDeviceToHost, D.stream));
  if (db) CU(fn_, cudaMemcpyAsync(dist, D.odist.p, db, cudaMemcpyDeviceToHost, D.stream));
  CU(fn_, cudaStreamSynchronize(D.stream));
}

// a std::unordered_set<size_t> built from a list: duplicates collapse; ascending order here
static std::vector<uint32_t> as_set(const char* fn, const uint64_t* f, int n, int rows, const char* what) {
  std::vector<uint32_t> v;
  v.reserve(n > 0 ? n : 0);
  for (int i = 0; i < n; ++i) {
    if (f[i] >= (uint64_t)rows) die(fn, "%s[%d] = %llu is not a row (< %d)", what, i, (unsigned long long)f[i], rows);
    v.push_back((uint32_t)f[i]);
  }
  std::sort(v.begin(), v.end());
  v.erase(std::unique(v.begin(), v.end()), v.end());
  return v;
}

void HostCall::lpk_filtered(const float* x, const float* y, int xrows, int yrows, int dim, int k, double p,
                            const uint64_t* filter_x, int nfx, const uint64_t* filter_y, int nfy, uint64_t* idx, float* dist) {
  int prev = 0;
  cudaGetDevice(&prev);
  std::vector<FilterGroup> groups(1);
  groups[0].xs = as_set(fn_, filter_x, nfx, xrows, "filter_x");
  groups[0].ys = as_set(fn_, filter_y, nfy, yrows, "filter_y");
  filt_begin(x, y, xrows, yrows, dim, k, idx, dist);
  filt_run(groups, xrows, yrows, dim, k, p);
  filt_end(yrows, k, idx, dist);
  cudaSetDevice(prev);
}

void HostCall::kmedians_nn(const float* x, const float* y, int xrows, int yrows, int dim, const float* medians_x,
                           const int32_t* labels_x, int nmx, const float* medians_y, const int32_t* labels_y, int nmy, int c,
                           int k, uint64_t* idx, float* dist) {
  int prev = 0;
  cudaGetDevice(&prev);
  // KMedians.h:268-273: the c nearest `other` medians of every median, L1 (K1f on the GPU, exact)
  std::vector<uint64_t> near((size_t)nmy * c);
  std::vector<float> near_d((size_t)nmy * c);
  lpk(medians_x, medians_y, false, nmx, nmy, dim, c, 1.0, near.data(), near_d.data());
  // cluster member lists (m_median_idx), ascending row id
  std::vector<std::vector<uint32_t>> xl(nmx), yl(nmy);
  for (int i = 0; i < xrows; ++i) {
    if (labels_x[i] < 0 || labels_x[i] >= nmx) die(fn_, "labels_x[%d] = %d outside [0, %d)", i, labels_x[i], nmx);
    xl[labels_x[i]].push_back((uint32_t)i);
  }
  for (int i = 0; i < yrows; ++i) {
    if (labels_y[i] < 0 || labels_y[i] >= nmy) die(fn_, "labels_y[%d] = %d outside [0, %d)", i, labels_y[i], nmy);
    yl[labels_y[i]].push_back((uint32_t)i);
  }
  filt_begin(x, y, xrows, yrows, dim, k, idx, dist);
  // KMedians.h:275-293, one filtered search per query cluster, in cluster order.  Clusters are disjoint, so a run of
  // groups is ONE launch; a group whose query list or candidate union is empty means "all rows" to BruteForceNn
  // (BruteForceNn.h:45-46) and overlaps the others, so it is launched on its own, in sequence, like the reference.
  std::vector<FilterGroup> batch;
  for (int g = 0; g < nmy; ++g) {
    FilterGroup fg;
    fg.ys = yl[g];
    for (int j = 0; j < c; ++j) {
      const std::vector<uint32_t>& m = xl[near[(size_t)g * c + j]];
      fg.xs.insert(fg.xs.end(), m.begin(), m.end());
    }
    if (fg.ys.empty() || fg.xs.empty()) {
      if (!batch.empty()) { filt_run(batch, xrows, yrows, dim, k, 1.0); batch.clear(); }
      filt_run(std::vector<FilterGroup>(1, fg), xrows, yrows, dim, k, 1.0);
    } else {
      batch.push_back(std::move(fg));
    }
  }
  if (!batch.empty()) filt_run(batch, xrows, yrows, dim, k, 1.0);
  filt_end(yrows, k, idx, dist);
  cudaSetDevice(prev);
}

}  // namespace spv
