// host_ctx.cu -- see host_ctx.h.
#include "host_ctx.h"
#include <algorithm>
#include "spv_common.h"

#include <dlfcn.h>
#include <nccl.h>   // types only; the library is dlopen()ed when SPECTAVI_GPUS > 1

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <random>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

namespace spv {
namespace {

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

// pinned host scratch (results come back device -> pinned by DMA, then pinned -> the caller's pageable memory)
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    const size_t want = n + n / 8 + 4096;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocPortable);
    if (e == cudaSuccess) cap = want;
    return e;
  }
};

// One upload lane: a copy stream and two pinned staging buffers.  The caller's arrays are ordinary pageable
// numpy memory (Spectavi.cpp:284-298 hands over raw pointers), so every byte goes pageable -> pinned (CPU
// memcpy) -> device (DMA); with two buffers the memcpy of sub-chunk i + 1 runs under the DMA of sub-chunk i,
// and -- because the DMA sits on its own stream -- under whatever the compute stream is doing.
constexpr size_t LANE_BYTES = (size_t)4 << 20;
constexpr int NLANE = 16;                      // lanes a context can have; they are created on first use (lanes_init)
struct Lane {
  cudaStream_t copy = nullptr;
  void* pin[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};   // the DMA out of pin[b] has finished
  cudaEvent_t ready = nullptr;                // the last piece handed to this lane is on the device
  bool used[2] = {false, false};
  int turn = 0;
};

struct Gpu {
  int dev = -1;
  cudaStream_t stream = nullptr;
  Lane lane[NLANE];
  int lanes_ready = 0;          // lanes [0, lanes_ready) exist
  cudaEvent_t ev[16] = {};      // scratch events (per-chunk completion of the final pass)
  Buf x, y, keys, gather, oidx, odist, xf, yf, dict, ws, rws, sel;
  PinBuf res;
};

struct Nccl {
  void* lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::vector<ncclComm_t> comms;
};

}  // namespace

// A context = the streams, staging lanes and scratch buffers one call works with.  The reference's functions are
// re-entrant (no globals), so concurrent callers must not serialise on one set of buffers: single-GPU calls take a
// context from a pool (a new one is made when all are busy); the multi-GPU layout (SPECTAVI_GPUS > 1) has one
// context, because its NCCL communicators admit one collective sequence at a time.
struct SpinBarrier {               // the per-GPU host threads of one call meet here (short waits, dedicated threads)
  std::atomic<int> count{0}, sense{0};
  int n = 1;
  void wait() {
    const int s = sense.load(std::memory_order_acquire);
    if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
      count.store(0, std::memory_order_relaxed);
      sense.store(s ^ 1, std::memory_order_release);
    } else {
      while (sense.load(std::memory_order_acquire) == s) std::this_thread::yield();
    }
  }
};

struct Ctx {
  std::vector<Gpu> gpus;
  PinBuf ypin;                     // multi-GPU: the query rows staged ONCE (portable pinned memory), DMA'd by every GPU
  SpinBarrier bar;
};

// ---- K0 on the host (see upload_f32_as_u8 below) ----
void f32_to_u8_avx2(const float* s, uint8_t* d, size_t n);   // host_cvt.cu (compiled with -mavx2)

static inline uint8_t f32_to_u8_one(float v) {
  // cvt.rzi.s32.f32: NaN -> 0, saturating; then + 128 in wrapping arithmetic, low byte
  int32_t i;
  if (!(v == v)) i = 0;
  else if (v >= 2147483648.0f) i = INT32_MAX;
  else if (v <= -2147483648.0f) i = INT32_MIN;
  else i = (int32_t)v;
  return (uint8_t)(((uint32_t)i + 128u) & 0xFFu);
}

void f32_to_u8_host(const float* s, uint8_t* d, size_t n) {
  static const bool avx2 = __builtin_cpu_supports("avx2");
  size_t done = 0;
  if (avx2) {
    done = n & ~(size_t)31;
    f32_to_u8_avx2(s, d, done);
  }
  for (size_t i = done; i < n; ++i) d[i] = f32_to_u8_one(s[i]);
}

namespace {

struct Runtime {
  std::mutex mu;                                  // guards init, the pool and nothing else
  std::vector<int> devs;
  std::vector<std::unique_ptr<Ctx>> pool;         // idle single-GPU contexts
  std::unique_ptr<Ctx> multi;                     // the multi-GPU context
  std::mutex multi_mu;                            // held for the whole of a multi-GPU call
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
  SYM(AllGather, "ncclAllGather") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
  n.comms.resize(ngpu);
  NC(fn, n.CommInitAll(n.comms.data(), ngpu, g_rt.devs.data()));
}

std::unique_ptr<Ctx> make_ctx(const char* fn) {
  std::unique_ptr<Ctx> c(new Ctx);
  c->gpus.resize(g_rt.devs.size());
  for (size_t g = 0; g < g_rt.devs.size(); ++g) {
    Gpu& G = c->gpus[g];
    G.dev = g_rt.devs[g];
    CU(fn, cudaSetDevice(G.dev));
    CU(fn, cudaStreamCreateWithFlags(&G.stream, cudaStreamNonBlocking));
  }
  CU(fn, cudaSetDevice(g_rt.devs[0]));
  return c;
}

void init_runtime(const char* fn) {   // g_rt.mu held
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
  g_rt.devs.resize(want);
  for (int g = 0; g < want; ++g) g_rt.devs[g] = (want == 1) ? first : g;
  if (want > 1) {
    g_rt.multi = make_ctx(fn);
    load_nccl(fn, want);
    CU(fn, cudaSetDevice(g_rt.devs[0]));
  }
  g_rt.ready = true;
}

// ---- optional host-side timeline (SPECTAVI_TRACE=1): one stderr line per step, milliseconds since the call began
struct Trace {
  bool on;
  std::chrono::steady_clock::time_point t0;
  Trace() : on(getenv("SPECTAVI_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {}
  void operator()(const char* what, long long a = 0, long long b = 0) const {
    if (!on) return;
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    fprintf(stderr, "[spv trace] %8.3f ms  %s %lld %lld\n", ms, what, a, b);
  }
};

// ---- staged uploads ------------------------------------------------------------------------------------
void lanes_init(const char* fn, Gpu& D, int n = 8) {
  if (D.lanes_ready >= n) return;
  if (D.lanes_ready == 0)
    for (cudaEvent_t& e : D.ev) CU(fn, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (int li = D.lanes_ready; li < n && li < NLANE; ++li) {
    Lane& L = D.lane[li];
    CU(fn, cudaStreamCreateWithFlags(&L.copy, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      CU(fn, cudaHostAlloc(&L.pin[b], LANE_BYTES, cudaHostAllocDefault));
      CU(fn, cudaEventCreateWithFlags(&L.done[b], cudaEventDisableTiming));
    }
    CU(fn, cudaEventCreateWithFlags(&L.ready, cudaEventDisableTiming));
  }
  D.lanes_ready = n < NLANE ? n : NLANE;
}

// pageable host -> device through the lane's pinned buffers; returns once the last sub-chunk is QUEUED.
// The device's current-device must be D's.  `ready` is recorded behind the piece.
void lane_upload(const char* fn, Lane& L, void* dst, const void* src, size_t bytes, size_t sub = LANE_BYTES) {
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  for (size_t off = 0; off < bytes; off += sub) {
    const size_t n = bytes - off < sub ? bytes - off : sub;
    const int b = L.turn;
    L.turn ^= 1;
    if (L.used[b]) CU(fn, cudaEventSynchronize(L.done[b]));
    memcpy(L.pin[b], s + off, n);
    CU(fn, cudaMemcpyAsync(d + off, L.pin[b], n, cudaMemcpyHostToDevice, L.copy));
    CU(fn, cudaEventRecord(L.done[b], L.copy));
    L.used[b] = true;
  }
  CU(fn, cudaEventRecord(L.ready, L.copy));
}

void copy_out(void* dst, const void* src, size_t bytes);

// Piece sizes of a streamed matrix: the first piece is small, so that compute starts after a fraction of a
// millisecond of upload, and the pieces double up to `cap` rows so that the launch count stays small.
std::vector<int> piece_rows(int rows, int first, int cap, int align) {
  std::vector<int> v;
  int n = first;
  while (rows > 0) {
    int take = n < rows ? n : rows;
    if (rows - take < first / 2) take = rows;          // no sliver at the end
    else take = take / align * align;
    v.push_back(take);
    rows -= take;
    if (n < cap) n *= 2;
  }
  return v;
}

// K1 over one GPU's train shard, inputs streamed from pageable host memory (the body of HostCall::l1k2 for one
// device; with SPECTAVI_GPUS > 1 one host thread per device runs it).  Query rows go up first in doubling pieces,
// each matched against the first, small train piece as soon as it has landed; then the remaining train pieces,
// each against all query rows.  K1 folds into one key buffer across launches (xoffset) and seeds its thresholds
// from it (l1_top2.cu), so the chunking costs no extra slow-path work.  Nothing is synchronised here.
// With hidx / hdist (single-GPU calls) the shard also delivers: the LAST pass over a slice of query rows makes that
// slice final, so the last train piece is matched slice by slice (up to 4), each followed by keys_finalize on the
// slice and a DMA into pinned memory on a second stream; the host copies slice c out while slice c + 1 computes, and
// only the last slice's delivery is exposed (traced on C3: finalize + D2H + copy-out of all 4.8 MB cost 1.4 ms behind K1).
// Multi-GPU (coop != nullptr): the query rows are identical for every GPU, so the G host threads stage each query
// piece TOGETHER into one portable pinned buffer (each copies 1 / G of it), meet at a barrier, and every GPU then
// DMAs the piece from there -- the pageable -> pinned memcpy, the slow half of the path, is done once instead of G
// times (measured at 8 GPUs with per-GPU staging: 8 x 128 MB of memcpy starved the GPUs for ~9 ms of a 238 ms call).
struct Coop { Ctx* c; int g, G; };
void l1k2_stream_shard(const char* fn, Gpu& D, const uint8_t* x, const uint8_t* y, int x0, int xn, int yrows, int dim,
                       uint64_t* hidx = nullptr, int32_t* hdist = nullptr, const Coop* coop = nullptr) {
  const Trace tr;
  CU(fn, cudaSetDevice(D.dev));
  lanes_init(fn, D);
  tr("lanes ready");
  const size_t ybytes = (size_t)yrows * dim, kbytes = (size_t)2 * yrows * sizeof(unsigned long long);
  CU(fn, D.y.ensure(ybytes));
  CU(fn, D.keys.ensure(kbytes));
  CU(fn, keys_reset((unsigned long long*)D.keys.p, yrows, D.stream));
  if (xn <= 0 && !coop) return;
  if (xn > 0) CU(fn, D.x.ensure((size_t)xn * dim));
  Lane& L = D.lane[0];
  // first train piece sized so that matching it against all query rows takes about as long as uploading them
  // (x0 * yrows * dim / 4 sad-ops at ~17 T/s  vs  yrows * dim bytes at ~10 GB/s  =>  ~7 k rows, any shape)
  const std::vector<int> xp = piece_rows(xn > 0 ? xn : 1, 8192, 65536, 64), yp = piece_rows(yrows, 8192, 131072, 256);
  uint8_t* dx = (uint8_t*)D.x.p;
  uint8_t* dy = (uint8_t*)D.y.p;
  unsigned long long* dk = (unsigned long long*)D.keys.p;
  tr("buffers ready");
  if (xn > 0) lane_upload(fn, L, dx, x + (size_t)x0 * dim, (size_t)xp[0] * dim);
  tr("x piece queued", xp[0]);
  int yo = 0;
  for (int rows : yp) {
    if (coop) {
      uint8_t* ypin = static_cast<uint8_t*>(coop->c->ypin.p);
      const size_t r0 = (size_t)rows * coop->g / coop->G, r1 = (size_t)rows * (coop->g + 1) / coop->G;
      memcpy(ypin + ((size_t)yo + r0) * dim, y + ((size_t)yo + r0) * dim, (r1 - r0) * dim);
      coop->c->bar.wait();                               // the whole piece is in pinned memory
      if (xn > 0) {
        CU(fn, cudaMemcpyAsync(dy + (size_t)yo * dim, ypin + (size_t)yo * dim, (size_t)rows * dim, cudaMemcpyHostToDevice, L.copy));
        CU(fn, cudaEventRecord(L.ready, L.copy));
      }
    } else {
      lane_upload(fn, L, dy + (size_t)yo * dim, y + (size_t)yo * dim, (size_t)rows * dim);
    }
    if (xn > 0) {
      CU(fn, cudaStreamWaitEvent(D.stream, L.ready, 0));
      CU(fn, l1_top2_u8(dx, dy + (size_t)yo * dim, xp[0], rows, dim, (uint32_t)x0, dk + yo, D.stream, yrows));
    }
    tr("y piece queued + K1 launched", rows);
    yo += rows;
  }
  if (xn <= 0) return;
  int xo = xp[0];
  const size_t last = xp.size() - 1;
  const bool deliver = hidx != nullptr && last >= 1;      // (one train piece only: the caller finalises the whole buffer)
  for (size_t i = 1; i < xp.size(); ++i) {
    lane_upload(fn, L, dx + (size_t)xo * dim, x + (size_t)(x0 + xo) * dim, (size_t)xp[i] * dim);
    CU(fn, cudaStreamWaitEvent(D.stream, L.ready, 0));
    if (!(deliver && i == last)) {
      CU(fn, l1_top2_u8(dx + (size_t)xo * dim, dy, xp[i], yrows, dim, (uint32_t)(x0 + xo), dk, D.stream, yrows));
      tr("x piece queued + K1 launched", xp[i]);
      xo += xp[i];
      continue;
    }
    // the final pass, slice by slice
    int nq = yrows / 8192;
    nq = nq < 1 ? 1 : (nq > 4 ? 4 : nq);
    const int per = ((yrows + nq - 1) / nq + 255) / 256 * 256;
    const size_t ib = (size_t)yrows * 2 * sizeof(uint64_t), db = (size_t)yrows * 2 * sizeof(int32_t), doff = (ib + 255) & ~(size_t)255;
    CU(fn, D.oidx.ensure(ib));
    CU(fn, D.odist.ensure(db));
    CU(fn, D.res.ensure(doff + db));
    char* pin = static_cast<char*>(D.res.p);
    cudaStream_t down = D.lane[1].copy;
    int nslice = 0, y0s[4], yns[4];
    for (int c0 = 0; c0 < yrows; c0 += per) {
      const int rows = yrows - c0 < per ? yrows - c0 : per;
      CU(fn, l1_top2_u8(dx + (size_t)xo * dim, dy + (size_t)c0 * dim, xp[i], rows, dim, (uint32_t)(x0 + xo), dk + c0, D.stream, yrows));
      CU(fn, keys_finalize_i32(dk + c0, rows, 2, (uint64_t*)D.oidx.p + (size_t)c0 * 2, (int32_t*)D.odist.p + (size_t)c0 * 2, D.stream, yrows));
      CU(fn, cudaEventRecord(D.ev[nslice], D.stream));
      CU(fn, cudaStreamWaitEvent(down, D.ev[nslice], 0));
      CU(fn, cudaMemcpyAsync(pin + (size_t)c0 * 16, (char*)D.oidx.p + (size_t)c0 * 16, (size_t)rows * 16, cudaMemcpyDeviceToHost, down));
      CU(fn, cudaMemcpyAsync(pin + doff + (size_t)c0 * 8, (char*)D.odist.p + (size_t)c0 * 8, (size_t)rows * 8, cudaMemcpyDeviceToHost, down));
      CU(fn, cudaEventRecord(D.ev[8 + nslice], down));
      y0s[nslice] = c0; yns[nslice] = rows;
      ++nslice;
    }
    tr("final pass queued in slices", nslice);
    for (int c = 0; c < nslice; ++c) {
      CU(fn, cudaEventSynchronize(D.ev[8 + c]));
      copy_out((char*)hidx + (size_t)y0s[c] * 16, pin + (size_t)y0s[c] * 16, (size_t)yns[c] * 16);
      copy_out((char*)hdist + (size_t)y0s[c] * 8, pin + doff + (size_t)y0s[c] * 8, (size_t)yns[c] * 8);
    }
    CU(fn, cudaStreamSynchronize(D.stream));
    tr("results on the host (sliced)");
    xo += xp[i];
  }
  if (hidx && !deliver) {
    // a single train piece: finalise the whole buffer behind the query pieces
    CU(fn, D.oidx.ensure((size_t)yrows * 2 * sizeof(uint64_t)));
    CU(fn, D.odist.ensure((size_t)yrows * 2 * sizeof(int32_t)));
    CU(fn, keys_finalize_i32(dk, yrows, 2, (uint64_t*)D.oidx.p, (int32_t*)D.odist.p, D.stream));
  }
}

// Host threads (= lanes) of one pageable upload.  Measured on the box (a 16-core VM; tools/bench_e2e_lanes.py on C5's
// 1 GB of float rows): 1 lane 11.5 GB/s, 2 lanes 21.6, 4 lanes 36.7, 6 lanes 38.8, 8 lanes 37.7.  The scheme in
// isolation (tools/ubench_stage_mt.cu): a memcpy thread alone moves 12-19 GB/s, the DMA alone 50-55 GB/s, together they
// share the host's memory system -- 256 MB: 40 GB/s with 4 threads, 43 with 6; 21 MB: 0.87 ms with 4 threads and
// 1 MB sub-chunks, no better with more (the DMAs only drain once the copy threads are done).  SPECTAVI_UPLOAD_LANES
// overrides.
int upload_lanes() {
  static const int n = [] {
    const char* e = getenv("SPECTAVI_UPLOAD_LANES");
    int v = e ? atoi(e) : 4;
    return v < 1 ? 1 : (v > 8 ? 8 : v);
  }();
  return n;
}

// ---- float32 rows -> device uint8 (value + 128), the cascade's inputs (CascadingHashNn.h:236-239) ----------------
// The rows have to pass through the CPU anyway (pageable -> pinned), so the conversion happens THERE: a lane's thread
// reads 4 bytes per element and writes 1 into its pinned buffer, and 1 byte per element crosses PCIe -- against 4 read,
// 4 written, 4 sent when the floats are staged as they are and K0 converts on the device.  Measured on the box
// (tools/ubench_stage_cvt.cu, 512 MB of floats): staged as floats 12.0 ms at best (4-8 threads, 45 GB/s: the host's
// memory system is the limit); converted while staged 11.6 / 8.0 / 6.4 ms with 4 / 6 / 8 threads (84 GB/s of floats) --
// it scales with the threads because each one moves 2.5x fewer bytes.  Semantics are K0's, bit for bit
// (f32_to_u8_host below; tests/test_f32_to_u8_host.py).  SPECTAVI_CVT=device keeps the old path (A/B runs).
// threads of one converting upload: SPECTAVI_CVT_LANES (1 .. NLANE), default 3/4 of the host's cores within [4, NLANE],
// shared between the GPUs of a call.  C5 end to end on the 16-core box (tools/bench_e2e_cvt.py): K0 on the device
// 28.8 ms; converted while staged 31.9 / 20.0 / 18.3 / 17.8 ms with 4 / 8 / 12 / 16 threads.
int cvt_lanes(int gpus) {
  static const int n = [] {
    const char* e = getenv("SPECTAVI_CVT_LANES");
    int v = e ? atoi(e) : (int)(std::thread::hardware_concurrency() * 3 / 4);
    if (!e && v < 4) v = 4;
    return v < 1 ? 1 : (v > NLANE ? NLANE : v);
  }();
  const int per = n / (gpus > 0 ? gpus : 1);
  return per < 2 ? 2 : per;
}

void upload_f32_as_u8(const char* fn, Gpu& D, const float* src, size_t count, uint8_t* dst, float* dtmp, int gpus) {
  static const bool on_device = [] { const char* e = getenv("SPECTAVI_CVT"); return e && !strcmp(e, "device"); }();
  const size_t per_chunk = on_device ? LANE_BYTES / 4 : ((size_t)1 << 20);   // elements per sub-chunk
  const size_t nchunks = (count + per_chunk - 1) / per_chunk;
  const size_t want = on_device ? (size_t)upload_lanes() : (size_t)cvt_lanes(gpus);
  const int nl = (int)(nchunks < want ? (nchunks ? nchunks : 1) : want);
  lanes_init(fn, D, nl > 8 ? nl : 8);
  auto work = [&](int li) {
    cudaSetDevice(D.dev);
    Lane& L = D.lane[li];
    const size_t c0 = nchunks * li / nl, c1 = nchunks * (li + 1) / nl;
    for (size_t c = c0; c < c1; ++c) {
      const size_t off = c * per_chunk, n = count - off < per_chunk ? count - off : per_chunk;
      const int b = L.turn;
      L.turn ^= 1;
      if (L.used[b]) CU(fn, cudaEventSynchronize(L.done[b]));
      if (on_device) {
        memcpy(L.pin[b], src + off, n * 4);
        float* dt = dtmp + ((size_t)li * 2 + b) * per_chunk;     // the device-side twin of pin[b]
        CU(fn, cudaMemcpyAsync(dt, L.pin[b], n * 4, cudaMemcpyHostToDevice, L.copy));
        CU(fn, f32_to_u8(dt, dst + off, n, L.copy));             // K0
      } else {
        f32_to_u8_host(src + off, static_cast<uint8_t*>(L.pin[b]), n);
        CU(fn, cudaMemcpyAsync(dst + off, L.pin[b], n, cudaMemcpyHostToDevice, L.copy));
      }
      CU(fn, cudaEventRecord(L.done[b], L.copy));
      L.used[b] = true;
    }
    CU(fn, cudaEventRecord(L.ready, L.copy));
  };
  if (nl == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int li = 1; li < nl; ++li) th.emplace_back(work, li);
    work(0);
    for (std::thread& t : th) t.join();
  }
  for (int li = 0; li < nl; ++li) CU(fn, cudaStreamWaitEvent(D.stream, D.lane[li].ready, 0));
}

void sync_all(const char* fn, Ctx& c) {
  for (Gpu& D : c.gpus) {
    CU(fn, cudaSetDevice(D.dev));
    CU(fn, cudaStreamSynchronize(D.stream));
  }
}

// pinned -> the caller's (pageable, usually never-touched: numpy zeros) memory.  A direct device -> pageable copy ran at
// ~2 GB/s on the box (4.8 MB of results = 2.3 ms, traced): the driver's own staging plus first-touch page faults in
// one thread.  Here the DMA lands in pinned memory at link speed and the final copy is split over a few threads.
void copy_out(void* dst, const void* src, size_t bytes) {
  const size_t slice = (size_t)1 << 20;
  int nt = (int)(bytes / slice);
  nt = nt > 6 ? 6 : nt;
  if (nt <= 1) { memcpy(dst, src, bytes); return; }
  std::vector<std::thread> th;
  const size_t per = ((bytes / nt) + 4095) & ~(size_t)4095;
  for (int t = 1; t < nt; ++t) {
    const size_t o = per * t, n = o >= bytes ? 0 : (bytes - o < per || t == nt - 1 ? bytes - o : per);
    if (n) th.emplace_back([=] { memcpy(static_cast<char*>(dst) + o, static_cast<const char*>(src) + o, n); });
  }
  memcpy(dst, src, per < bytes ? per : bytes);
  for (std::thread& t : th) t.join();
}

// results of GPU R: two device arrays -> the caller's arrays, through R.res
void fetch_results(const char* fn, Ctx& c, Gpu& R, void* h0, const void* d0, size_t b0, void* h1, const void* d1, size_t b1) {
  CU(fn, cudaSetDevice(R.dev));
  const size_t o1 = (b0 + 255) & ~(size_t)255;
  CU(fn, R.res.ensure(o1 + b1));
  char* pin = static_cast<char*>(R.res.p);
  CU(fn, cudaMemcpyAsync(pin, d0, b0, cudaMemcpyDeviceToHost, R.stream));
  CU(fn, cudaMemcpyAsync(pin + o1, d1, b1, cudaMemcpyDeviceToHost, R.stream));
  sync_all(fn, c);
  copy_out(h0, pin, b0);
  copy_out(h1, pin + o1, b1);
}

// run f(g) for every GPU of the context: inline for one, one host thread per GPU otherwise
template <typename F>
void per_gpu(Ctx& c, F f) {
  const int G = (int)c.gpus.size();
  if (G == 1) { f(0); return; }
  std::vector<std::thread> th;
  for (int g = 1; g < G; ++g) th.emplace_back(f, g);
  f(0);
  for (std::thread& t : th) t.join();
}


// all-gather the per-GPU key buffers, K5 merge on GPU 0; no-op for one GPU
void merge_shards(const char* fn, Ctx& c, int yrows) {
  const int G = (int)c.gpus.size();
  if (G == 1) return;
  const size_t kbytes = (size_t)2 * yrows * sizeof(unsigned long long);
  for (Gpu& D : c.gpus) {
    CU(fn, cudaSetDevice(D.dev));
    CU(fn, D.gather.ensure(kbytes * G));
  }
  NC(fn, g_rt.nccl.GroupStart());
  for (int g = 0; g < G; ++g)
    NC(fn, g_rt.nccl.AllGather(c.gpus[g].keys.p, c.gpus[g].gather.p, (size_t)2 * yrows, ncclUint64, g_rt.nccl.comms[g],
                               c.gpus[g].stream));
  NC(fn, g_rt.nccl.GroupEnd());
  Gpu& R = c.gpus[0];
  CU(fn, cudaSetDevice(R.dev));
  CU(fn, keys_merge((unsigned long long*)R.keys.p, (const unsigned long long*)R.gather.p, G, yrows, R.stream));
}

}  // namespace

HostCall::HostCall(const char* fn) : fn_(fn) {
  cudaGetDevice(&prev_dev_);
  std::unique_lock<std::mutex> lk(g_rt.mu);
  init_runtime(fn);
  if (g_rt.multi) {
    lk.unlock();
    multi_lock_ = std::unique_lock<std::mutex>(g_rt.multi_mu);
    ctx_ = g_rt.multi.get();
    return;
  }
  if (!g_rt.pool.empty()) {
    ctx_ = g_rt.pool.back().release();
    g_rt.pool.pop_back();
  } else {
    ctx_ = make_ctx(fn).release();
  }
}

HostCall::~HostCall() {
  if (!multi_lock_.owns_lock() && ctx_) {
    std::lock_guard<std::mutex> lk(g_rt.mu);
    g_rt.pool.emplace_back(ctx_);
  }
  cudaSetDevice(prev_dev_);
}

// K6 on the merged key buffer of GPU 0: ratio test + compaction on the device, then only
// count x 16 bytes come back (instead of yrows x 2 x 12)
void HostCall::deliver_matches(const MatchSink& ms, int yrows, bool dist_f32) {
  Gpu& R = ctx_->gpus[0];
  CU(fn_, cudaSetDevice(R.dev));
  const size_t wsb = ratio_ws_bytes(yrows);
  CU(fn_, R.rws.ensure(wsb + 16));
  CU(fn_, R.oidx.ensure((size_t)yrows * 2 * sizeof(uint64_t)));
  unsigned* dcount = reinterpret_cast<unsigned*>(static_cast<char*>(R.rws.p) + ((wsb + 7) / 8) * 8);
  CU(fn_, ratio_matches((const unsigned long long*)R.keys.p, yrows, ms.min_ratio, dist_f32, nullptr, (uint64_t*)R.oidx.p,
                        dcount, R.rws.p, wsb, R.stream));
  unsigned count = 0;
  CU(fn_, cudaMemcpyAsync(&count, dcount, sizeof(count), cudaMemcpyDeviceToHost, R.stream));
  sync_all(fn_, *ctx_);
  CU(fn_, cudaSetDevice(R.dev));
  uint64_t* out = ms.alloc(count);
  if (count) {
    CU(fn_, cudaMemcpyAsync(out, R.oidx.p, (size_t)count * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, R.stream));
    CU(fn_, cudaStreamSynchronize(R.stream));
  }
}

void HostCall::l1k2(const uint8_t* x, const uint8_t* y, int xrows, int yrows, int dim, uint64_t* idx, int32_t* dist,
                    const MatchSink* ms) {
  Ctx& c = *ctx_;
  const int G = (int)c.gpus.size();
  const int per = (xrows + G - 1) / G;   // train rows per GPU
  const char* fn = fn_;
  // every GPU streams its own train shard and its own copy of the query rows over its own PCIe link
  const bool self_deliver = G == 1 && !ms && xrows > 0;
  if (G > 1) {
    CU(fn_, cudaSetDevice(c.gpus[0].dev));
    CU(fn_, c.ypin.ensure((size_t)yrows * dim));
    c.bar.n = G;
  }
  per_gpu(c, [&](int g) {
    const int x0 = (long long)g * per < xrows ? g * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    const Coop coop = {&c, g, G};
    l1k2_stream_shard(fn, c.gpus[g], x, y, x0, xn, yrows, dim, self_deliver ? idx : nullptr, self_deliver ? dist : nullptr,
                      G > 1 ? &coop : nullptr);
  });
  const Trace tr;
  if (self_deliver) {
    Gpu& R0 = c.gpus[0];
    const std::vector<int> xp = piece_rows(xrows, 8192, 65536, 64);
    if (xp.size() >= 2) return;                       // delivered slice by slice inside the shard
    fetch_results(fn_, c, R0, idx, R0.oidx.p, (size_t)yrows * 2 * sizeof(uint64_t), dist, R0.odist.p, (size_t)yrows * 2 * sizeof(int32_t));
    return;
  }
  merge_shards(fn_, c, yrows);
  if (ms) {
    deliver_matches(*ms, yrows, false);
    return;
  }
  Gpu& R = c.gpus[0];
  CU(fn_, cudaSetDevice(R.dev));
  if (tr.on) { CU(fn_, cudaStreamSynchronize(R.stream)); tr("(trace only) K1 launches drained"); }
  CU(fn_, R.oidx.ensure((size_t)yrows * 2 * sizeof(uint64_t)));
  CU(fn_, R.odist.ensure((size_t)yrows * 2 * sizeof(int32_t)));
  CU(fn_, keys_finalize_i32((const unsigned long long*)R.keys.p, yrows, 2, (uint64_t*)R.oidx.p, (int32_t*)R.odist.p,
                            R.stream));
  fetch_results(fn_, c, R, idx, R.oidx.p, (size_t)yrows * 2 * sizeof(uint64_t), dist, R.odist.p, (size_t)yrows * 2 * sizeof(int32_t));
  tr("results on the host");
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
  Ctx& c = *ctx_;
  const int G = (int)c.gpus.size();
  const size_t ybytes = (size_t)yrows * dim;
  const int per = (xrows + G - 1) / G;   // train rows per GPU, as for brute force
  std::vector<float> hdict((size_t)n * dim * m);
  cascade_dictionary(cascade_seed(), dim, m, n, true, hdict.data());
  const char* fn = fn_;
  const Trace tr;
  tr("cascade: dictionary drawn");
  // per GPU: its train shard and all query rows are converted to uint8 by the staging lanes' threads on their way
  // into pinned memory (upload_f32_as_u8), then K2-K4 on the shard with the SAME dictionary
  per_gpu(c, [&](int gi) {
    Gpu& D = c.gpus[gi];
    CU(fn, cudaSetDevice(D.dev));
    const int x0 = (long long)gi * per < xrows ? gi * per : xrows;
    const int xn = (x0 + per <= xrows) ? per : xrows - x0;
    const size_t wsb = cascade_ws_bytes(xn, yrows, dim, m, n, g);
    if (wsb == 0) die(fn, "unsupported parameters dim=%d m=%d n=%d g=%d (need dim %% 16 == 0, 1 <= m <= 31, min(g, m) <= 12, n * 2^g <= 2^24, n * rows < 2^31)", dim, m, n, g);
    CU(fn, D.y.ensure(ybytes));
    CU(fn, D.keys.ensure((size_t)2 * yrows * sizeof(unsigned long long)));
    CU(fn, D.dict.ensure(hdict.size() * 4));
    CU(fn, D.ws.ensure(wsb));
    CU(fn, D.xf.ensure((size_t)8 * 2 * LANE_BYTES));       // device twins of the pinned staging buffers (SPECTAVI_CVT=device)
    CU(fn, cudaMemcpyAsync(D.dict.p, hdict.data(), hdict.size() * 4, cudaMemcpyHostToDevice, D.stream));
    CU(fn, keys_reset((unsigned long long*)D.keys.p, yrows, D.stream));
    if (xn > 0) {
      CU(fn, D.x.ensure((size_t)xn * dim));
      upload_f32_as_u8(fn, D, x + (size_t)x0 * dim, (size_t)xn * dim, (uint8_t*)D.x.p, (float*)D.xf.p, (int)c.gpus.size());   // CascadingHashNn.h:236-239
    }
    tr("cascade: train rows queued", gi);
    upload_f32_as_u8(fn, D, y, ybytes, (uint8_t*)D.y.p, (float*)D.xf.p, (int)c.gpus.size());
    tr("cascade: query rows queued", gi);
    CU(fn, cascade_match((const uint8_t*)D.x.p, (const uint8_t*)D.y.p, xn, yrows, dim, m, n, g, (const float*)D.dict.p, 0,
                         (uint32_t)x0, (unsigned long long*)D.keys.p, D.ws.p, D.ws.cap, D.stream));
  });
  sync_all(fn_, c);        // hdict is about to go away; the collective below starts from quiescent streams
  tr("cascade: K2-K4 done");
  merge_shards(fn_, c, yrows);
  if (ms) {
    deliver_matches(*ms, yrows, true);
    return;
  }
  Gpu& R = c.gpus[0];
  CU(fn_, cudaSetDevice(R.dev));
  const size_t ib = (size_t)yrows * kcols * sizeof(uint64_t), db = (size_t)yrows * kcols * sizeof(float);
  CU(fn_, R.oidx.ensure(ib));
  CU(fn_, R.odist.ensure(db));
  CU(fn_, cudaMemsetAsync(R.oidx.p, 0, ib, R.stream));    // columns >= 2 keep the caller's zeros (Spectavi.cpp:329-334)
  CU(fn_, cudaMemsetAsync(R.odist.p, 0, db, R.stream));
  CU(fn_, keys_finalize_f32((const unsigned long long*)R.keys.p, yrows, kcols, (uint64_t*)R.oidx.p, (float*)R.odist.p, R.stream));
  fetch_results(fn_, c, R, idx, R.oidx.p, ib, dist, R.odist.p, db);
  tr("cascade: results delivered", (long long)ib, (long long)db);
}

// ---- single-GPU entry points: pageable <-> device without the driver's slow paths ----------------------------
// Upload: arrays of a few MB go through the driver's own staging (one call, nothing to coordinate); larger ones are
// cut into one contiguous part per upload lane (thread + copy stream + two pinned buffers), and the compute stream
// waits for every lane.  Download: the DMAs land in pinned memory and copy_out moves them on (a direct device ->
// pageable copy runs at ~2 GB/s, see copy_out).
struct Down { void* h; const void* d; size_t bytes; };
struct Up { void* d; const void* h; size_t bytes; };

// The arrays of `items` are treated as one run of bytes and cut into one part per lane, so that two 10 MB matrices
// keep every lane busy; a part goes up in 1 MB sub-chunks (a single memcpy thread moves 5-8 GB/s on the box: with
// 4 MB sub-chunks the first DMA would only start after ~0.6 ms).
void upload(const char* fn, Gpu& D, std::initializer_list<Up> items) {
  size_t total = 0;
  for (const Up& it : items) total += it.bytes;
  if (total < ((size_t)1 << 20)) {
    for (const Up& it : items)
      if (it.bytes) CU(fn, cudaMemcpyAsync(it.d, it.h, it.bytes, cudaMemcpyHostToDevice, D.stream));
    return;
  }
  lanes_init(fn, D);
  const size_t grain = (size_t)256 << 10, ngrains = (total + grain - 1) / grain, want = (size_t)upload_lanes();
  const int nl = (int)(ngrains / 2 < want ? (ngrains / 2 ? ngrains / 2 : 1) : want);   // at least 512 KB per lane
  const size_t sub = total / nl >= 2 * LANE_BYTES ? LANE_BYTES : ((size_t)1 << 20);
  const std::vector<Up> v(items);
  const Trace tr;
  auto work = [&](int li) {
    cudaSetDevice(D.dev);
    tr("upload lane starts", li);
    const size_t b0 = ngrains * li / nl * grain, b1 = li + 1 == nl ? total : ngrains * (li + 1) / nl * grain;
    size_t pos = 0;                                       // start of the current array in the run
    for (const Up& it : v) {
      const size_t lo = b0 > pos ? b0 : pos, hi = b1 < pos + it.bytes ? b1 : pos + it.bytes;
      if (lo < hi) lane_upload(fn, D.lane[li], static_cast<char*>(it.d) + (lo - pos), static_cast<const char*>(it.h) + (lo - pos), hi - lo, sub);
      pos += it.bytes;
    }
    tr("upload lane queued its last sub-chunk", li, (long long)(b1 - b0));
  };
  std::vector<std::thread> th;
  for (int li = 1; li < nl; ++li) th.emplace_back(work, li);
  work(0);
  for (std::thread& t : th) t.join();
  for (int li = 0; li < nl; ++li) CU(fn, cudaStreamWaitEvent(D.stream, D.lane[li].ready, 0));
  tr("upload joined");
}
void upload(const char* fn, Gpu& D, void* dst, const void* src, size_t bytes) { upload(fn, D, {{dst, src, bytes}}); }

// queues the DMAs behind D.stream's work, waits for the stream, delivers
void download(const char* fn, Gpu& D, std::initializer_list<Down> items) {
  size_t total = 0;
  for (const Down& it : items) total += (it.bytes + 255) & ~(size_t)255;
  if (!total) { CU(fn, cudaStreamSynchronize(D.stream)); return; }
  CU(fn, D.res.ensure(total));
  char* pin = static_cast<char*>(D.res.p);
  size_t o = 0;
  for (const Down& it : items) {
    if (it.bytes) CU(fn, cudaMemcpyAsync(pin + o, it.d, it.bytes, cudaMemcpyDeviceToHost, D.stream));
    o += (it.bytes + 255) & ~(size_t)255;
  }
  CU(fn, cudaStreamSynchronize(D.stream));
  o = 0;
  for (const Down& it : items) {
    if (it.bytes) copy_out(it.h, pin + o, it.bytes);
    o += (it.bytes + 255) & ~(size_t)255;
  }
}

int HostCall::ransac(const double* x0, const double* x1, int npt, double required, double err_allowed, bool find_best,
                     double sv_ratio, const double* F, int nF, bool stop_at_success, int32_t* ok, int32_t* ninlier,
                     double* cams, uint32_t* masks) {
  // single GPU: hypotheses are independent, so a multi-GPU split would be trivial, but the matcher's callers
  // (example/ex01:134-152) stop at the first success, which a few thousand hypotheses usually reach
  Gpu& D = ctx_->gpus[0];
  CU(fn_, cudaSetDevice(D.dev));
  const int W = (npt + 31) / 32;
  const int B = nF < 16384 ? nF : 16384;                       // hypotheses per launch
  const size_t pb = (size_t)npt * 3 * 8, wsb = ransac_ws_bytes(B, npt);
  CU(fn_, D.x.ensure(pb));
  CU(fn_, D.y.ensure(pb));
  CU(fn_, D.xf.ensure((size_t)B * 9 * 8));
  CU(fn_, D.ws.ensure(wsb));
  CU(fn_, D.oidx.ensure((size_t)B * 8));                       // ok, ninlier
  CU(fn_, D.odist.ensure((size_t)B * 12 * 8));
  if (masks) CU(fn_, D.gather.ensure((size_t)B * W * 4));
  upload(fn_, D, {{D.x.p, x0, pb}, {D.y.p, x1, pb}});
  int done = 0;
  while (done < nF) {
    const int nb = nF - done < B ? nF - done : B;
    int32_t* dok = (int32_t*)D.oidx.p;
    int32_t* dcnt = dok + B;
    upload(fn_, D, D.xf.p, F + (size_t)done * 9, (size_t)nb * 9 * 8);
    CU(fn_, ransac_score((const double*)D.xf.p, nb, (const double*)D.x.p, (const double*)D.y.p, npt, required, err_allowed,
                         find_best, sv_ratio, dok, dcnt, (double*)D.odist.p, masks ? (uint32_t*)D.gather.p : nullptr, D.ws.p,
                         D.ws.cap, D.stream));
    download(fn_, D, {{ok + done, dok, (size_t)nb * 4}, {ninlier + done, dcnt, (size_t)nb * 4},
                      {cams + (size_t)done * 12, D.odist.p, (size_t)nb * 12 * 8},
                      {masks ? masks + (size_t)done * W : nullptr, D.gather.p, masks ? (size_t)nb * W * 4 : 0}});
    done += nb;
    if (stop_at_success) {
      bool hit = false;
      for (int i = done - nb; i < done && !hit; ++i) hit = ok[i] && (double)ninlier[i] / (double)npt > required;
      if (hit) break;
    }
  }
  for (int i = done; i < nF; ++i) {
    ok[i] = 0; ninlier[i] = 0;
    for (int k = 0; k < 12; ++k) cams[(size_t)i * 12 + k] = 0.0;
    if (masks) for (int w = 0; w < W; ++w) masks[(size_t)i * W + w] = 0u;
  }
  return done;
}

void HostCall::normalize(const float* x, int rows, int dim, float* out) {
  Gpu& D = ctx_->gpus[0];
  CU(fn_, cudaSetDevice(D.dev));
  const size_t ib = (size_t)rows * dim * 4, ob = (size_t)rows * normalize_new_dim(dim) * 4;
  CU(fn_, D.xf.ensure(ib));
  CU(fn_, D.yf.ensure(ob));
  CU(fn_, D.rws.ensure(normalize_ws_bytes(dim)));
  upload(fn_, D, D.xf.p, x, ib);
  CU(fn_, normalize_rows((const float*)D.xf.p, rows, dim, (float*)D.yf.p, nullptr, D.rws.p, D.rws.cap, D.stream));
  download(fn_, D, {{out, D.yf.p, ob}});
}

// K1f on one GPU: every train row, the query rows [0, yrows) of `y`, results to idx / dist (host)
static void lpk_on(const char* fn, Gpu& D, const void* x, const void* y, bool is_int, int xrows, int yrows, int dim, int k,
                   double p, float mu, uint64_t* idx, void* dist) {
  int sms = 0;
  CU(fn, cudaSetDevice(D.dev));
  CU(fn, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, D.dev));
  const size_t xb = (size_t)xrows * dim * 4, yb = (size_t)yrows * dim * 4;
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  CU(fn, D.x.ensure(xb ? xb : 4));
  CU(fn, D.y.ensure(yb ? yb : 4));
  const bool approx = is_int ? ((int)mu != 0) : (mu != 0.f);      // Scalar m_mu: truncated for the int instantiation
  CU(fn, D.gather.ensure(approx ? lp_mu_ws_bytes(yrows, k) : lp_ws_bytes(xrows, yrows, dim, k, sms)));
  CU(fn, D.oidx.ensure(ib));
  CU(fn, D.odist.ensure(db));
  Trace tr;
  tr("lpk buffers ready");
  upload(fn, D, {{D.x.p, x, xb}, {D.y.p, y, yb}});
  if (tr.on) { cudaStreamSynchronize(D.stream); tr("lpk uploads done (traced runs wait here)", (long long)xb, (long long)yb); }
  if (approx)
    CU(fn, lp_topk_mu(D.x.p, D.y.p, is_int, xrows, yrows, dim, k, p, mu, D.gather.p, D.gather.cap, (uint64_t*)D.oidx.p,
                      D.odist.p, D.stream));
  else
    CU(fn, lp_topk(D.x.p, D.y.p, is_int, xrows, yrows, dim, k, p, D.gather.p, D.gather.cap, (uint64_t*)D.oidx.p, D.odist.p,
                   D.stream));
  if (tr.on) { cudaStreamSynchronize(D.stream); tr("lpk kernels done"); }
  download(fn, D, {{idx, D.oidx.p, ib}, {dist, D.odist.p, db}});
  tr("lpk results delivered", (long long)ib, (long long)db);
}

void HostCall::lpk(const void* x, const void* y, bool is_int, int xrows, int yrows, int dim, int k, double p,
                   uint64_t* idx, void* dist, float mu) {
  // Query rows are independent (BruteForceNn.h:51-119 loops over them), so with SPECTAVI_GPUS > 1 every GPU takes a
  // slice of the queries against ALL train rows: no merge, no collective.  Each GPU needs its own copy of x, so the
  // split only pays when a slice's compute outweighs that upload -- at least 8192 query rows per GPU
  // (test/test_feature.py:67-81's shapes stay on one GPU).
  const int G = (int)ctx_->gpus.size();
  int use = yrows / 8192;
  use = use < 1 ? 1 : (use > G ? G : use);
  if (use == 1) {
    lpk_on(fn_, ctx_->gpus[0], x, y, is_int, xrows, yrows, dim, k, p, mu, idx, dist);
    return;
  }
  const int per = (yrows + use - 1) / use;
  const char* fn = fn_;
  std::vector<std::thread> th;
  auto slice = [&](int g) {
    const int y0 = g * per < yrows ? g * per : yrows, yn = y0 + per <= yrows ? per : yrows - y0;
    if (yn <= 0) return;
    lpk_on(fn, ctx_->gpus[g], x, static_cast<const char*>(y) + (size_t)y0 * dim * 4, is_int, xrows, yn, dim, k, p, mu,
           idx + (size_t)y0 * k, static_cast<char*>(dist) + (size_t)y0 * k * 4);
  };
  for (int g = 1; g < use; ++g) th.emplace_back(slice, g);
  slice(0);
  for (std::thread& t : th) t.join();
}

// ---- the filter overload and its caller --------------------------------
// One GPU, like lpk.  filt_begin uploads the rows and the caller's in-out arrays once; filt_run runs
// any number of groups in ONE launch (their query lists must be disjoint); filt_end downloads.
void HostCall::filt_begin(const float* x, const float* y, int xrows, int yrows, int dim, int k, const uint64_t* idx,
                          const float* dist) {
  Gpu& D = ctx_->gpus[0];
  CU(fn_, cudaSetDevice(D.dev));
  const size_t xb = (size_t)xrows * dim * 4, yb = (size_t)yrows * dim * 4;
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  CU(fn_, D.x.ensure(xb ? xb : 4));
  CU(fn_, D.y.ensure(yb ? yb : 4));
  CU(fn_, D.oidx.ensure(ib ? ib : 8));
  CU(fn_, D.odist.ensure(db ? db : 4));
  upload(fn_, D, {{D.x.p, x, xb}, {D.y.p, y, yb}, {D.oidx.p, idx, ib}, {D.odist.p, dist, db}});
}

void HostCall::filt_run(const std::vector<FilterGroup>& groups, int xrows, int yrows, int dim, int k, double p) {
  Gpu& D = ctx_->gpus[0];
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
  Gpu& D = ctx_->gpus[0];
  const size_t ib = (size_t)yrows * k * sizeof(uint64_t), db = (size_t)yrows * k * 4;
  download(fn_, D, {{idx, D.oidx.p, ib}, {dist, D.odist.p, db}});
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
  std::vector<FilterGroup> groups(1);
  groups[0].xs = as_set(fn_, filter_x, nfx, xrows, "filter_x");
  groups[0].ys = as_set(fn_, filter_y, nfy, yrows, "filter_y");
  filt_begin(x, y, xrows, yrows, dim, k, idx, dist);
  filt_run(groups, xrows, yrows, dim, k, p);
  filt_end(yrows, k, idx, dist);
}

void HostCall::kmedians_nn(const float* x, const float* y, int xrows, int yrows, int dim, const float* medians_x,
                           const int32_t* labels_x, int nmx, const float* medians_y, const int32_t* labels_y, int nmy, int c,
                           int k, uint64_t* idx, float* dist) {
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
}

}  // namespace spv
