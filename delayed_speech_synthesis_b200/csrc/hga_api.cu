// hga_api.cu — the C ABI of libhga_b200.so (include/hga_b200.h) over the sm_100a kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -shared -Xcompiler -fPIC (see __graft_entry__.py)
#include "../../include/hga_b200.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <complex>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "hga_common.cuh"
#include "hga_kernels_exact.cuh"
#include "hga_kernels_fused.cuh"
#include "hga_kernels_pre.cuh"
#include "hga_lane_host.cuh"

using namespace hga;

// ------------------------------------------------------------------------------------------------ errors / misc
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CU(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) return fail(HGA_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                       __FILE__, __LINE__);                                             \
  } while (0)
#define LAUNCHED()                                                                                       \
  do {                                                                                                   \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                                  \
    cudaError_t e_ = cudaGetLastError();                                                                 \
    if (e_ != cudaSuccess) return fail(HGA_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), \
                                       __FILE__, __LINE__);                                              \
  } while (0)

extern "C" const char* hga_last_error(void) { return g_err.c_str(); }
extern "C" int hga_abi_version(void) { return HGA_ABI_VERSION; }
extern "C" int64_t hga_launch_count(void) { return g_launches.load(); }
extern "C" int hga_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// Growable device buffer.
// Guard bands (test instrumentation, hga_debug_guards): when enabled, every device buffer the library allocates gets
// kGuardBytes of a fill pattern in front of and behind the bytes asked for; hga_debug_check_guards reads them back.
// compute-sanitizer is not available on every pool; an overrun of the library's own scratch shows up here instead.
constexpr size_t kGuardBytes = 64 << 10;
constexpr int kGuardFill = 0xA5;
static std::atomic<int> g_guards{0};
struct DevBuf;
static std::mutex g_guard_mu;
static std::vector<DevBuf*>& guard_registry() {
  static std::vector<DevBuf*> r;
  return r;
}
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  void* base = nullptr;   // what cudaMalloc returned (== p without guard bands)
  bool guarded = false;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { unregister(); }
  void unregister() {
    if (!guarded) return;
    std::lock_guard<std::mutex> lk(g_guard_mu);
    auto& r = guard_registry();
    r.erase(std::remove(r.begin(), r.end(), this), r.end());
    guarded = false;
  }
  int reserve(size_t bytes) {
    if (bytes <= cap) return HGA_OK;
    release();
    const bool guard = g_guards.load(std::memory_order_relaxed) != 0;
    const size_t total = bytes + (guard ? 2 * kGuardBytes : 0);
    cudaError_t e = cudaMalloc(&base, total);
    if (e != cudaSuccess) {
      cudaGetLastError();
      base = nullptr;
      return fail(HGA_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", total, cudaGetErrorString(e));
    }
    p = base;
    if (guard) {
      p = static_cast<char*>(base) + kGuardBytes;
      cudaMemset(base, kGuardFill, kGuardBytes);
      cudaMemset(static_cast<char*>(p) + bytes, kGuardFill, kGuardBytes);
      std::lock_guard<std::mutex> lk(g_guard_mu);
      guard_registry().push_back(this);
      guarded = true;
    }
    cap = bytes;
    return HGA_OK;
  }
  void release() {
    unregister();
    if (base) cudaFree(base);
    base = p = nullptr;
    cap = 0;
  }
  template <class T>
  T* as() const { return static_cast<T*>(p); }
};
static void swap_bufs(DevBuf& a, DevBuf& b) {   // the registry keeps pointing at a and b: both stay registered or neither
  std::swap(a.p, b.p);
  std::swap(a.cap, b.cap);
  std::swap(a.base, b.base);
  if (a.guarded != b.guarded) {
    std::lock_guard<std::mutex> lk(g_guard_mu);
    auto& r = guard_registry();
    DevBuf* was = a.guarded ? &a : &b;
    DevBuf* now = a.guarded ? &b : &a;
    std::replace(r.begin(), r.end(), was, now);
    std::swap(a.guarded, b.guarded);
  }
}
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return HGA_OK;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocMapped | cudaHostAllocPortable);  // mapped: the zero-copy packet path reads / writes it from the kernel
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(HGA_ERR_NOMEM, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    }
    cap = bytes;
    return HGA_OK;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  template <class T>
  T* as() const { return static_cast<T*>(p); }
  template <class T>
  T* dev() const {  // the same buffer as the current device sees it
    void* d = nullptr;
    if (!p || cudaHostGetDevicePointer(&d, p, 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return static_cast<T*>(d);
  }
};
// Packets of at most this many rows go through the zero-copy path: the kernel reads the packet from mapped pinned host
// memory (staged into shared memory) and writes the frames to mapped pinned host memory: one launch + one poll of per-warp
// completion slots per packet instead of H2D copy + launch + D2H copy + synchronisation.
constexpr int64_t kMappedRows = 64;

// ------------------------------------------------------------------------------------------------ cascade set-up
// Fill a Cascade from one or two scipy SOS matrices (rows [b0 b1 b2 a0 a1 a2], a0 == 1).
static int build_cascade(Cascade& q, const double* sos_a, int na, const double* zi_a, const double* sos_b, int nb,
                         const double* zi_b) {
  if (na < 0 || nb < 0 || na + nb < 1 || na + nb > kMaxSec) return fail(HGA_ERR_INVALID, "bad section count %d+%d", na, nb);
  memset(&q, 0, sizeof q);
  q.ns = na + nb;
  q.ns_a = na;
  q.unit = 1;
  for (int s = 0; s < q.ns; ++s) {
    const double* r = s < na ? sos_a + 6 * s : sos_b + 6 * (s - na);
    const double* z = s < na ? (zi_a ? zi_a + 2 * s : nullptr) : (zi_b ? zi_b + 2 * (s - na) : nullptr);
    if (r[3] != 1.0) return fail(HGA_ERR_INVALID, "section %d: a0 must be 1 (got %g)", s, r[3]);
    q.b0[s] = r[0]; q.b1[s] = r[1]; q.b2[s] = r[2]; q.a1[s] = r[4]; q.a2[s] = r[5];
    q.zi0[s] = z ? z[0] : 0.0;
    q.zi1[s] = z ? z[1] : 0.0;
    if (r[0] == r[2] && r[0] != 0.0) {
      q.g[s] = r[0];
      q.b1u[s] = r[1] / r[0];
      if (r[0] != 1.0) q.has_gain |= 1u << s;
    } else {
      q.unit = 0;
    }
  }
  for (int s = q.ns; s < kMaxSec; ++s) q.g[s] = 1.0;
  return HGA_OK;
}
// "standard" layout the fully specialised kernels assume: 8 + 8 sections, gains only at sections 0 and 8.
static bool is_unit_std(const Cascade& q) {
  return q.unit && q.ns == 16 && q.ns_a == 8 && (q.has_gain & ~0x101u) == 0;
}
// Largest pole radius of the cascade (roots of z^2 + a1 z + a2).
static double max_pole_radius(const Cascade& q) {
  double r = 0.0;
  for (int s = 0; s < q.ns; ++s) {
    const double disc = q.a1[s] * q.a1[s] - 4.0 * q.a2[s];
    if (disc < 0) r = std::max(r, std::sqrt(std::fabs(q.a2[s])));
    else r = std::max(r, (std::fabs(q.a1[s]) + std::sqrt(disc)) / 2.0);
  }
  return r;
}

// ------------------------------------------------------------------------------------------------ lane kernels, host side
static LaneEncodeFn encode_tiled_fn() {
  static LaneEncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return (LaneEncodeFn)p;
  }();
  return fn;
}
static int upload_log_table() {
  double tab[2 * kLogTabN];
  fill_log_table(tab);
  CU(cudaMemcpyToSymbol(g_logtab, tab, sizeof tab));
  return HGA_OK;
}
// The lane kernels start a segment from the fresh-extractor state in their own section basis, which they derive from the
// meaning of sosfilt_zi (steady state of a unit step): check that the caller's zi IS that state of the caller's band-pass
// (one step with x = 1 must leave it unchanged), else the generic kernels run with the caller's zi as given.
static bool zi_is_step_steady_state(const Cascade& q) {
  double x = 1.0;
  for (int s = 0; s < q.ns_a; ++s) {
    const double y = q.b0[s] * x + q.zi0[s];
    const double n0 = (q.b1[s] * x - q.a1[s] * y) + q.zi1[s], n1 = q.b2[s] * x - q.a2[s] * y;
    const double scale = std::max({1e-300, std::fabs(q.zi0[s]), std::fabs(q.zi1[s]), std::fabs(x)});
    if (std::fabs(n0 - q.zi0[s]) > 1e-9 * scale || std::fabs(n1 - q.zi1[s]) > 1e-9 * scale) return false;
    x = y;
  }
  return true;
}
// x must be 16-byte aligned with an even leading dimension for the TMA descriptor
static bool tma_layout_ok(const double* x, int64_t ldx) { return ((uintptr_t)x % 16) == 0 && (ldx % 2) == 0; }

// Kernel shapes (tools/microbench/lane_bench.cu, profiles/): two warps per (chunk, channel group) item and two items per
// SM for the float64 and the packed float32 arithmetic; the mixed arithmetic keeps every warp on both pipes (one warp per
// item, four per SM).
struct LaneShape {
  int cw;        // channels per item
  int workers;   // items an SM runs at a time
};
static LaneShape lane_shape(int precision) {
  if (precision == HGA_PREC_F64) return {32, 2};
  if (precision == HGA_PREC_F32) return {64, 2};
  return {64, 4};
}
// the fix-up kernels alone (after partial launches with LaneDev::skip_fix)
static int lane_fix_only(int precision, int hop, const LaneConst& k, const LanePlan& p, const LaneDev& d, int C, const double* mu, const double* sigma,
                         double* out, int64_t ldo, cudaStream_t st) {
  int n;
  if (precision == HGA_PREC_F32) n = hop == 10 ? lane_launch_fix<float, 10>(k, p, d, C, mu, sigma, out, ldo, st) : lane_launch_fix<float, 20>(k, p, d, C, mu, sigma, out, ldo, st);
  else n = hop == 10 ? lane_launch_fix<double, 10>(k, p, d, C, mu, sigma, out, ldo, st) : lane_launch_fix<double, 20>(k, p, d, C, mu, sigma, out, ldo, st);
  if (cudaGetLastError() != cudaSuccess) return fail(HGA_ERR_CUDA, "lane fix-up launch failed");
  g_launches.fetch_add(n, std::memory_order_relaxed);
  return HGA_OK;
}
static int lane_launch(int precision, int hop, const LaneConst& k, const LanePlan& p, const LaneDev& d, const CUtensorMap& tm, int C, const double* mu,
                       const double* sigma, double* out, int64_t ldo, int n_sm, cudaStream_t st, bool stats = false) {
  int n = -1;
  if (stats && precision == HGA_PREC_F32) {        // the same kernels, leaving (n, mean, M2) partials of the frames they emit
    // (packed float32 only: there the float64 pipe is idle and the accumulation is free - K2: 1.289 ms against 1.268 ms without
    // statistics and 1.493 ms with the separate pass; in the float64 kernels it competes with the filter for the same pipe and
    // costs more than the second read it saves, tools/stats_ab.py)
    n = hop == 10 ? lanes_launch_t<LaneF32, 10, 60, 2, 2, 2, 3, 5, false, true>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st)
                  : lanes_launch_t<LaneF32, 20, 60, 2, 2, 2, 3, 5, false, true>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st);
  } else if (precision == HGA_PREC_F64) {
    n = hop == 10 ? lanes_launch_t<LaneF64, 10, 80, 2, 2, 2, 3, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st)
                  : lanes_launch_t<LaneF64, 20, 80, 2, 2, 2, 3, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st);
  } else if (precision == HGA_PREC_F32) {
    n = hop == 10 ? lanes_launch_t<LaneF32, 10, 60, 2, 2, 2, 3, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st)
                  : lanes_launch_t<LaneF32, 20, 60, 2, 2, 2, 3, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st);
  } else {
    n = hop == 10 ? lane_launch_t<LaneMix, 10, 50, 4, 2, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st)
                  : lane_launch_t<LaneMix, 20, 40, 4, 2, 5>(k, p, d, tm, C, mu, sigma, out, ldo, n_sm, st);
  }
  if (n < 0) return fail(HGA_ERR_CUDA, "lane kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
  g_launches.fetch_add(n, std::memory_order_relaxed);
  return HGA_OK;
}
static int lane_tile_rows(int precision, int hop) { return precision == HGA_PREC_F64 ? 80 : precision == HGA_PREC_F32 ? 60 : (hop == 10 ? 50 : 40); }

// ------------------------------------------------------------------------------------------------ the handle
struct hga_handle {
  hga_config cfg;
  Cascade q;
  // lane kernels (offline extraction of fresh segments): operands, fix-up pieces, cached plan and its scratch
  bool lane_ok = false;
  LaneConst lk;
  int lane_nsub = 0;
  DevBuf lane_subs, lane_P;
  LanePlan lane_plan_host;
  std::vector<int64_t> lane_key;
  DevBuf lane_work, lane_fix, lane_segs, lane_order, lane_counter, lane_y, lane_bs, lane_E, lane_S0, lane_M, lane_st;
  bool lane_use_M = false;
  int lane_window = 0;   // > 0: windowed parallel carry (lane_carry_window), else the per-segment chain
  int n_sm = 0;
  int warmup = 0;
  int overlap = 0;   // L - hop
  int nb = 0;        // L / hop when L % hop == 0, else 0 (fused kernels unavailable)
  int cap_rows = 0;  // rows of the streaming frame buffer
  // streaming state (device)
  DevBuf state, state2;        // (ns, 2, C); state2 = ping-pong partner for bulk pushes
  DevBuf framebuf, framebuf2;  // (cap_rows, C): [remainder | new rows]
  int first_frame = 1;
  // scratch
  DevBuf work, segs, ybuf, stats_part, seg_state, seg_rem;
  std::vector<Work> work_host;
  std::vector<Segment> segs_host;
  std::vector<int64_t> plan_key;  // seg_start/len/out_row of the cached plan
  int plan_variant = -1;
  // host-buffer paths
  DevBuf xdev, outdev, mudev, sigdev, rawdev, xpad;
  PinBuf pin_in, pin_out, pin_flag;   // pin_flag: per-warp completion slots of the zero-copy packet path
  PinBuf stage_up[2], stage_down;     // offline host path: staging of pageable caller buffers
  cudaEvent_t stage_ev[2] = {nullptr, nullptr};
  unsigned long long push_seq = 0;
  bool push_inflight = false;         // a host push returned on its completion slots; its kernel may still be storing state
  cudaEvent_t dev_push_ev = nullptr;  // recorded on the caller's stream behind every device push (hga_stream_push)
  bool dev_push_pending = false;      // ... and not yet waited for by the handle's own stream
  std::vector<double> mu_cache, sig_cache;
  bool musig_valid = false;
  cudaStream_t s_own = nullptr, s_copy = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_main = nullptr;
  bool timed_main = false;
  std::vector<cudaEvent_t> slab_ev;
  int timing = 0;
  bool timed = false;
};

// Host pushes that poll completion slots return before their kernel has stored the new remainder / filter state on the
// handle's own stream; anything that touches that state from another stream (or frees it) waits for it first.
// The streaming state lives on the device and is updated by whichever stream the push ran on: a device push on the
// caller's stream, host / packet pushes on s_own.  Work on s_own is ordered behind the last device push.
static void order_after_dev_push(hga_handle* h) {
  if (h->dev_push_pending) {
    cudaStreamWaitEvent(h->s_own, h->dev_push_ev, 0);
    h->dev_push_pending = false;
  }
}
static void drain_push(hga_handle* h) {
  if (h->push_inflight) {
    cudaStreamSynchronize(h->s_own);
    h->push_inflight = false;
  }
}
static int tile_zi(hga_handle* h, cudaStream_t st) {
  const int C = h->cfg.n_channels, ns = h->q.ns;
  std::vector<double> z((size_t)ns * 2 * C);
  for (int s = 0; s < ns; ++s)
    for (int c = 0; c < C; ++c) {
      z[((size_t)2 * s) * C + c] = h->q.zi0[s];
      z[((size_t)2 * s + 1) * C + c] = h->q.zi1[s];
    }
  CU(cudaMemcpyAsync(h->state.p, z.data(), z.size() * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(h->framebuf.p, 0, (size_t)h->cap_rows * C * 8, st));
  CU(cudaStreamSynchronize(st));  // z is a stack-owned vector
  h->first_frame = 1;
  return HGA_OK;
}

extern "C" int hga_create(const hga_config* cfg, hga_handle** out) {
  if (!cfg || !out) return fail(HGA_ERR_INVALID, "null argument");
  *out = nullptr;
  if (cfg->abi_version != HGA_ABI_VERSION) return fail(HGA_ERR_INVALID, "ABI version %d != %d", cfg->abi_version, HGA_ABI_VERSION);
  if (cfg->n_channels < 1 || cfg->frame_len < 1 || cfg->hop < 1 || cfg->frame_len < cfg->hop || cfg->fs < 1)
    return fail(HGA_ERR_INVALID, "bad geometry: C=%d L=%d hop=%d fs=%d", cfg->n_channels, cfg->frame_len, cfg->hop, cfg->fs);
  if (cfg->n_sections_bp < 1 || cfg->n_sections_bp > HGA_MAX_SECTIONS || cfg->n_sections_notch < 0 ||
      cfg->n_sections_notch > HGA_MAX_SECTIONS)
    return fail(HGA_ERR_INVALID, "bad section counts %d/%d", cfg->n_sections_bp, cfg->n_sections_notch);
  if (cfg->precision < HGA_PREC_EXACT || cfg->precision > HGA_PREC_MIXED) return fail(HGA_ERR_INVALID, "bad precision %d", cfg->precision);
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= hga_device_count()) return fail(HGA_ERR_INVALID, "bad device %d", cfg->device);
  CU(cudaSetDevice(cfg->device));
  hga_handle* h = new hga_handle();
  h->cfg = *cfg;
  int rc = build_cascade(h->q, cfg->sos_bp, cfg->n_sections_bp, cfg->zi_bp, cfg->sos_notch, cfg->n_sections_notch, cfg->zi_notch);
  if (rc) { delete h; return rc; }
  if ((rc = upload_log_table())) { delete h; return rc; }
  cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, cfg->device);
  h->overlap = cfg->frame_len - cfg->hop;
  h->nb = (cfg->frame_len % cfg->hop == 0 && cfg->frame_len / cfg->hop <= kMaxBlocksPerFrame) ? cfg->frame_len / cfg->hop : 0;
  const double r = max_pole_radius(h->q);
  if (r >= 1.0) { delete h; return fail(HGA_ERR_INVALID, "unstable filter: pole radius %g", r); }
  if (cfg->warmup > 0) h->warmup = cfg->warmup;
  else {
    // Only the generic kernel (k_hga_fused_seq: non-standard cascades, odd leading dimensions) still re-filters history from
    // zero state.  The dropped zero-input response scales with the filter STATE, which follows out-of-band content (offsets,
    // drifts) and resonates between neighbouring poles, hence the 1e-4 margin on the requested tolerance.
    const double tol = (cfg->warmup_tol > 0 ? cfg->warmup_tol : 1e-12) * 1e-4;
    h->warmup = r > 0 ? (int)std::ceil(std::log(tol) / std::log(r)) : 1;
  }
  h->warmup = std::max(h->warmup, h->overlap);
  h->warmup = (h->warmup + cfg->hop - 1) / cfg->hop * cfg->hop;
  // lane kernels: the standard cascade, L = 5 hops, hop 10 or 20, the caller's zi = sosfilt_zi
  h->lane_ok = (cfg->precision == HGA_PREC_F64 || cfg->precision == HGA_PREC_F32 || cfg->precision == HGA_PREC_MIXED) && lane_cascade_ok(h->q) &&
               h->nb == kLaneNB && (cfg->hop == 10 || cfg->hop == 20) && zi_is_step_steady_state(h->q) && encode_tiled_fn() != nullptr;
  if (h->lane_ok) {
    const double tol = cfg->warmup_tol > 0 ? cfg->warmup_tol : (cfg->precision == HGA_PREC_F64 ? 1e-11 : cfg->precision == HGA_PREC_F32 ? 1e-6 : 1e-7);
    lane_build_const(h->q, cfg->frame_len, cfg->hop, cfg->eps, tol, h->lk);
    LaneFixPlan fp;
    lane_build_fix(h->lk, cfg->hop, fp);
    h->lane_nsub = (int)fp.subs.size();
    if ((rc = h->lane_subs.reserve(fp.subs.size() * sizeof(LaneSub))) || (rc = h->lane_P.reserve(fp.P.size() * 8))) { hga_destroy(h); return rc; }
    CU(cudaMemcpy(h->lane_subs.p, fp.subs.data(), fp.subs.size() * sizeof(LaneSub), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->lane_P.p, fp.P.data(), fp.P.size() * 8, cudaMemcpyHostToDevice));
  }
  const int max_push = cfg->max_push_rows > 0 ? cfg->max_push_rows : 4096;
  h->cap_rows = std::max(cfg->frame_len, h->overlap + max_push);
  const int C = cfg->n_channels;
  if ((rc = h->state.reserve((size_t)h->q.ns * 2 * C * 8)) || (rc = h->framebuf.reserve((size_t)h->cap_rows * C * 8)) ||
      (rc = h->state2.reserve((size_t)h->q.ns * 2 * C * 8)) || (rc = h->framebuf2.reserve((size_t)h->cap_rows * C * 8))) {
    hga_destroy(h);
    return rc;
  }
  cudaStreamCreateWithFlags(&h->s_own, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking);
  cudaEventCreate(&h->ev0);
  cudaEventCreate(&h->ev1);
  cudaEventCreate(&h->ev_main);
  if ((rc = tile_zi(h, h->s_own))) { hga_destroy(h); return rc; }
  *out = h;
  return HGA_OK;
}

extern "C" int hga_destroy(hga_handle* h) {
  if (!h) return HGA_OK;
  cudaSetDevice(h->cfg.device);
  drain_push(h);
  for (DevBuf* b : {&h->state, &h->framebuf, &h->state2, &h->framebuf2, &h->work, &h->segs, &h->ybuf, &h->stats_part, &h->seg_state, &h->seg_rem, &h->lane_subs, &h->lane_P, &h->lane_work, &h->lane_fix, &h->lane_segs, &h->lane_order, &h->lane_counter,
                    &h->lane_y, &h->lane_bs, &h->lane_E, &h->lane_S0, &h->lane_M, &h->lane_st,
                    &h->xdev, &h->outdev, &h->mudev, &h->sigdev, &h->rawdev, &h->xpad})
    b->release();
  h->pin_in.release();
  h->pin_out.release();
  h->pin_flag.release();
  h->stage_up[0].release();
  h->stage_up[1].release();
  h->stage_down.release();
  for (cudaEvent_t e : h->stage_ev) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : h->slab_ev) cudaEventDestroy(e);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev_main) cudaEventDestroy(h->ev_main);
  if (h->dev_push_ev) cudaEventDestroy(h->dev_push_ev);
  if (h->s_own) cudaStreamDestroy(h->s_own);
  if (h->s_copy) cudaStreamDestroy(h->s_copy);
  delete h;
  return HGA_OK;
}

extern "C" int hga_reset(hga_handle* h, void* stream) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  CU(cudaSetDevice(h->cfg.device));
  drain_push(h);
  return tile_zi(h, (cudaStream_t)stream);
}
extern "C" int hga_warmup(const hga_handle* h) { return h ? (h->lane_ok ? lane_carry_rows(h->lk) : h->warmup) : 0; }
extern "C" int hga_last_plan(const hga_handle* h, int64_t* chunk_rows, int32_t* n_chunks, int32_t* n_carried, int32_t* carry_rows) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  const bool lane = h->lane_ok && !h->lane_key.empty();
  if (chunk_rows) *chunk_rows = lane ? h->lane_plan_host.Lc : 0;
  if (n_chunks) *n_chunks = lane ? (int32_t)h->lane_plan_host.work.size() : (int32_t)h->work_host.size();
  if (n_carried) *n_carried = lane ? (int32_t)h->lane_plan_host.fix.size() : 0;
  if (carry_rows) *carry_rows = hga_warmup(h);
  return HGA_OK;
}
extern "C" int hga_set_timing(hga_handle* h, int32_t enable) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  h->timing = enable;
  h->timed = false;
  h->timed_main = false;
  return HGA_OK;
}
extern "C" double hga_last_kernel_ms(hga_handle* h) {
  if (!h || !h->timed) return -1.0;
  float ms = 0;
  if (cudaEventSynchronize(h->ev1) != cudaSuccess || cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) {
    cudaGetLastError();
    return -1.0;
  }
  return ms;
}
extern "C" int hga_debug_guards(int32_t enable) {
  g_guards.store(enable ? 1 : 0, std::memory_order_relaxed);
  return HGA_OK;
}
extern "C" int64_t hga_debug_check_guards(void) {
  if (cudaDeviceSynchronize() != cudaSuccess) return fail(HGA_ERR_CUDA, "synchronize failed: %s", cudaGetErrorString(cudaGetLastError()));
  std::lock_guard<std::mutex> lk(g_guard_mu);
  std::vector<unsigned char> host(kGuardBytes);
  int64_t bad = 0;
  for (DevBuf* b : guard_registry()) {
    const char* bands[2] = {static_cast<const char*>(b->base), static_cast<const char*>(b->p) + b->cap};
    for (const char* band : bands) {
      if (cudaMemcpy(host.data(), band, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(HGA_ERR_CUDA, "guard read-back failed");
      for (size_t i = 0; i < kGuardBytes; ++i)
        if (host[i] != kGuardFill) { ++bad; break; }
    }
  }
  return bad;   // number of overwritten guard bands
}
extern "C" int64_t hga_debug_guarded_buffers(void) {
  std::lock_guard<std::mutex> lk(g_guard_mu);
  return (int64_t)guard_registry().size();
}
extern "C" int hga_last_kernel_split_ms(hga_handle* h, double* main_ms, double* rest_ms) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  if (!h->timed || !h->timed_main) return fail(HGA_ERR_INVALID, "no timed lane extraction on this handle (hga_set_timing, then a device extraction)");
  float a = 0, b = 0;
  if (cudaEventSynchronize(h->ev1) != cudaSuccess || cudaEventElapsedTime(&a, h->ev0, h->ev_main) != cudaSuccess ||
      cudaEventElapsedTime(&b, h->ev_main, h->ev1) != cudaSuccess) {
    return fail(HGA_ERR_CUDA, "event query failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  if (main_ms) *main_ms = a;
  if (rest_ms) *rest_ms = b;
  return HGA_OK;
}
extern "C" int64_t hga_num_frames(const hga_handle* h, int64_t rows) {
  if (!h || rows < h->cfg.frame_len) return 0;
  return (rows - h->cfg.frame_len) / h->cfg.hop + 1;
}

// ------------------------------------------------------------------------------------------------ offline path
// Build the time-chunk work list.  Cost model: items run `slots` warps at a time, each costing (chunk + warm-up)
// samples; pick the chunk length that minimises waves * (chunk + warmup).
// slots_ctas: CTAs (work item x channel group) the device runs at a time; groups: channel groups per work item.
static void plan_chunks(hga_handle* h, int n_seg, const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                        int64_t slots_ctas, int64_t groups, bool allow_chunking) {
  const int hop = h->cfg.hop, K = h->warmup, nb = h->nb;
  int64_t total = 0, maxT = 0;
  for (int i = 0; i < n_seg; ++i) { total += seg_len[i]; maxT = std::max(maxT, seg_len[i]); }
  int64_t best_Lc = (maxT + hop - 1) / hop * hop;
  if (allow_chunking && slots_ctas > 0) {
    double best = 1e300;
    const int64_t min_Lc = std::max<int64_t>((int64_t)h->cfg.frame_len * 4, 64 * hop);
    for (int64_t n = 1; n <= 1 << 20; n = n < 64 ? n + 1 : n + n / 16) {
      int64_t Lc = ((total + n - 1) / n + hop - 1) / hop * hop;
      if (Lc < min_Lc) break;
      int64_t items = 0;
      for (int i = 0; i < n_seg; ++i) items += (seg_len[i] + Lc - 1) / Lc;
      const int64_t waves = (items * groups + slots_ctas - 1) / slots_ctas;  // e.g. K4 float64: 9 chunks x 32 groups = 288 CTAs = 2 waves
                                                                              // of 148 beat 4 x 32 = 128 CTAs in 1 wave (86 % of the SMs)
      const double per = (double)std::min(Lc, maxT) + (items > n_seg ? (double)K : 0.0);
      const double cost = waves * per;
      if (cost < best * 0.999) { best = cost; best_Lc = Lc; }
    }
  }
  h->work_host.clear();
  h->segs_host.clear();
  for (int i = 0; i < n_seg; ++i) {
    const int64_t T = seg_len[i];
    h->segs_host.push_back(Segment{seg_start[i], T, out_row[i]});
    const int64_t n_items = std::max<int64_t>(1, (T + best_Lc - 1) / best_Lc);
    for (int64_t k = 0; k < n_items; ++k) {
      Work w;
      const int64_t cs = k * best_Lc;
      w.row0 = seg_start[i];
      w.T = T;
      w.t_begin = std::max<int64_t>(0, cs - K);
      w.t_acc = std::max<int64_t>(0, cs - (int64_t)(nb - 1) * hop);
      if (w.t_begin > w.t_acc) w.t_begin = w.t_acc;
      w.last = k == n_items - 1;
      w.t_end = w.last ? T : cs + best_Lc;
      w.out_row0 = out_row[i];
      w.seg = i;
      h->work_host.push_back(w);
    }
  }
}

static int upload_plan(hga_handle* h, cudaStream_t st) {
  int rc;
  if ((rc = h->work.reserve(std::max<size_t>(1, h->work_host.size()) * sizeof(Work)))) return rc;
  if ((rc = h->segs.reserve(std::max<size_t>(1, h->segs_host.size()) * sizeof(Segment)))) return rc;
  if (!h->work_host.empty())
    CU(cudaMemcpyAsync(h->work.p, h->work_host.data(), h->work_host.size() * sizeof(Work), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(h->segs.p, h->segs_host.data(), h->segs_host.size() * sizeof(Segment), cudaMemcpyHostToDevice, st));
  CU(cudaStreamSynchronize(st));  // pageable source: keep it simple and safe; plans are cached across calls
  return HGA_OK;
}

// Which fused kernel: 2 = lane kernels (chunked scan with exact state carry; fresh segments, the standard cascade, TMA-able x),
// 0 = thread-per-(item, channel) sequential cascade (float64 only: any cascade, any layout, carried state), -3 = none.
static int choose_variant(const hga_handle* h, const double* x, int64_t ldx, bool plain) {
  if (h->lane_ok && plain && tma_layout_ok(x, ldx)) return 2;
  return h->cfg.precision == HGA_PREC_F64 ? 0 : -3;
}
static const char* kLaneOnlyMsg = "this precision needs the standard cascade (8 + 8 unit-numerator sections, sosfilt_zi start), L = 5 hops, hop 10 or 20, "
                                  "16-byte-aligned x with an even leading dimension, and no carried state";

// Plan (cached on the segment table) and scratch of the lane kernels.  by_time: additionally build the item list sorted by
// the last x row an item reads (host-buffer path: items become runnable as their rows arrive); it goes to `items_by_time`.
static int lane_prepare(hga_handle* h, int n_seg, const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row, int64_t xlo, cudaStream_t st,
                        std::vector<int>* items_by_time = nullptr) {
  const int C = h->cfg.n_channels, hop = h->cfg.hop;
  const LaneShape sh = lane_shape(h->cfg.precision);
  std::vector<int64_t> key;
  key.reserve(3 * (size_t)n_seg + 1);
  key.insert(key.end(), seg_start, seg_start + n_seg);
  key.insert(key.end(), seg_len, seg_len + n_seg);
  key.insert(key.end(), out_row, out_row + n_seg);
  key.push_back(xlo);
  int rc;
  LanePlan& p = h->lane_plan_host;
  if (key != h->lane_key) {
    h->lane_key.clear();
    lane_plan(h->lk, hop, C, sh.cw, n_seg, seg_start, seg_len, out_row, xlo, (int64_t)h->n_sm * sh.workers, true, p);
    for (const LaneWork& w : p.work)
      if ((int64_t)w.x_row + w.rows > INT32_MAX) return fail(HGA_ERR_UNSUPPORTED, "more than 2^31 rows in one call");
    const size_t CP = (size_t)p.groups * p.cw, nf = p.fix.size();
    const size_t ysz = h->cfg.precision == HGA_PREC_F32 ? 4 : 8;
    if ((rc = h->lane_work.reserve(std::max<size_t>(1, p.work.size()) * sizeof(LaneWork))) || (rc = h->lane_fix.reserve(std::max<size_t>(1, nf) * sizeof(LaneFix))) ||
        (rc = h->lane_segs.reserve(std::max<size_t>(1, p.segs.size()) * sizeof(LaneCarrySeg))) || (rc = h->lane_order.reserve(std::max<size_t>(1, p.order.size()) * 2 * sizeof(int))) ||
        (rc = h->lane_counter.reserve(16)) || (rc = h->lane_y.reserve(std::max<size_t>(16, (size_t)p.y_rows * CP * ysz))) ||
        (rc = h->lane_bs.reserve(std::max<size_t>(16, (size_t)p.bs_rows * CP * 8))) || (rc = h->lane_E.reserve(std::max<size_t>(16, nf * kLaneState * CP * 8))) ||
        (rc = h->lane_M.reserve(2 * kLaneState * kLaneState * 8)))
      return rc;
    // the carry is propagated across a chunk only where it survives it (chunks shorter than the carry's reach)
    std::vector<double> M(kLaneState * kLaneState);
    const double tol = h->cfg.warmup_tol > 0 ? h->cfg.warmup_tol : (h->cfg.precision == HGA_PREC_F64 ? 1e-11 : h->cfg.precision == HGA_PREC_F32 ? 1e-6 : 1e-7);
    h->lane_use_M = nf > 1 && lane_transition(h->lk, p.Lc, M.data()) > 1e-3 * tol;
    h->lane_window = h->lane_use_M ? lane_carry_window(M.data(), 1e-3 * tol) : 0;
    if (h->lane_use_M) {   // M, then its transpose
      M.resize(2 * kLaneState * kLaneState);
      for (int i = 0; i < kLaneState; ++i)
        for (int j = 0; j < kLaneState; ++j) M[kLaneState * kLaneState + j * kLaneState + i] = M[i * kLaneState + j];
    }
    if (h->lane_use_M && (rc = h->lane_S0.reserve(nf * kLaneState * CP * 8))) return rc;
    CU(cudaMemcpyAsync(h->lane_work.p, p.work.data(), p.work.size() * sizeof(LaneWork), cudaMemcpyHostToDevice, st));
    if (nf) CU(cudaMemcpyAsync(h->lane_fix.p, p.fix.data(), nf * sizeof(LaneFix), cudaMemcpyHostToDevice, st));
    if (!p.segs.empty()) CU(cudaMemcpyAsync(h->lane_segs.p, p.segs.data(), p.segs.size() * sizeof(LaneCarrySeg), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->lane_order.p, p.order.data(), p.order.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    if (h->lane_use_M) CU(cudaMemcpyAsync(h->lane_M.p, M.data(), M.size() * 8, cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));  // pageable sources; plans are cached across calls
    h->lane_key = key;
  }
  if (items_by_time) {
    std::vector<int>& v = *items_by_time;
    v.resize(p.order.size());
    for (size_t i = 0; i < v.size(); ++i) v[i] = (int)i;
    std::stable_sort(v.begin(), v.end(), [&](int a, int b) {
      const LaneWork &wa = p.work[a / p.groups], &wb = p.work[b / p.groups];
      return (int64_t)wa.x_row + wa.rows < (int64_t)wb.x_row + wb.rows;
    });
    CU(cudaMemcpyAsync(h->lane_order.as<int>() + p.order.size(), v.data(), v.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
  }
  return HGA_OK;
}
static LaneDev lane_dev(const hga_handle* h) {
  LaneDev d;
  d.work = h->lane_work.as<LaneWork>();
  d.fix = h->lane_fix.as<LaneFix>();
  d.segs = h->lane_segs.as<LaneCarrySeg>();
  d.order = h->lane_order.as<int>();
  d.counter = h->lane_counter.as<int>();
  d.ybuf = h->lane_y.p;
  d.bs = h->lane_bs.as<double>();
  d.E = h->lane_E.as<double>();
  d.S0 = h->lane_S0.as<double>();
  d.M = h->lane_use_M ? h->lane_M.as<double>() : nullptr;
  d.MT = h->lane_use_M ? h->lane_M.as<double>() + kLaneState * kLaneState : nullptr;
  d.carry_window = h->lane_window;
  d.subs = h->lane_subs.as<LaneSub>();
  d.n_sub = h->lane_nsub;
  d.P = h->lane_P.as<double>();
  return d;
}
static int lane_tensor(const hga_handle* h, const double* x, int64_t ldx, int64_t xlo, int64_t xhi, CUtensorMap* tm) {
  const LaneShape sh = lane_shape(h->cfg.precision);
  const CUresult r = lane_tensor_map(encode_tiled_fn(), x + xlo * ldx, ldx, h->cfg.n_channels, xhi - xlo, sh.cw, lane_tile_rows(h->cfg.precision, h->cfg.hop), tm);
  if (r != CUDA_SUCCESS) return fail(HGA_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return HGA_OK;
}

template <int ARITH>
static int launch_fused_seq(hga_handle* h, const FusedArgs& a, cudaStream_t st) {
  const int cg = (a.C + 31) / 32;
  const int64_t warps = (int64_t)a.n_work * cg;
  const unsigned grid = (unsigned)((warps + 3) / 4);
  if (a.hop == 10) k_hga_fused_seq<ARITH, 10><<<grid, 128, 0, st>>>(h->q, a);
  else if (a.hop == 20) k_hga_fused_seq<ARITH, 20><<<grid, 128, 0, st>>>(h->q, a);
  else k_hga_fused_seq<ARITH, 0><<<grid, 128, 0, st>>>(h->q, a);
  LAUNCHED();
  return HGA_OK;
}

static int extract_offline_dev(hga_handle* h, const double* x, int64_t ldx, int n_seg, const int64_t* seg_start,
                               const int64_t* seg_len, const int64_t* out_row, const double* mu, const double* sigma,
                               double* out, int64_t ldo, cudaStream_t st, double* state_out, double* rem_out,
                               const double* state_in = nullptr, const double* rem_in = nullptr, bool allow_short = false) {
  const int C = h->cfg.n_channels, L = h->cfg.frame_len, hop = h->cfg.hop;
  if (n_seg < 1) return HGA_OK;
  if ((mu == nullptr) != (sigma == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  for (int i = 0; i < n_seg; ++i)
    if (seg_len[i] < L && !allow_short) return fail(HGA_ERR_INVALID, "segment %d has %lld rows < frame_len %d", i, (long long)seg_len[i], L);
  int rc;
  const bool fused = h->cfg.precision != HGA_PREC_EXACT && h->nb > 0;
  const bool plain = !state_in && !rem_in && !state_out && !rem_out;
  int variant = fused ? choose_variant(h, x, ldx, plain) : -2;
  int64_t xlo = INT64_MAX, xhi = 0;
  for (int i = 0; i < n_seg; ++i) { xlo = std::min(xlo, seg_start[i]); xhi = std::max(xhi, seg_start[i] + seg_len[i]); }
  if (variant == -3 && h->lane_ok && plain) {
    // packed / mixed arithmetic on a layout the TMA cannot describe (odd leading dimension, unaligned base): one 2-D copy
    // of the rows in use to an aligned scratch, then the lane kernels (float64 takes the generic kernel instead)
    const int64_t ldp = (C + 1) / 2 * 2;
    if ((rc = h->xpad.reserve((size_t)(xhi - xlo) * ldp * 8))) return rc;
    CU(cudaMemcpy2DAsync(h->xpad.p, (size_t)ldp * 8, x + xlo * ldx, (size_t)ldx * 8, (size_t)C * 8, (size_t)(xhi - xlo), cudaMemcpyDeviceToDevice, st));
    x = h->xpad.as<double>() - xlo * ldp;
    ldx = ldp;
    variant = 2;
  }
  if (variant == -3) return fail(HGA_ERR_UNSUPPORTED, "%s", kLaneOnlyMsg);
  if (variant == 2) {
    if ((rc = lane_prepare(h, n_seg, seg_start, seg_len, out_row, xlo, st))) return rc;
    CUtensorMap tm;
    if ((rc = lane_tensor(h, x, ldx, xlo, xhi, &tm))) return rc;
    LaneDev ld = lane_dev(h);
    ld.x0 = x + xlo * ldx;
    ld.ldx = ldx;
    h->timed_main = false;
    if (h->timing) { CU(cudaEventRecord(h->ev0, st)); ld.ev_main = h->ev_main; }
    if ((rc = lane_launch(h->cfg.precision, hop, h->lk, h->lane_plan_host, ld, tm, C, mu, sigma, out, ldo, h->n_sm, st))) return rc;
    if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; h->timed_main = true; }
    return HGA_OK;
  }
  if (variant == 0) {
    // plan (cached on the segment table)
    std::vector<int64_t> key;
    key.reserve(3 * (size_t)n_seg + 1);
    key.insert(key.end(), seg_start, seg_start + n_seg);
    key.insert(key.end(), seg_len, seg_len + n_seg);
    key.insert(key.end(), out_row, out_row + n_seg);
    if (key != h->plan_key || variant != h->plan_variant) {
      plan_chunks(h, n_seg, seg_start, seg_len, out_row, std::max<int64_t>(1, (int64_t)h->n_sm * 12), (C + 31) / 32, true);
      if ((rc = upload_plan(h, st))) return rc;
      h->plan_key = key;
      h->plan_variant = variant;
    }
    FusedArgs a{x, ldx, h->work.as<Work>(), (int)h->work_host.size(), C, L, hop, h->nb, h->cfg.eps, mu, sigma, out, ldo, state_in, rem_in, state_out, rem_out, xlo, xhi};
    if (h->timing) CU(cudaEventRecord(h->ev0, st));
    if (is_unit_std(h->q)) rc = launch_fused_seq<kUnitStd>(h, a, st);
    else if (h->q.unit) rc = launch_fused_seq<kUnit>(h, a, st);
    else rc = launch_fused_seq<kFused>(h, a, st);
    if (rc) return rc;
    if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; }
    return HGA_OK;
  }
  // exact: filtered rows to scratch, then row-order windows.  Every segment gets its own rows of the scratch (laid end to
  // end): trial slices may overlap in x (trials closer than the 40 ms tail, prepare_corpus.py:49), and each is filtered
  // from its own fresh state.
  int64_t total = 0, maxF = 0;
  h->segs_host.clear();
  std::vector<Segment> ysegs((size_t)n_seg);
  for (int i = 0; i < n_seg; ++i) {
    h->segs_host.push_back(Segment{seg_start[i], seg_len[i], out_row[i]});
    ysegs[(size_t)i] = Segment{total, seg_len[i], out_row[i]};
    total += seg_len[i];
    maxF = std::max(maxF, (seg_len[i] - L) / hop + 1);
  }
  if (n_seg > 65535) return fail(HGA_ERR_UNSUPPORTED, "more than 65535 segments per call");
  if ((rc = h->ybuf.reserve((size_t)std::max<int64_t>(1, total) * C * 8)) || (rc = h->segs.reserve(2 * (size_t)n_seg * sizeof(Segment)))) return rc;
  Segment* segs_x = h->segs.as<Segment>();
  Segment* segs_y = segs_x + n_seg;
  CU(cudaMemcpyAsync(segs_x, h->segs_host.data(), (size_t)n_seg * sizeof(Segment), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(segs_y, ysegs.data(), (size_t)n_seg * sizeof(Segment), cudaMemcpyHostToDevice, st));
  CU(cudaStreamSynchronize(st));  // pageable sources
  h->plan_key.clear();            // the generic kernel's plan shares h->segs
  double* y = h->ybuf.as<double>();
  const int cg = (C + 31) / 32;
  const unsigned grid = (unsigned)(((int64_t)n_seg * cg + 3) / 4);
  if (h->timing) CU(cudaEventRecord(h->ev0, st));
  k_iir_seq<kExact><<<grid, 128, 0, st>>>(h->q, x, ldx, y, C, segs_x, Segment{}, n_seg, C, nullptr, state_out, segs_y);
  LAUNCHED();
  if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; }
  if (L == 5 * hop) {
    const int R = roll_run_length(maxF * n_seg, cg, h->n_sm);
    dim3 g2((unsigned)cg, (unsigned)((maxF + (int64_t)R * kRollWarps - 1) / ((int64_t)R * kRollWarps)), (unsigned)n_seg);
    k_frames_roll<5><<<g2, 32 * kRollWarps, 0, st>>>(y, C, segs_y, Segment{}, n_seg, C, hop, R, h->cfg.eps, mu, sigma, out, ldo);
  } else {
    dim3 g2((unsigned)cg, (unsigned)((maxF + kFramesPerBlock - 1) / kFramesPerBlock), (unsigned)n_seg);
    k_frames_seq<<<g2, 256, 0, st>>>(y, C, 0, segs_y, Segment{}, n_seg, C, maxF, L, hop, nullptr, nullptr, h->cfg.eps, mu, sigma, out, ldo);
  }
  LAUNCHED();
  if (rem_out) {
    for (int i = 0; i < n_seg; ++i)
      CU(cudaMemcpy2DAsync(rem_out + (size_t)i * h->overlap * C, (size_t)C * 8, y + (ysegs[(size_t)i].row0 + seg_len[i] - h->overlap) * C,
                           (size_t)C * 8, (size_t)C * 8, h->overlap, cudaMemcpyDeviceToDevice, st));
  }
  return HGA_OK;
}

extern "C" int hga_extract_offline(hga_handle* h, const double* x_dev, int64_t ldx, int32_t n_segments, const int64_t* seg_start,
                                   const int64_t* seg_len, const int64_t* out_row, const double* mu_dev, const double* sigma_dev,
                                   double* out_dev, int64_t ldo, void* stream) {
  if (!h || !x_dev || !out_dev || !seg_start || !seg_len || !out_row) return fail(HGA_ERR_INVALID, "null argument");
  if (ldx < h->cfg.n_channels || ldo < h->cfg.n_channels) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  CU(cudaSetDevice(h->cfg.device));
  return extract_offline_dev(h, x_dev, ldx, n_segments, seg_start, seg_len, out_row, mu_dev, sigma_dev, out_dev, ldo,
                             (cudaStream_t)stream, nullptr, nullptr);
}

// ------------------------------------------------------------------------------------------------ streaming path
// Frame-buffer geometry of the next push of n rows (WarmStartFrameBuffer.insert cases 1-3, hga_optimized.pyx:96-131).
struct PushGeom {
  int y_row0, zero_rows, rows_total;
  int64_t frames;
};
static PushGeom push_geom(const hga_handle* h, int64_t n) {
  const int L = h->cfg.frame_len, hop = h->cfg.hop;
  PushGeom g;
  if (h->first_frame && n >= L) g = {0, 0, (int)n, 0};                               // case 1
  else if (h->first_frame) g = {(int)(L - n), (int)(L - n), L, 0};                   // case 2
  else g = {h->overlap, 0, (int)(h->overlap + n), 0};                                // case 3
  g.frames = g.rows_total >= L ? (g.rows_total - L) / hop + 1 : 0;
  return g;
}
extern "C" int64_t hga_stream_frames(const hga_handle* h, int64_t n) { return h && n > 0 ? push_geom(h, n).frames : 0; }

// Pushes larger than kBulkRows go through the throughput kernels with the handle's state as initial condition:
// thread-per-channel k_stream_push is a latency path.
constexpr int64_t kBulkRows = 1024;
static int bulk_push(hga_handle* h, const double* x, int64_t n, int64_t ldx, const double* mu, const double* sigma, double* out,
                     int64_t ldo, int64_t* n_frames, cudaStream_t st, const PushGeom& g) {
  const int C = h->cfg.n_channels, L = h->cfg.frame_len, hop = h->cfg.hop, ov = h->overlap;
  const bool first = h->first_frame != 0;
  int rc;
  if (h->cfg.precision == HGA_PREC_F64 && h->nb > 0) {
    // [remainder ; y] framing == fused kernel with (nb-1) preloaded hop blocks; frame w of the buffer is emitted when new
    // block w closes, i.e. (nb-1) rows later than for a fresh stream: shift the output origin instead.
    const int64_t s0 = 0, o0 = 0;
    double* out_shift = first ? out : out + (int64_t)(h->nb - 1) * ldo;
    rc = extract_offline_dev(h, x, ldx, 1, &s0, &n, &o0, mu, sigma, out_shift, ldo, st, h->state2.as<double>(), h->framebuf2.as<double>(),
                             first ? nullptr : h->state.as<double>(), first ? nullptr : h->framebuf.as<double>(), /*allow_short=*/true);
    if (rc) return rc;
    swap_bufs(h->state, h->state2);
    swap_bufs(h->framebuf, h->framebuf2);
  } else {
    if ((rc = h->ybuf.reserve((size_t)(ov + n) * C * 8))) return rc;
    double* yb = h->ybuf.as<double>();
    if (!first && ov) CU(cudaMemcpyAsync(yb, h->framebuf.p, (size_t)ov * C * 8, cudaMemcpyDeviceToDevice, st));
    const int cg = (C + 31) / 32;
    k_iir_seq<kExact><<<(cg + 3) / 4, 128, 0, st>>>(h->q, x, ldx, yb + (int64_t)g.y_row0 * C, C, nullptr, Segment{0, n, 0}, 1, C,
                                                     h->state.as<double>(), h->state.as<double>());
    LAUNCHED();
    if (g.frames > 0) {
      if ((rc = hga_log_power(h->cfg.device, yb, g.rows_total, 0, C, C, g.frames, L, hop, nullptr, nullptr, h->cfg.eps, mu, sigma, out, ldo, st)))
        return rc;
    }
    if (ov) CU(cudaMemcpyAsync(h->framebuf.p, yb + (int64_t)(g.rows_total - ov) * C, (size_t)ov * C * 8, cudaMemcpyDeviceToDevice, st));
  }
  h->first_frame = 0;
  if (n_frames) *n_frames = g.frames;
  return HGA_OK;
}

// One packet through the latency kernels.  xf != nullptr: float32 [C][n] payload in place of x.  Small packets take
// k_stream_push_sys (sections across lanes, frame rows and packet in shared memory), the rest k_stream_push.
// Returns true when the launched kernel publishes `done_seq` in done_slots (one slot per pair of channels).
static bool launch_stream(hga_handle* h, const PushGeom& g, const double* x, const float* xf, int64_t ldx, int64_t n, const double* mu,
                          const double* sigma, double* out, int64_t ldo, float* out_f32, cudaStream_t st,
                          unsigned long long* done_slots = nullptr, unsigned long long done_seq = 0) {
  const int C = h->cfg.n_channels;
  if (g.rows_total + n <= kStreamFastRows) {
    const size_t smem = (size_t)kSysWarps * (g.rows_total + n) * 2 * 8;  // <= 12 KB
    const unsigned grid = (unsigned)((C + 2 * kSysWarps - 1) / (2 * kSysWarps));
    if (xf)
      k_stream_push_sys<1><<<grid, 32 * kSysWarps, smem, st>>>(h->q, xf, ldx, (int)n, C, h->state.as<double>(), h->framebuf.as<double>(), g.y_row0,
                                                                g.zero_rows, g.rows_total, h->overlap, (int)g.frames, h->cfg.frame_len, h->cfg.hop,
                                                                h->cfg.eps, mu, sigma, out, ldo, out_f32, done_slots, done_seq);
    else
      k_stream_push_sys<0><<<grid, 32 * kSysWarps, smem, st>>>(h->q, x, ldx, (int)n, C, h->state.as<double>(), h->framebuf.as<double>(), g.y_row0,
                                                                g.zero_rows, g.rows_total, h->overlap, (int)g.frames, h->cfg.frame_len, h->cfg.hop,
                                                                h->cfg.eps, mu, sigma, out, ldo, out_f32, done_slots, done_seq);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    h->first_frame = 0;
    return done_slots != nullptr;
  } else {
    k_stream_push<<<(C + 31) / 32, 32, 0, st>>>(h->q, x, ldx, (int)n, C, h->state.as<double>(), h->framebuf.as<double>(), g.y_row0,
                                                g.zero_rows, g.rows_total, h->overlap, (int)g.frames, h->cfg.frame_len, h->cfg.hop,
                                                h->cfg.eps, mu, sigma, out, ldo, xf, out_f32);
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  h->first_frame = 0;
  return false;
}
// Wait until the frames of the push that published `seq` are in host memory: poll the per-warp slots (mapped pinned
// memory) for a bounded time, then fall back to synchronising the stream (also the path that surfaces a kernel fault).
static int wait_frames(hga_handle* h, bool flagged, unsigned long long seq, int n_slots, cudaStream_t st) {
  if (flagged) {
    const volatile unsigned long long* slots = h->pin_flag.as<const volatile unsigned long long>();
    for (int spin = 0; spin < 200000; ++spin) {
      int k = 0;
      while (k < n_slots && slots[k] == seq) ++k;
      if (k == n_slots) { std::atomic_thread_fence(std::memory_order_acquire); h->push_inflight = true; return HGA_OK; }
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    }
  }
  CU(cudaStreamSynchronize(st));
  h->push_inflight = false;
  return HGA_OK;
}
static int stream_push_dev(hga_handle* h, const double* x, int64_t n, int64_t ldx, const double* mu, const double* sigma,
                           double* out, int64_t ldo, int64_t* n_frames, cudaStream_t st) {
  const int C = h->cfg.n_channels;
  if (n < 1) return fail(HGA_ERR_INVALID, "push of %lld rows", (long long)n);
  if (n <= kBulkRows && n + h->overlap > h->cap_rows) return fail(HGA_ERR_INVALID, "push of %lld rows exceeds max_push_rows", (long long)n);
  if ((mu == nullptr) != (sigma == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  const PushGeom g = push_geom(h, n);
  if (n > kBulkRows && !(h->first_frame && n < h->cfg.frame_len)) return bulk_push(h, x, n, ldx, mu, sigma, out, ldo, n_frames, st, g);
  launch_stream(h, g, x, nullptr, ldx, n, mu, sigma, out, ldo, nullptr, st);
  CU(cudaGetLastError());
  if (n_frames) *n_frames = g.frames;
  return HGA_OK;
}

extern "C" int hga_stream_push(hga_handle* h, const double* x_dev, int64_t n, int64_t ldx, const double* mu_dev,
                               const double* sigma_dev, double* out_dev, int64_t ldo, int64_t* n_frames, void* stream) {
  if (!h || !x_dev || !out_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (ldx < h->cfg.n_channels || ldo < h->cfg.n_channels) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  CU(cudaSetDevice(h->cfg.device));
  drain_push(h);
  const int rc = stream_push_dev(h, x_dev, n, ldx, mu_dev, sigma_dev, out_dev, ldo, n_frames, (cudaStream_t)stream);
  if (rc == HGA_OK && (cudaStream_t)stream != h->s_own) {
    if (!h->dev_push_ev) CU(cudaEventCreateWithFlags(&h->dev_push_ev, cudaEventDisableTiming));
    CU(cudaEventRecord(h->dev_push_ev, (cudaStream_t)stream));
    h->dev_push_pending = true;
  }
  return rc;
}

// mu/sigma host vectors are cached on the device; re-uploaded only when their contents change.
static int sync_musigma(hga_handle* h, const double* mu, const double* sigma, cudaStream_t st, const double** mu_d,
                        const double** sig_d) {
  *mu_d = *sig_d = nullptr;
  if ((mu == nullptr) != (sigma == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  if (!mu) return HGA_OK;
  const int C = h->cfg.n_channels;
  int rc;
  if ((rc = h->mudev.reserve((size_t)C * 8)) || (rc = h->sigdev.reserve((size_t)C * 8))) return rc;
  if (!h->musig_valid || memcmp(h->mu_cache.data(), mu, (size_t)C * 8) || memcmp(h->sig_cache.data(), sigma, (size_t)C * 8)) {
    h->mu_cache.assign(mu, mu + C);
    h->sig_cache.assign(sigma, sigma + C);
    CU(cudaMemcpyAsync(h->mudev.p, h->mu_cache.data(), (size_t)C * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->sigdev.p, h->sig_cache.data(), (size_t)C * 8, cudaMemcpyHostToDevice, st));
    h->musig_valid = true;
  }
  *mu_d = h->mudev.as<double>();
  *sig_d = h->sigdev.as<double>();
  return HGA_OK;
}

extern "C" int hga_stream_push_host(hga_handle* h, const double* x_host, int64_t n, int64_t ldx, const double* mu_host,
                                    const double* sigma_host, double* out_host, int64_t ldo, int64_t* n_frames) {
  if (!h || !x_host || !out_host) return fail(HGA_ERR_INVALID, "null argument");
  const int C = h->cfg.n_channels;
  if (ldx < C || ldo < C) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  if (n < 1 || (n <= kBulkRows && n + h->overlap > h->cap_rows))
    return fail(HGA_ERR_INVALID, "push of %lld rows outside [1, max_push_rows]", (long long)n);
  CU(cudaSetDevice(h->cfg.device));
  cudaStream_t st = h->s_own;
  order_after_dev_push(h);
  int rc;
  const PushGeom g = push_geom(h, n);
  if (n > kBulkRows) {  // bulk: no pinned staging, straight 2-D copies from/to the caller's buffers
    const size_t rb = (size_t)C * 8;
    if ((rc = h->xdev.reserve((size_t)n * rb)) || (rc = h->outdev.reserve((size_t)std::max<int64_t>(1, g.frames) * rb))) return rc;
    const double *mu_b, *sig_b;
    if ((rc = sync_musigma(h, mu_host, sigma_host, st, &mu_b, &sig_b))) return rc;
    CU(cudaMemcpy2DAsync(h->xdev.p, rb, x_host, (size_t)ldx * 8, rb, (size_t)n, cudaMemcpyHostToDevice, st));
    int64_t nfb = 0;
    if ((rc = stream_push_dev(h, h->xdev.as<double>(), n, C, mu_b, sig_b, h->outdev.as<double>(), C, &nfb, st))) return rc;
    if (nfb > 0) CU(cudaMemcpy2DAsync(out_host, (size_t)ldo * 8, h->outdev.p, rb, rb, (size_t)nfb, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (n_frames) *n_frames = nfb;
    return HGA_OK;
  }
  const size_t in_bytes = (size_t)n * C * 8, out_bytes = (size_t)std::max<int64_t>(1, g.frames) * C * 8;
  if ((rc = h->pin_in.reserve(in_bytes)) || (rc = h->pin_out.reserve(out_bytes)) || (rc = h->xdev.reserve(in_bytes)) ||
      (rc = h->outdev.reserve(out_bytes)))
    return rc;
  if (!h->pin_flag.p) {
    if ((rc = h->pin_flag.reserve((size_t)((C + 1) / 2) * 8))) return rc;
    memset(h->pin_flag.p, 0, (size_t)((C + 1) / 2) * 8);
  }
  const double *mu_d, *sig_d;
  if ((rc = sync_musigma(h, mu_host, sigma_host, st, &mu_d, &sig_d))) return rc;
  // gather rows into pinned staging (compacts ldx -> C), one H2D, one kernel, one D2H
  double* pin = h->pin_in.as<double>();
  if (ldx == C) memcpy(pin, x_host, in_bytes);
  else for (int64_t r = 0; r < n; ++r) memcpy(pin + r * C, x_host + r * ldx, (size_t)C * 8);
  int64_t nf = 0;
  bool flagged = false;
  double* in_map = h->pin_in.dev<double>();
  double* out_map = h->pin_out.dev<double>();
  if (n <= kMappedRows && in_map && out_map) {  // zero-copy: see kMappedRows
    unsigned long long* slots = h->pin_flag.dev<unsigned long long>();
    flagged = launch_stream(h, g, in_map, nullptr, C, n, mu_d, sig_d, out_map, C, nullptr, st, slots, ++h->push_seq);
    CU(cudaGetLastError());
    nf = g.frames;
  } else {
    CU(cudaMemcpyAsync(h->xdev.p, pin, in_bytes, cudaMemcpyHostToDevice, st));
    if ((rc = stream_push_dev(h, h->xdev.as<double>(), n, C, mu_d, sig_d, h->outdev.as<double>(), C, &nf, st))) return rc;
    if (nf > 0) CU(cudaMemcpyAsync(h->pin_out.p, h->outdev.p, (size_t)nf * C * 8, cudaMemcpyDeviceToHost, st));
  }
  if ((rc = wait_frames(h, flagged, h->push_seq, (C + 1) / 2, st))) return rc;
  const double* po = h->pin_out.as<double>();
  if (ldo == C) memcpy(out_host, po, (size_t)nf * C * 8);
  else for (int64_t r = 0; r < nf; ++r) memcpy(out_host + r * ldo, po + r * C, (size_t)C * 8);
  if (n_frames) *n_frames = nf;
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ wire packets
// BCI2000 GenericSignal packet as the reference's ZMQ connector receives it (local/units.py:48,78-82; packed by
// development_amplifier.py:14-25): struct '=BBB HH' (descriptor, supplement, dtype, n_channels, n_samples; 7 bytes,
// native byte order, no padding) followed by n_channels * n_samples float32, channel-major.
static int parse_packet(const void* packet, int64_t nbytes, int32_t* n_ch, int32_t* n_samp, const unsigned char** payload) {
  if (!packet) return fail(HGA_ERR_INVALID, "null packet");
  if (nbytes < 7) return fail(HGA_ERR_INVALID, "packet of %lld bytes is shorter than its 7-byte header", (long long)nbytes);
  const unsigned char* b = static_cast<const unsigned char*>(packet);
  uint16_t c, n;
  memcpy(&c, b + 3, 2);
  memcpy(&n, b + 5, 2);
  if ((int64_t)c * n * 4 != nbytes - 7)
    return fail(HGA_ERR_INVALID, "packet payload is %lld bytes, header says %u channels x %u samples float32", (long long)(nbytes - 7), c, n);
  if (n_ch) *n_ch = c;
  if (n_samp) *n_samp = n;
  if (payload) *payload = b + 7;
  return HGA_OK;
}
extern "C" int hga_packet_header(const void* packet_host, int64_t nbytes, int32_t* n_channels, int32_t* n_samples) {
  return parse_packet(packet_host, nbytes, n_channels, n_samples, nullptr);
}
extern "C" int hga_packet_decode(int32_t device, const void* packet_host, int64_t nbytes, double* x_dev, int64_t ldx, void* stream) {
  int32_t C = 0, n = 0;
  const unsigned char* payload = nullptr;
  int rc = parse_packet(packet_host, nbytes, &C, &n, &payload);
  if (rc) return rc;
  if (!x_dev || ldx < C) return fail(HGA_ERR_INVALID, "bad output: ldx=%lld for %d channels", (long long)ldx, C);
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  if (C == 0 || n == 0) return HGA_OK;
  CU(cudaSetDevice(device));
  cudaStream_t st = (cudaStream_t)stream;
  float* tmp = nullptr;
  const size_t bytes = (size_t)C * n * 4;
  CU(cudaMallocAsync((void**)&tmp, bytes, st));
  CU(cudaMemcpyAsync(tmp, payload, bytes, cudaMemcpyHostToDevice, st));  // pageable, unaligned source: staged by the driver
  k_packet_rows<<<dim3((C + 31) / 32, (n + 31) / 32), 256, 0, st>>>(tmp, C, n, x_dev, ldx);
  LAUNCHED();
  CU(cudaFreeAsync(tmp, st));
  return HGA_OK;
}
// One packet through a long-lived extractor: payload -> pinned staging -> H2D -> ONE kernel that reads the float32
// channel-major payload directly (no transposed float64 copy is ever made) -> D2H of the frames.
extern "C" int hga_stream_push_packet(hga_handle* h, const void* packet_host, int64_t nbytes, const double* mu_host, const double* sigma_host,
                                      double* out_host, int64_t ldo, float* out_f32_dev, int64_t* n_frames) {
  if (!h || !out_host) return fail(HGA_ERR_INVALID, "null argument");
  int32_t Cp = 0, np_ = 0;
  const unsigned char* payload = nullptr;
  int rc = parse_packet(packet_host, nbytes, &Cp, &np_, &payload);
  if (rc) return rc;
  const int C = h->cfg.n_channels;
  const int64_t n = np_;
  if (Cp != C) return fail(HGA_ERR_INVALID, "packet has %d channels, extractor has %d", Cp, C);
  if (ldo < C) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  if (n < 1 || n > kBulkRows || n + h->overlap > h->cap_rows) return fail(HGA_ERR_INVALID, "packet of %lld samples outside [1, min(max_push_rows, %lld)]", (long long)n, (long long)kBulkRows);
  CU(cudaSetDevice(h->cfg.device));
  cudaStream_t st = h->s_own;
  order_after_dev_push(h);
  const PushGeom g = push_geom(h, n);
  const size_t in_bytes = (size_t)n * C * 4, out_bytes = (size_t)std::max<int64_t>(1, g.frames) * C * 8;
  if ((rc = h->pin_in.reserve(in_bytes)) || (rc = h->pin_out.reserve(out_bytes)) || (rc = h->xdev.reserve(in_bytes)) ||
      (rc = h->outdev.reserve(out_bytes)))
    return rc;
  if (!h->pin_flag.p) {
    if ((rc = h->pin_flag.reserve((size_t)((C + 1) / 2) * 8))) return rc;
    memset(h->pin_flag.p, 0, (size_t)((C + 1) / 2) * 8);
  }
  const double *mu_d, *sig_d;
  if ((rc = sync_musigma(h, mu_host, sigma_host, st, &mu_d, &sig_d))) return rc;
  memcpy(h->pin_in.p, payload, in_bytes);
  const int64_t nf = g.frames;
  bool flagged = false;
  float* in_map = h->pin_in.dev<float>();
  double* out_map = h->pin_out.dev<double>();
  if (n <= kMappedRows && in_map && out_map) {
    unsigned long long* slots = h->pin_flag.dev<unsigned long long>();
    flagged = launch_stream(h, g, nullptr, in_map, C, n, mu_d, sig_d, out_map, C, out_f32_dev, st, slots, ++h->push_seq);
    CU(cudaGetLastError());
  } else {
    CU(cudaMemcpyAsync(h->xdev.p, h->pin_in.p, in_bytes, cudaMemcpyHostToDevice, st));
    launch_stream(h, g, nullptr, h->xdev.as<float>(), C, n, mu_d, sig_d, h->outdev.as<double>(), C, out_f32_dev, st);
    CU(cudaGetLastError());
    if (nf > 0) CU(cudaMemcpyAsync(h->pin_out.p, h->outdev.p, (size_t)nf * C * 8, cudaMemcpyDeviceToHost, st));
  }
  h->first_frame = 0;
  if ((rc = wait_frames(h, flagged, h->push_seq, (C + 1) / 2, st))) return rc;
  const double* po = h->pin_out.as<double>();
  if (ldo == C) memcpy(out_host, po, (size_t)nf * C * 8);
  else for (int64_t r = 0; r < nf; ++r) memcpy(out_host + r * ldo, po + r * C, (size_t)C * 8);
  if (n_frames) *n_frames = nf;
  return HGA_OK;
}

extern "C" int hga_get_state(hga_handle* h, double* hg_state, double* fh_state, double* remainder, int32_t* first_frame) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  CU(cudaSetDevice(h->cfg.device));
  const size_t C = h->cfg.n_channels, na = h->q.ns_a, nb = h->q.ns - h->q.ns_a;
  CU(cudaDeviceSynchronize());
  if (hg_state) CU(cudaMemcpy(hg_state, h->state.p, na * 2 * C * 8, cudaMemcpyDeviceToHost));
  if (fh_state && nb) CU(cudaMemcpy(fh_state, h->state.as<double>() + na * 2 * C, nb * 2 * C * 8, cudaMemcpyDeviceToHost));
  if (remainder && h->overlap) CU(cudaMemcpy(remainder, h->framebuf.p, (size_t)h->overlap * C * 8, cudaMemcpyDeviceToHost));
  if (first_frame) *first_frame = h->first_frame;
  return HGA_OK;
}
extern "C" int hga_set_state(hga_handle* h, const double* hg_state, const double* fh_state, const double* remainder,
                             const int32_t* first_frame) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  CU(cudaSetDevice(h->cfg.device));
  const size_t C = h->cfg.n_channels, na = h->q.ns_a, nb = h->q.ns - h->q.ns_a;
  CU(cudaDeviceSynchronize());
  if (hg_state) CU(cudaMemcpy(h->state.p, hg_state, na * 2 * C * 8, cudaMemcpyHostToDevice));
  if (fh_state && nb) CU(cudaMemcpy(h->state.as<double>() + na * 2 * C, fh_state, nb * 2 * C * 8, cudaMemcpyHostToDevice));
  if (remainder && h->overlap) CU(cudaMemcpy(h->framebuf.p, remainder, (size_t)h->overlap * C * 8, cudaMemcpyHostToDevice));
  if (first_frame) h->first_frame = *first_frame ? 1 : 0;
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ building blocks
extern "C" int hga_sosfilt(int32_t device, const double* sos, int32_t n_sections, const double* x_dev, int64_t T, int32_t C,
                           int64_t ldx, double* state_dev, double* y_dev, int64_t ldy, void* stream) {
  if (!sos || !x_dev || !state_dev || !y_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (n_sections < 1 || n_sections > kMaxSec) return fail(HGA_ERR_INVALID, "n_sections %d outside [1, %d]", n_sections, kMaxSec);
  if (T < 0 || C < 1 || ldx < C || ldy < C) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  Cascade q;
  int rc = build_cascade(q, sos, n_sections, nullptr, nullptr, 0, nullptr);
  if (rc) return rc;
  if (T == 0) return HGA_OK;
  const int cg = (C + 31) / 32;
  k_iir_seq<kExact><<<(cg + 3) / 4, 128, 0, (cudaStream_t)stream>>>(q, x_dev, ldx, y_dev, ldy, nullptr, Segment{0, T, 0}, 1, C, state_dev,
                                                                    state_dev);
  LAUNCHED();
  return HGA_OK;
}

extern "C" int hga_log_power(int32_t device, const double* data_dev, int64_t rows, int64_t zero_rows, int32_t C, int64_t ld,
                             int64_t n_frames, int32_t frame_len, int32_t hop, const int64_t* start_dev, const int64_t* stop_dev,
                             double eps, const double* mu_dev, const double* sigma_dev, double* out_dev, int64_t ldo, void* stream) {
  if (!out_dev || (!data_dev && rows > 0)) return fail(HGA_ERR_INVALID, "null argument");
  if ((start_dev == nullptr) != (stop_dev == nullptr)) return fail(HGA_ERR_INVALID, "start and stop must both be given or both be NULL");
  if ((mu_dev == nullptr) != (sigma_dev == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  if (C < 1 || ld < C || ldo < C || rows < 0 || zero_rows < 0) return fail(HGA_ERR_INVALID, "bad shape");
  if (!start_dev) {
    if (frame_len < 1 || hop < 1) return fail(HGA_ERR_INVALID, "bad window geometry");
    const int64_t tot = rows + zero_rows;
    const int64_t F = tot >= frame_len ? (tot - frame_len) / hop + 1 : 0;
    if (n_frames > F) return fail(HGA_ERR_INVALID, "n_frames %lld > %lld available", (long long)n_frames, (long long)F);
  }
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  if (n_frames <= 0) return HGA_OK;
  if (!start_dev && zero_rows == 0 && frame_len == 5 * hop) {
    // the path's geometry: rolling sums, one load per sample (bit-identical to the per-window loop below)
    int n_sm = 148;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    const int cg = (C + 31) / 32, R = roll_run_length(n_frames, cg, n_sm);
    const int64_t rows_used = (n_frames - 1) * hop + frame_len;   // exactly n_frames windows
    dim3 gr((unsigned)cg, (unsigned)((n_frames + (int64_t)R * kRollWarps - 1) / ((int64_t)R * kRollWarps)), 1);
    if (gr.y > 65535u) return fail(HGA_ERR_UNSUPPORTED, "too many frames in one call");
    k_frames_roll<5><<<gr, 32 * kRollWarps, 0, (cudaStream_t)stream>>>(data_dev, ld, nullptr, Segment{0, rows_used, 0}, 1, C, hop, R, eps, mu_dev,
                                                                       sigma_dev, out_dev, ldo);
    LAUNCHED();
    return HGA_OK;
  }
  dim3 g((unsigned)((C + 31) / 32), (unsigned)((n_frames + 7) / 8), 1);
  if (g.y > 65535u * 64u) return fail(HGA_ERR_UNSUPPORTED, "too many frames in one call");
  // grid.y is limited to 65535: fold the excess into several launches
  const int64_t per = 65535LL * 8;
  for (int64_t w0 = 0; w0 < n_frames; w0 += per) {
    const int64_t nf = std::min(per, n_frames - w0);
    dim3 gg(g.x, (unsigned)((nf + kFramesPerBlock - 1) / kFramesPerBlock), 1);
    // window w0+w: shift the data origin for regular windows, the index arrays for irregular ones
    if (!start_dev) {
      const int64_t shift = w0 * hop;  // logical rows to skip
      const int64_t zr = std::max<int64_t>(0, zero_rows - shift), skip = std::max<int64_t>(0, shift - zero_rows);
      k_frames_seq<<<gg, 256, 0, (cudaStream_t)stream>>>(data_dev + skip * ld, ld, zr, nullptr, Segment{0, rows - skip, w0}, 1, C, nf,
                                                         frame_len, hop, nullptr, nullptr, eps, mu_dev, sigma_dev, out_dev, ldo);
    } else {
      k_frames_seq<<<gg, 256, 0, (cudaStream_t)stream>>>(data_dev, ld, zero_rows, nullptr, Segment{0, rows, w0}, 1, C, nf, frame_len, hop,
                                                         start_dev + w0, stop_dev + w0, eps, mu_dev, sigma_dev, out_dev, ldo);
    }
    LAUNCHED();
  }
  return HGA_OK;
}

extern "C" int hga_zscore(int32_t device, const double* f_dev, int64_t F, int32_t C, int64_t ld, const double* mu_dev,
                          const double* sigma_dev, double* out_dev, int64_t ldo, void* stream) {
  if (!f_dev || !mu_dev || !sigma_dev || !out_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (C < 1 || ld < C || ldo < C || F < 0) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  if (F == 0) return HGA_OK;
  int n_sm = 148;
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
  if (f_dev != out_dev && (C & 3) == 0 && (ld & 3) == 0 && (ldo & 3) == 0 &&
      (((uintptr_t)f_dev | (uintptr_t)out_dev | (uintptr_t)mu_dev | (uintptr_t)sigma_dev) & 31) == 0) {
    const unsigned cg4 = (unsigned)((C / 4 + 31) / 32);
    const unsigned gy4 = (unsigned)std::max<int64_t>(1, std::min<int64_t>((F + 31) / 32, (int64_t)n_sm * 8 / cg4 + 1));
    k_zscore4<<<dim3(cg4, gy4), 256, 0, (cudaStream_t)stream>>>(f_dev, F, C, ld, mu_dev, sigma_dev, out_dev, ldo);
    LAUNCHED();
    return HGA_OK;
  }
  const unsigned cgz = (unsigned)((C + 31) / 32);
  const unsigned gy = (unsigned)std::max<int64_t>(1, std::min<int64_t>((F + 31) / 32, (int64_t)n_sm * 8 / cgz + 1));
  k_zscore<<<dim3(cgz, gy), 256, 0, (cudaStream_t)stream>>>(f_dev, F, C, ld, mu_dev, sigma_dev, out_dev, ldo);
  LAUNCHED();
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ statistics
extern "C" int hga_stats_accumulate(int32_t device, const double* feats_dev, int64_t F, int32_t C, int64_t ld, double* acc_dev,
                                    void* stream) {
  if (!feats_dev || !acc_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (C < 1 || ld < C || F < 0) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  if (F == 0) return HGA_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int n_sm = 148;
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
  const int cg = (C + 31) / 32;
  int64_t nby = std::max<int64_t>(1, std::min<int64_t>((int64_t)n_sm * 8 / cg, (F + 63) / 64));
  nby = std::min<int64_t>(nby, 65535);
  const int64_t rpb = (F + nby - 1) / nby;
  nby = (F + rpb - 1) / rpb;
  // block partials: one grow-only buffer per device and host thread (calls on one stream are ordered; a per-call
  // cudaMallocAsync made the duration vary from run to run)
  static thread_local DevBuf part_buf[16];
  DevBuf& pb = part_buf[device & 15];
  int rc;
  if ((rc = pb.reserve((size_t)nby * 3 * C * 8))) return rc;
  double* part = pb.as<double>();
  k_stats_partial<<<dim3((unsigned)cg, (unsigned)nby), 256, 0, st>>>(feats_dev, F, C, ld, rpb, part);
  LAUNCHED();
  k_stats_merge<<<(C + 127) / 128, 128, 0, st>>>(acc_dev, part, (int)nby, C);
  LAUNCHED();
  return HGA_OK;
}
// ---- hga_stats_accumulate without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The block partials of k_stats_partial live in a file-scope device array instead of the grow-only DevBuf above.  Every GPU
// that loads the library holds its 1 MiB until exit.  In host code the array's name is only a stand-in: its device address
// is looked up once per device, with that device current (cudaGetSymbolAddress), and that pointer goes to the kernels.
// All calls on a device share the array: calls that overlap on the GPU (different streams, no ordering) race on it.
// Nothing in it has to start at zero: k_stats_partial writes every partial k_stats_merge reads (nby x 3 x C doubles), the
// rest keeps whatever the last call left.
constexpr size_t kStatsStaticDoubles = (1u << 20) / 8;   // 131 072 doubles = 1 MiB
static __device__ __align__(256) double g_stats_static[kStatsStaticDoubles];
static std::mutex g_stats_static_mu;
static double* g_stats_static_ptr[64] = {};
extern "C" int hga_stats_accumulate_static(int32_t device, const double* feats_dev, int64_t F, int32_t C, int64_t ld, double* acc_dev,
                                           void* stream) {
  if (!feats_dev || !acc_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (C < 1 || ld < C || F < 0) return fail(HGA_ERR_INVALID, "bad shape");
  if (device < 0 || device >= 64) return fail(HGA_ERR_INVALID, "device ordinal %d outside [0, 64)", device);
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  int n_sm = 148;
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
  // the same tiling as hga_stats_accumulate (same partials, same merge order: the two entry points return the same bits)
  const int cg = (C + 31) / 32;
  int64_t nby = std::max<int64_t>(1, std::min<int64_t>((int64_t)n_sm * 8 / cg, (std::max<int64_t>(F, 1) + 63) / 64));
  nby = std::min<int64_t>(nby, 65535);
  const int64_t rpb = (std::max<int64_t>(F, 1) + nby - 1) / nby;
  nby = (std::max<int64_t>(F, 1) + rpb - 1) / rpb;
  const size_t need = (size_t)nby * 3 * (size_t)C;
  if (need > kStatsStaticDoubles)
    return fail(HGA_ERR_INVALID, "hga_stats_accumulate_static: %lld row tiles x 3 x %d channels = %zu doubles of block partials exceed the static scratch of %zu "
                "doubles (1 MiB; at most 43690 channels) - use hga_stats_accumulate", (long long)nby, C, need, kStatsStaticDoubles);
  if (F == 0) return HGA_OK;
  double* part = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_stats_static_mu);
    if (!g_stats_static_ptr[device]) {   // set-up, once per device, with that device current
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_stats_static);
      if (e != cudaSuccess || !p) {
        cudaGetLastError();
        return fail(HGA_ERR_CUDA, "cudaGetSymbolAddress(g_stats_static) failed on device %d: %s", device, cudaGetErrorString(e));
      }
      g_stats_static_ptr[device] = static_cast<double*>(p);
    }
    part = g_stats_static_ptr[device];
  }
  cudaStream_t st = (cudaStream_t)stream;
  k_stats_partial<<<dim3((unsigned)cg, (unsigned)nby), 256, 0, st>>>(feats_dev, F, C, ld, rpb, part);
  LAUNCHED();
  k_stats_merge<<<(C + 127) / 128, 128, 0, st>>>(acc_dev, part, (int)nby, C);
  LAUNCHED();
  return HGA_OK;
}
extern "C" int hga_stats_merge_days(int32_t device, const double* gathered_dev, int32_t n_src, int32_t n_days, int32_t C, double* mu_dev,
                                    double* sigma_dev, void* stream) {
  if (!gathered_dev || !mu_dev || !sigma_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (n_src < 1 || n_days < 1 || n_days > 65535 || C < 1) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  k_stats_merge_days<<<dim3((unsigned)((C + 127) / 128), (unsigned)n_days), 128, 0, (cudaStream_t)stream>>>(gathered_dev, n_src, n_days, C, mu_dev, sigma_dev);
  LAUNCHED();
  return HGA_OK;
}
extern "C" int hga_stats_merge(int32_t device, double* acc_a_dev, const double* acc_b_dev, int32_t count, int32_t C, void* stream) {
  if (!acc_a_dev || !acc_b_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (C < 1 || count < 0) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  if (count == 0) return HGA_OK;
  k_stats_merge<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(acc_a_dev, acc_b_dev, count, C);
  LAUNCHED();
  return HGA_OK;
}
extern "C" int hga_stats_finalize(int32_t device, const double* acc_dev, int32_t C, double* mu_dev, double* sigma_dev, void* stream) {
  if (!acc_dev || !mu_dev || !sigma_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (C < 1) return fail(HGA_ERR_INVALID, "bad shape");
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  CU(cudaSetDevice(device));
  k_stats_finalize<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(acc_dev, C, mu_dev, sigma_dev);
  LAUNCHED();
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ channel maps
// Gather / common-average-reference / bad-channel patch along the channel axis (hga_kernels_pre.cuh).
struct hga_chanmap {
  int device = 0, c_in = 0, c_out = 0, n_groups = 0, n_idx = 0, n_sm = 0;
  DevBuf tab;  // sel | grp | gptr | gidx (int32) followed by alpha | beta (double)
  const int32_t *sel = nullptr, *grp = nullptr, *gptr = nullptr, *gidx = nullptr;
  const double *alpha = nullptr, *beta = nullptr;
  DevBuf xdev, odev;  // staging of the host-buffer variant
  bool identity = false;  // out[t, j] = x[t, j]: raw ingest takes the flat k_rawscale stream
};

extern "C" int hga_chanmap_create(int32_t device, int32_t c_in, int32_t c_out, const int32_t* sel, const int32_t* grp, const double* alpha,
                                  const double* beta, int32_t n_groups, const int32_t* gptr, const int32_t* gidx, hga_chanmap** out) {
  if (!out) return fail(HGA_ERR_INVALID, "null argument");
  *out = nullptr;
  if (!sel || !grp || !alpha || !beta || (n_groups > 0 && (!gptr || !gidx))) return fail(HGA_ERR_INVALID, "null argument");
  if (c_in < 1 || c_out < 1 || n_groups < 0 || n_groups > kChanMapMaxGroups) return fail(HGA_ERR_INVALID, "bad shape: c_in=%d c_out=%d groups=%d (max %d)", c_in, c_out, n_groups, kChanMapMaxGroups);
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  if (device < 0 || device >= hga_device_count()) return fail(HGA_ERR_INVALID, "bad device %d", device);
  const int n_idx = n_groups > 0 ? gptr[n_groups] : 0;
  for (int j = 0; j < c_out; ++j) {
    if (sel[j] < -1 || sel[j] >= c_in) return fail(HGA_ERR_INVALID, "sel[%d] = %d outside [0, %d)", j, sel[j], c_in);
    if (grp[j] < -1 || grp[j] >= n_groups) return fail(HGA_ERR_INVALID, "grp[%d] = %d outside [0, %d)", j, grp[j], n_groups);
  }
  for (int g = 0; g < n_groups; ++g)
    if (gptr[g + 1] <= gptr[g]) return fail(HGA_ERR_INVALID, "group %d is empty", g);
  for (int k = 0; k < n_idx; ++k)
    if (gidx[k] < 0 || gidx[k] >= c_in) return fail(HGA_ERR_INVALID, "group column %d outside [0, %d)", gidx[k], c_in);
  CU(cudaSetDevice(device));
  hga_chanmap* m = new hga_chanmap();
  m->device = device; m->c_in = c_in; m->c_out = c_out; m->n_groups = n_groups; m->n_idx = n_idx;
  m->identity = c_in == c_out;
  for (int j = 0; j < c_out && m->identity; ++j) m->identity = sel[j] == j && grp[j] < 0 && alpha[j] == 1.0;
  cudaDeviceGetAttribute(&m->n_sm, cudaDevAttrMultiProcessorCount, device);
  const size_t n_i32 = (size_t)2 * c_out + (n_groups + 1) + n_idx, off_d = (n_i32 * 4 + 15) / 16 * 16;
  std::vector<unsigned char> hbuf(off_d + (size_t)2 * c_out * 8, 0);
  int32_t* hi = reinterpret_cast<int32_t*>(hbuf.data());
  memcpy(hi, sel, (size_t)c_out * 4);
  memcpy(hi + c_out, grp, (size_t)c_out * 4);
  if (n_groups > 0) {
    memcpy(hi + 2 * c_out, gptr, (size_t)(n_groups + 1) * 4);
    memcpy(hi + 2 * c_out + n_groups + 1, gidx, (size_t)n_idx * 4);
  }
  double* hd = reinterpret_cast<double*>(hbuf.data() + off_d);
  memcpy(hd, alpha, (size_t)c_out * 8);
  memcpy(hd + c_out, beta, (size_t)c_out * 8);
  int rc = m->tab.reserve(hbuf.size());
  if (rc) { delete m; return rc; }
  if (cudaMemcpy(m->tab.p, hbuf.data(), hbuf.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
    m->tab.release();
    delete m;
    return fail(HGA_ERR_CUDA, "upload of the channel-map tables failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  const int32_t* di = m->tab.as<int32_t>();
  m->sel = di; m->grp = di + c_out; m->gptr = di + 2 * c_out; m->gidx = di + 2 * c_out + n_groups + 1;
  m->alpha = reinterpret_cast<const double*>(static_cast<unsigned char*>(m->tab.p) + off_d);
  m->beta = m->alpha + c_out;
  *out = m;
  return HGA_OK;
}
extern "C" int hga_chanmap_destroy(hga_chanmap* m) {
  if (!m) return HGA_OK;
  cudaSetDevice(m->device);
  m->tab.release(); m->xdev.release(); m->odev.release();
  delete m;
  return HGA_OK;
}
template <typename IN, bool GAIN>
static int chanmap_launch(hga_chanmap* m, const ChanMapArgs& a, size_t smem, cudaStream_t stream) {
  static size_t smem_set[64] = {};  // per instantiation and device
  size_t& set = smem_set[m->device & 63];
  if (smem > 48 * 1024 && smem > set) {
    CU(cudaFuncSetAttribute(k_chanmap<IN, GAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = smem;
  }
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / std::max<size_t>(smem, 1)));
  const int64_t need = (a.T + kChanMapWarps * kChanMapRows - 1) / (kChanMapWarps * kChanMapRows);
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)m->n_sm * per_sm));
  k_chanmap<IN, GAIN><<<grid, 32 * kChanMapWarps, smem, stream>>>(a);
  LAUNCHED();
  return HGA_OK;
}
template <typename IN>
static int chanmap_launch_g(hga_chanmap* m, const ChanMapArgs& a, size_t smem, cudaStream_t stream) {
  return a.gain ? chanmap_launch<IN, true>(m, a, smem, stream) : chanmap_launch<IN, false>(m, a, smem, stream);
}

// Identity map over compact rows: the flat k_rawscale stream.
static bool rawscale_applies(const hga_chanmap* m, const void* x, int32_t dtype, const double* gain, int64_t ldx, const double* out, int64_t ldo) {
  if (!m->identity || ldx != m->c_in || ldo != m->c_out || (m->c_in & 3)) return false;
  const size_t quad = 4 * (dtype == HGA_RAW_F64 ? 8 : dtype == HGA_RAW_I16 ? 2 : 4);
  return ((uintptr_t)x % (quad > 16 ? 16 : quad)) == 0 && ((uintptr_t)out & 31) == 0 && ((uintptr_t)gain & 15) == 0;
}
template <typename IN>
static int rawscale_launch(hga_chanmap* m, const void* x, const double* gain, int64_t T, double* out, cudaStream_t stream) {
  const int qpr = m->c_in / 4;
  const int64_t n_quads = T * qpr;
  // total threads a multiple of quads_per_row: each thread keeps its four gain factors in registers
  int64_t q = qpr, t = kRawScaleThreads;
  while (t) { const int64_t r = q % t; q = t; t = r; }   // q = gcd(qpr, threads)
  const int64_t mult = qpr / q;
  const int64_t need = (n_quads + (int64_t)kRawScaleThreads * kRawScaleUnroll - 1) / ((int64_t)kRawScaleThreads * kRawScaleUnroll);
  int64_t grid = std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)m->n_sm * 4));   // 56 registers: 4 CTAs per SM
  grid = (grid + mult - 1) / mult * mult;
  if (gain)
    k_rawscale<IN, true><<<(unsigned)grid, kRawScaleThreads, 0, stream>>>(static_cast<const IN*>(x), gain, out, n_quads, qpr);
  else
    k_rawscale<IN, false><<<(unsigned)grid, kRawScaleThreads, 0, stream>>>(static_cast<const IN*>(x), nullptr, out, n_quads, qpr);
  LAUNCHED();
  return HGA_OK;
}

extern "C" int hga_chanmap_apply_raw(hga_chanmap* m, const void* x_dev, int32_t dtype, const double* gain_dev, int64_t T, int64_t ldx,
                                     double* out_dev, int64_t ldo, void* stream) {
  if (!m || !x_dev || !out_dev) return fail(HGA_ERR_INVALID, "null argument");
  if (T < 0 || ldx < m->c_in || ldo < m->c_out) return fail(HGA_ERR_INVALID, "bad shape: T=%lld ldx=%lld ldo=%lld", (long long)T, (long long)ldx, (long long)ldo);
  if (dtype < HGA_RAW_F64 || dtype > HGA_RAW_I16) return fail(HGA_ERR_INVALID, "unknown raw dtype %d", dtype);
  if (T == 0) return HGA_OK;
  CU(cudaSetDevice(m->device));
  ChanMapArgs a{x_dev, gain_dev, ldx, T, m->c_in, m->c_out, m->n_groups, m->n_idx, m->sel, m->grp, m->alpha, m->beta, m->gptr, m->gidx, out_dev, ldo};
  const size_t smem = chanmap_smem_bytes(m->c_in, m->c_out, m->n_groups, m->n_idx);
  if (smem > 227 * 1024) return fail(HGA_ERR_UNSUPPORTED, "channel map needs %zu bytes of shared memory (c_in=%d c_out=%d)", smem, m->c_in, m->c_out);
  cudaStream_t st = (cudaStream_t)stream;
  if (rawscale_applies(m, x_dev, dtype, gain_dev, ldx, out_dev, ldo)) {
    switch (dtype) {
      case HGA_RAW_F64: return rawscale_launch<double>(m, x_dev, gain_dev, T, out_dev, st);
      case HGA_RAW_F32: return rawscale_launch<float>(m, x_dev, gain_dev, T, out_dev, st);
      case HGA_RAW_I32: return rawscale_launch<int32_t>(m, x_dev, gain_dev, T, out_dev, st);
      default: return rawscale_launch<int16_t>(m, x_dev, gain_dev, T, out_dev, st);
    }
  }
  switch (dtype) {
    case HGA_RAW_F64: return chanmap_launch_g<double>(m, a, smem, st);
    case HGA_RAW_F32: return chanmap_launch_g<float>(m, a, smem, st);
    case HGA_RAW_I32: return chanmap_launch_g<int32_t>(m, a, smem, st);
    default: return chanmap_launch_g<int16_t>(m, a, smem, st);
  }
}
extern "C" int hga_chanmap_apply(hga_chanmap* m, const double* x_dev, int64_t T, int64_t ldx, double* out_dev, int64_t ldo, void* stream) {
  return hga_chanmap_apply_raw(m, x_dev, HGA_RAW_F64, nullptr, T, ldx, out_dev, ldo, stream);
}
extern "C" int hga_chanmap_apply_host(hga_chanmap* m, const double* x_host, int64_t T, int64_t ldx, double* out_host, int64_t ldo) {
  if (!m || !x_host || !out_host) return fail(HGA_ERR_INVALID, "null argument");
  if (T < 0 || ldx < m->c_in || ldo < m->c_out) return fail(HGA_ERR_INVALID, "bad shape");
  if (T == 0) return HGA_OK;
  CU(cudaSetDevice(m->device));
  int rc;
  if ((rc = m->xdev.reserve((size_t)T * m->c_in * 8)) || (rc = m->odev.reserve((size_t)T * m->c_out * 8))) return rc;
  CU(cudaMemcpy2D(m->xdev.p, (size_t)m->c_in * 8, x_host, (size_t)ldx * 8, (size_t)m->c_in * 8, (size_t)T, cudaMemcpyHostToDevice));
  if ((rc = hga_chanmap_apply(m, m->xdev.as<double>(), T, m->c_in, m->odev.as<double>(), m->c_out, nullptr))) return rc;
  CU(cudaMemcpy2D(out_host, (size_t)ldo * 8, m->odev.p, (size_t)m->c_out * 8, (size_t)m->c_out * 8, (size_t)T, cudaMemcpyDeviceToHost));
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ offline + baseline statistics
// The day-baseline flow of prepare_corpus.py:190-199 / baseline_offline.py:45-60 - extract the baseline recording's trials,
// then mean / std over all their frames - in one pass: the kernels that emit frames (k_hga_lanes, k_lane_fix,
// k_lane_fix_frames) also leave (n, mean, M2) of what they emitted, and one small kernel Chan-merges those partials into
// acc in a fixed order.  The features are never read back.  Packed-float32 handles only (see lane_launch); every other handle
// (EXACT, F64, MIXED, non-standard cascades, layouts the TMA cannot describe) extracts first and accumulates in a second pass.
extern "C" int hga_extract_offline_stats(hga_handle* h, const double* x_dev, int64_t ldx, int32_t n_segments, const int64_t* seg_start,
                                         const int64_t* seg_len, const int64_t* out_row, const double* mu_dev, const double* sigma_dev,
                                         double* out_dev, int64_t ldo, double* acc_dev, void* stream, int32_t* fused) {
  if (fused) *fused = 0;
  if (!h || !x_dev || !out_dev || !acc_dev || !seg_start || !seg_len || !out_row) return fail(HGA_ERR_INVALID, "null argument");
  const int C = h->cfg.n_channels, L = h->cfg.frame_len, hop = h->cfg.hop;
  if (ldx < C || ldo < C) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  if ((mu_dev == nullptr) != (sigma_dev == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  if (n_segments < 1) return HGA_OK;
  for (int i = 0; i < n_segments; ++i)
    if (seg_len[i] < L) return fail(HGA_ERR_INVALID, "segment %d has %lld rows < frame_len %d", i, (long long)seg_len[i], L);
  CU(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  const bool lanes = h->cfg.precision == HGA_PREC_F32 && h->nb > 0 && choose_variant(h, x_dev, ldx, true) == 2;
  if (lanes) {
    int64_t xlo = INT64_MAX, xhi = 0;
    for (int i = 0; i < n_segments; ++i) { xlo = std::min(xlo, seg_start[i]); xhi = std::max(xhi, seg_start[i] + seg_len[i]); }
    if ((rc = lane_prepare(h, n_segments, seg_start, seg_len, out_row, xlo, st))) return rc;
    const LanePlan& p = h->lane_plan_host;
    const size_t CP = (size_t)p.groups * p.cw, n_items = p.work.size() * p.groups, nf = p.fix.size();
    const size_t n_main = n_items * 3 * p.cw, n_fix = nf * h->lane_nsub * 3 * CP, n_seam = nf * 3 * CP;
    if ((rc = h->lane_st.reserve(std::max<size_t>(16, (n_main + n_fix + n_seam) * 8)))) return rc;
    CUtensorMap tm;
    if ((rc = lane_tensor(h, x_dev, ldx, xlo, xhi, &tm))) return rc;
    LaneDev ld = lane_dev(h);
    ld.x0 = x_dev + xlo * ldx;
    ld.ldx = ldx;
    ld.st_main = h->lane_st.as<double>();
    ld.st_fix = ld.st_main + n_main;
    ld.st_seam = ld.st_fix + n_fix;
    if ((rc = lane_launch(h->cfg.precision, hop, h->lk, p, ld, tm, C, mu_dev, sigma_dev, out_dev, ldo, h->n_sm, st, true))) return rc;
    k_lane_stats_merge<<<C, 128, 0, st>>>(acc_dev, ld.st_main, (int)p.work.size(), p.groups, p.cw, ld.st_fix, (int)(nf * h->lane_nsub), ld.st_seam,
                                                   (int)nf, C, (int)CP);
    LAUNCHED();
    if (fused) *fused = 1;
    return HGA_OK;
  }
  int64_t F = 0;
  for (int i = 0; i < n_segments; ++i) F = std::max<int64_t>(F, out_row[i] + (seg_len[i] - L) / hop + 1);
  if ((rc = extract_offline_dev(h, x_dev, ldx, n_segments, seg_start, seg_len, out_row, mu_dev, sigma_dev, out_dev, ldo, st, nullptr, nullptr))) return rc;
  // second pass over every segment's frames (the out_row ranges need not tile one block)
  for (int i = 0; i < n_segments; ++i) {
    const int64_t nf = (seg_len[i] - L) / hop + 1;
    if ((rc = hga_stats_accumulate(h->cfg.device, out_dev + out_row[i] * ldo, nf, C, ldo, acc_dev, stream))) return rc;
  }
  (void)F;
  return HGA_OK;
}

// ------------------------------------------------------------------------------------------------ offline, channel map fused
// extract_features with its pre-transform chain (local/units.py:147-152): the folded channel map runs inside the extraction
// kernel's load stage (mapper warps, hga_kernels_lane.cuh) when the handle computes in packed float32 at hop 10 and the map
// fits the kernel (raw rows of up to ~180 columns, at most 2 averaging groups); otherwise the map runs as a pass of its own
// into scratch and the plain extraction follows.  Same results either way to the precision's tolerance.
extern "C" int hga_extract_offline_mapped(hga_handle* h, hga_chanmap* map, const double* x_raw_dev, int64_t ldx, int32_t n_segments,
                                          const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row, const double* mu_dev,
                                          const double* sigma_dev, double* out_dev, int64_t ldo, void* stream, int32_t* fused) {
  if (fused) *fused = 0;
  if (!h || !map || !x_raw_dev || !out_dev || !seg_start || !seg_len || !out_row) return fail(HGA_ERR_INVALID, "null argument");
  const int C = h->cfg.n_channels, L = h->cfg.frame_len, hop = h->cfg.hop;
  if (map->c_out != C || map->device != h->cfg.device)
    return fail(HGA_ERR_INVALID, "channel map yields %d columns on device %d, the extractor wants %d on device %d", map->c_out, map->device, C, h->cfg.device);
  if (ldx < map->c_in || ldo < C) return fail(HGA_ERR_INVALID, "leading dimension < number of columns");
  if ((mu_dev == nullptr) != (sigma_dev == nullptr)) return fail(HGA_ERR_INVALID, "mu and sigma must both be given or both be NULL");
  if (n_segments < 1) return HGA_OK;
  for (int i = 0; i < n_segments; ++i)
    if (seg_len[i] < L) return fail(HGA_ERR_INVALID, "segment %d has %lld rows < frame_len %d", i, (long long)seg_len[i], L);
  CU(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  int64_t xlo = INT64_MAX, xhi = 0;
  for (int i = 0; i < n_segments; ++i) { xlo = std::min(xlo, seg_start[i]); xhi = std::max(xhi, seg_start[i] + seg_len[i]); }
  int rc;
  if (h->cfg.precision == HGA_PREC_F32 && h->lane_ok && hop == 10 && h->nb > 0) {
    if ((rc = lane_prepare(h, n_segments, seg_start, seg_len, out_row, xlo, st))) return rc;
    LaneMapDev m;
    m.raw = x_raw_dev + xlo * ldx;
    m.ld_raw = ldx;
    m.raw_rows = xhi - xlo;
    m.C_in = map->c_in;
    m.n_groups = map->n_groups;
    m.n_idx = map->n_idx;
    m.sel = map->sel; m.grp = map->grp; m.gptr = map->gptr; m.gidx = map->gidx; m.alpha = map->alpha; m.beta = map->beta;
    LaneDev ld = lane_dev(h);
    h->timed_main = false;
    if (h->timing) { CU(cudaEventRecord(h->ev0, st)); ld.ev_main = h->ev_main; }
    const int n = lanes_launch_map_t<LaneF32, 10, 20, 2, 2, 2, 3, 5>(h->lk, h->lane_plan_host, ld, m, C, mu_dev, sigma_dev, out_dev, ldo, h->n_sm, st);
    if (n == -1) return fail(HGA_ERR_CUDA, "fused lane kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (n > 0) {
      g_launches.fetch_add(n, std::memory_order_relaxed);
      if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; h->timed_main = true; }
      if (fused) *fused = 1;
      return HGA_OK;
    }
  }
  // the map as a pass of its own over the rows in use, then the plain extraction
  const int64_t ldp = (C + 1) / 2 * 2;
  if ((rc = h->xpad.reserve((size_t)(xhi - xlo) * ldp * 8))) return rc;
  if ((rc = hga_chanmap_apply(map, x_raw_dev + xlo * ldx, xhi - xlo, ldx, h->xpad.as<double>(), ldp, stream))) return rc;
  return extract_offline_dev(h, h->xpad.as<double>() - xlo * ldp, ldp, n_segments, seg_start, seg_len, out_row, mu_dev, sigma_dev, out_dev, ldo, st, nullptr,
                             nullptr);
}

// ------------------------------------------------------------------------------------------------ offline, host buffers
// Upload x in time slabs on the copy stream; after each slab, launch the work items whose rows are resident; download
// each group's frame rows on the handle's own stream while the next slab uploads.
//
// Raw ingest (map != nullptr): the host rows are amplifier counts of `dtype`, c_in columns wide; each slab is uploaded at its
// raw element size and expanded on the copy stream by the channel-map kernel (signal * gain, then the folded
// pre-transform chain) into the float64 rows the extraction kernels read.
static size_t raw_elem_size(int32_t dtype) { return dtype == HGA_RAW_F64 ? 8 : dtype == HGA_RAW_I16 ? 2 : 4; }

// Host memory the DMA engine can read where it lies (cudaHostAlloc / cudaHostRegister)?
static bool host_is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}
// rows x row_bytes from a strided host matrix into a compact buffer (or back), split over a few threads: one core moves
// ~10 GB/s, the PCIe link 50+
static void host_rows_copy(unsigned char* dst, size_t dst_pitch, const unsigned char* src, size_t src_pitch, size_t row_bytes, int64_t rows) {
  const int64_t bytes = rows * (int64_t)row_bytes;
  const int nt = (int)std::max<int64_t>(1, std::min<int64_t>({16, (int64_t)std::thread::hardware_concurrency(), bytes >> 22}));
  auto part = [=](int64_t r0, int64_t r1) {
    if (dst_pitch == row_bytes && src_pitch == row_bytes) memcpy(dst + r0 * row_bytes, src + r0 * row_bytes, (size_t)(r1 - r0) * row_bytes);
    else for (int64_t r = r0; r < r1; ++r) memcpy(dst + r * dst_pitch, src + r * src_pitch, row_bytes);
  };
  if (nt == 1) { part(0, rows); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) th.emplace_back(part, rows * t / nt, rows * (t + 1) / nt);
  for (auto& t : th) t.join();
}
constexpr size_t kStageBytes = 64u << 20;   // pinned staging buffers of the pageable path (two, ping-pong)

static int extract_offline_host_impl(hga_handle* h, hga_chanmap* map, const void* x_host_v, int32_t dtype, const double* gain_host, int64_t ldx,
                                     int32_t n_segments, const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                                     const double* mu_host, const double* sigma_host, double* out_host, int64_t ldo) {
  if (!h || !x_host_v || !out_host || !seg_start || !seg_len || !out_row) return fail(HGA_ERR_INVALID, "null argument");
  const int C = h->cfg.n_channels, L = h->cfg.frame_len, hop = h->cfg.hop;
  const int c_in = map ? map->c_in : C;
  if (map && (map->c_out != C || map->device != h->cfg.device))
    return fail(HGA_ERR_INVALID, "channel map yields %d columns on device %d, the extractor wants %d on device %d", map->c_out, map->device, C, h->cfg.device);
  if (dtype < HGA_RAW_F64 || dtype > HGA_RAW_I16) return fail(HGA_ERR_INVALID, "unknown raw dtype %d", dtype);
  if (ldx < c_in || ldo < C) return fail(HGA_ERR_INVALID, "leading dimension < n_channels");
  if (n_segments < 1) return HGA_OK;
  CU(cudaSetDevice(h->cfg.device));
  int rc;
  const size_t esz = raw_elem_size(dtype), raw_row = (size_t)c_in * esz;
  const unsigned char* x_raw = static_cast<const unsigned char*>(x_host_v);
  int64_t lo = INT64_MAX, hi = 0, olo = INT64_MAX, ohi = 0;
  for (int i = 0; i < n_segments; ++i) {
    if (seg_len[i] < L) return fail(HGA_ERR_INVALID, "segment %d has %lld rows < frame_len %d", i, (long long)seg_len[i], L);
    lo = std::min(lo, seg_start[i]);
    hi = std::max(hi, seg_start[i] + seg_len[i]);
    olo = std::min(olo, out_row[i]);
    ohi = std::max(ohi, out_row[i] + (seg_len[i] - L) / hop + 1);
  }
  const size_t row_bytes = (size_t)C * 8;
  const int64_t ldd = h->lane_ok ? (C + 1) / 2 * 2 : C;   // device leading dimension: even, so that the rows are a TMA tensor
  if ((rc = h->xdev.reserve((size_t)(hi - lo) * ldd * 8)) || (rc = h->outdev.reserve((size_t)(ohi - olo) * row_bytes))) return rc;
  cudaStream_t sc = h->s_copy, sk = h->s_own;
  const double *mu_d, *sig_d;
  if ((rc = sync_musigma(h, mu_host, sigma_host, sk, &mu_d, &sig_d))) return rc;
  double* xd = h->xdev.as<double>() - lo * ldd;   // device row r <-> host row r
  double* od = h->outdev.as<double>() - olo * C;  // device out row r <-> host out row r
  const bool fused = h->cfg.precision != HGA_PREC_EXACT && h->nb > 0;
  const double* gain_d = nullptr;
  if (map) {
    if ((rc = h->rawdev.reserve((size_t)(hi - lo) * raw_row + (gain_host ? (size_t)c_in * 8 : 0) + 16))) return rc;
    if (gain_host) {
      const size_t off = ((size_t)(hi - lo) * raw_row + 15) / 16 * 16;
      gain_d = reinterpret_cast<const double*>(static_cast<unsigned char*>(h->rawdev.p) + off);
      CU(cudaMemcpyAsync(const_cast<double*>(gain_d), gain_host, (size_t)c_in * 8, cudaMemcpyHostToDevice, sc));
    }
  }
  // Pageable caller buffers (what BCI2000MatFile.signals() returns, local/common.py:458-461) go through two pinned staging
  // buffers: the host copy of piece k + 1 overlaps the DMA of piece k.  Page-locked buffers are read where they lie.
  const bool in_pinned = host_is_pinned(x_host_v), out_pinned = host_is_pinned(out_host);
  if (!in_pinned) {
    for (int b = 0; b < 2; ++b) {
      if ((rc = h->stage_up[b].reserve(kStageBytes))) return rc;
      if (!h->stage_ev[b]) CU(cudaEventCreateWithFlags(&h->stage_ev[b], cudaEventDisableTiming));
    }
  }
  if (!out_pinned && (rc = h->stage_down.reserve((size_t)(ohi - olo) * row_bytes))) return rc;
  int stage_b = 0;
  // rows [r0, r1) of the host recording -> device rows (raw element type when a map follows), enqueued on stream st
  auto host_to_dev = [&](unsigned char* dev, size_t dpitch, int64_t r0, int64_t r1, cudaStream_t st) -> int {
    const unsigned char* src = x_raw + (size_t)r0 * ldx * esz;
    const size_t spitch = (size_t)ldx * esz;
    if (in_pinned) {
      if (spitch == raw_row && dpitch == raw_row) CU(cudaMemcpyAsync(dev, src, (size_t)(r1 - r0) * raw_row, cudaMemcpyHostToDevice, st));   // one linear copy
      else CU(cudaMemcpy2DAsync(dev, dpitch, src, spitch, raw_row, (size_t)(r1 - r0), cudaMemcpyHostToDevice, st));
      return HGA_OK;
    }
    const int64_t piece = std::max<int64_t>(1, (int64_t)(kStageBytes / raw_row));
    for (int64_t r = r0; r < r1; r += piece) {
      const int64_t n = std::min(piece, r1 - r);
      CU(cudaEventSynchronize(h->stage_ev[stage_b]));   // the previous DMA out of this staging buffer has finished
      unsigned char* pin = h->stage_up[stage_b].as<unsigned char>();
      host_rows_copy(pin, raw_row, src + (size_t)(r - r0) * spitch, spitch, raw_row, n);
      if (dpitch == raw_row) CU(cudaMemcpyAsync(dev + (size_t)(r - r0) * raw_row, pin, (size_t)n * raw_row, cudaMemcpyHostToDevice, st));
      else CU(cudaMemcpy2DAsync(dev + (size_t)(r - r0) * dpitch, dpitch, pin, raw_row, raw_row, (size_t)n, cudaMemcpyHostToDevice, st));
      CU(cudaEventRecord(h->stage_ev[stage_b], st));
      stage_b ^= 1;
    }
    return HGA_OK;
  };
  auto upload = [&](int64_t r0, int64_t r1, cudaStream_t st) -> int {
    int rc2;
    if (!map) return host_to_dev(reinterpret_cast<unsigned char*>(xd + r0 * ldd), (size_t)ldd * 8, r0, r1, st);
    unsigned char* rd = static_cast<unsigned char*>(h->rawdev.p) + (size_t)(r0 - lo) * raw_row;
    if ((rc2 = host_to_dev(rd, raw_row, r0, r1, st))) return rc2;
    return hga_chanmap_apply_raw(map, rd, dtype, gain_d, r1 - r0, c_in, xd + r0 * ldd, ldd, st);
  };
  // frame rows [f0, f1) of the device result -> the caller's buffer (pageable: into the pinned staging, copied out at the end)
  unsigned char* down = out_pinned ? nullptr : h->stage_down.as<unsigned char>();
  auto download = [&](int64_t f0, int64_t f1, cudaStream_t st) -> int {
    if (f1 <= f0) return HGA_OK;
    if (out_pinned) CU(cudaMemcpy2DAsync(out_host + f0 * ldo, (size_t)ldo * 8, od + f0 * C, row_bytes, row_bytes, (size_t)(f1 - f0), cudaMemcpyDeviceToHost, st));
    else CU(cudaMemcpyAsync(down + (size_t)(f0 - olo) * row_bytes, od + f0 * C, (size_t)(f1 - f0) * row_bytes, cudaMemcpyDeviceToHost, st));
    return HGA_OK;
  };
  // the same for a set of ranges: touching ranges travel as one copy, gaps between segments' outputs are never written
  auto download_ranges = [&](std::vector<std::pair<int64_t, int64_t>>& r, cudaStream_t st) -> int {
    std::sort(r.begin(), r.end());
    for (size_t i = 0; i < r.size();) {
      int64_t a0 = r[i].first, a1 = r[i].second;
      size_t j = i + 1;
      while (j < r.size() && r[j].first <= a1) { a1 = std::max(a1, r[j].second); ++j; }
      const int rc2 = download(a0, a1, st);
      if (rc2) return rc2;
      i = j;
    }
    return HGA_OK;
  };
  auto finish = [&]() -> int {
    CU(cudaStreamSynchronize(sk));
    CU(cudaStreamSynchronize(sc));
    if (!out_pinned)   // per segment: rows between the segments' output ranges are not the caller's to overwrite
      for (int i = 0; i < n_segments; ++i) {
        const int64_t f0 = out_row[i], nf = (seg_len[i] - L) / hop + 1;
        host_rows_copy(reinterpret_cast<unsigned char*>(out_host + f0 * ldo), (size_t)ldo * 8, down + (size_t)(f0 - olo) * row_bytes, row_bytes, row_bytes, nf);
      }
    return HGA_OK;
  };
  const int variant = fused ? choose_variant(h, xd + lo * ldd, ldd, true) : -2;
  if (variant == -3) return fail(HGA_ERR_UNSUPPORTED, "%s", kLaneOnlyMsg);

  if (variant == -2) {  // exact arithmetic: one upload, the unfused kernels, per-segment download
    if (map) CU(cudaStreamSynchronize(sc));   // the gain vector went up on the copy stream
    if ((rc = upload(lo, hi, sk))) return rc;
    if ((rc = extract_offline_dev(h, xd, ldd, n_segments, seg_start, seg_len, out_row, mu_d, sig_d, od, C, sk, nullptr, nullptr))) return rc;
    for (int i = 0; i < n_segments; ++i)
      if ((rc = download(out_row[i], out_row[i] + (seg_len[i] - L) / hop + 1, sk))) return rc;
    return finish();
  }
  // Time slabs: slab k + 1 goes up while the kernels run on the chunks of slab k and finished frames go down.  Each partial
  // launch runs its items at the full per-item duration, so the slab count stays small (float64: 12, int16 counts: 5).
  const int64_t total_bytes = (hi - lo) * (int64_t)row_bytes;
  const int64_t max_slabs = map ? std::max<int64_t>(2, std::min<int64_t>(12, (int64_t)(2.7 * (double)esz))) : 12;
  int64_t slab_bytes = std::min<int64_t>(std::max<int64_t>(total_bytes / max_slabs, 16LL << 20), (512LL << 20) * 8 / (int64_t)esz);
  const int64_t slab_rows = std::max<int64_t>(hop, slab_bytes / (int64_t)row_bytes / hop * hop);
  const size_t n_slabs = (size_t)((hi - lo + slab_rows - 1) / slab_rows);
  while (h->slab_ev.size() < 2 * n_slabs + 2) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->slab_ev.push_back(e);
  }
  size_t ev = 0;

  if (variant == 2) {
    std::vector<int> by_time;
    if ((rc = lane_prepare(h, n_segments, seg_start, seg_len, out_row, lo, sk, &by_time))) return rc;
    const LanePlan& p = h->lane_plan_host;
    CUtensorMap tm;
    if ((rc = lane_tensor(h, xd, ldd, lo, hi, &tm))) return rc;
    LaneDev d = lane_dev(h);
    d.x0 = xd + lo * ldd;
    d.ldx = ldd;
    const int* order_time = h->lane_order.as<int>() + p.order.size();
    d.skip_fix = true;
    size_t next = 0;
    for (int64_t r0 = lo; r0 < hi; r0 += slab_rows) {
      const int64_t r1 = std::min(hi, r0 + slab_rows);
      if ((rc = upload(r0, r1, sc))) return rc;
      size_t end = next;
      while (end < by_time.size()) {
        const LaneWork& w = p.work[(size_t)by_time[end] / p.groups];
        if (lo + (int64_t)w.x_row + w.rows > r1) break;
        ++end;
      }
      if (r1 == hi) end = by_time.size();
      if (end == next) continue;
      cudaEvent_t e_up = h->slab_ev[ev++];
      CU(cudaEventRecord(e_up, sc));
      CU(cudaStreamWaitEvent(sk, e_up, 0));
      d.order = order_time + next;
      d.n_items = (int)(end - next);
      if ((rc = lane_launch(h->cfg.precision, hop, h->lk, p, d, tm, C, mu_d, sig_d, od, C, h->n_sm, sk))) return rc;
      std::vector<std::pair<int64_t, int64_t>> ranges;   // frames these items finished themselves
      for (size_t i = next; i < end; ++i) {
        const LaneWork& w = p.work[(size_t)by_time[i] / p.groups];
        const int64_t a0 = w.out_row0 + w.blk0 + w.first_clean, a1 = w.out_row0 + w.blk0 + w.rows / hop - (kLaneNB - 1);
        if (a1 > a0) ranges.emplace_back(a0, a1);
      }
      if ((rc = download_ranges(ranges, sk))) return rc;
      next = end;
    }
    if (!p.fix.empty()) {  // carried heads: their frames go down after the fix-up
      d.skip_fix = false;
      if ((rc = lane_fix_only(h->cfg.precision, hop, h->lk, p, d, C, mu_d, sig_d, od, C, sk))) return rc;
      std::vector<std::pair<int64_t, int64_t>> ranges;
      for (const LaneFix& f : p.fix) {
        const int64_t a0 = f.out_row0 + f.blk0 - (kLaneNB - 1);
        ranges.emplace_back(a0, a0 + f.n_frames);
      }
      if ((rc = download_ranges(ranges, sk))) return rc;
    }
    return finish();
  }

  // generic float64 kernel: plan once for the whole table, launch sub-ranges of the work list
  {
    std::vector<int64_t> key;
    key.insert(key.end(), seg_start, seg_start + n_segments);
    key.insert(key.end(), seg_len, seg_len + n_segments);
    key.insert(key.end(), out_row, out_row + n_segments);
    if (key != h->plan_key || variant != h->plan_variant) {
      plan_chunks(h, n_segments, seg_start, seg_len, out_row, std::max<int64_t>(1, (int64_t)h->n_sm * 12), (C + 31) / 32, true);
      if ((rc = upload_plan(h, sk))) return rc;
      h->plan_key = key;
      h->plan_variant = variant;
    }
  }
  const std::vector<Work>& W = h->work_host;
  size_t next = 0;
  for (int64_t r0 = lo; r0 < hi; r0 += slab_rows) {
    const int64_t r1 = std::min(hi, r0 + slab_rows);
    if ((rc = upload(r0, r1, sc))) return rc;
    size_t end = next;
    while (end < W.size() && W[end].row0 + W[end].t_end <= r1 && W[end].row0 + W[end].t_begin >= lo) ++end;
    if (r1 == hi) end = W.size();
    if (end == next) continue;
    cudaEvent_t e_up = h->slab_ev[ev++];
    CU(cudaEventRecord(e_up, sc));
    CU(cudaStreamWaitEvent(sk, e_up, 0));
    FusedArgs a{xd, ldd, h->work.as<Work>() + next, (int)(end - next), C, L, hop, h->nb, h->cfg.eps, mu_d, sig_d, od, C, nullptr, nullptr, nullptr, nullptr, lo, hi};
    if (is_unit_std(h->q)) rc = launch_fused_seq<kUnitStd>(h, a, sk);
    else if (h->q.unit) rc = launch_fused_seq<kUnit>(h, a, sk);
    else rc = launch_fused_seq<kFused>(h, a, sk);
    if (rc) return rc;
    std::vector<std::pair<int64_t, int64_t>> ranges;
    for (size_t i = next; i < end; ++i) {
      const Work& w = W[i];
      ranges.emplace_back(w.out_row0 + w.t_acc / hop, w.out_row0 + std::max<int64_t>(0, w.t_end / hop - (h->nb - 1)));
    }
    if ((rc = download_ranges(ranges, sk))) return rc;
    next = end;
  }
  return finish();
}
extern "C" int hga_extract_offline_host(hga_handle* h, const double* x_host, int64_t ldx, int32_t n_segments, const int64_t* seg_start,
                                        const int64_t* seg_len, const int64_t* out_row, const double* mu_host, const double* sigma_host,
                                        double* out_host, int64_t ldo) {
  return extract_offline_host_impl(h, nullptr, x_host, HGA_RAW_F64, nullptr, ldx, n_segments, seg_start, seg_len, out_row, mu_host, sigma_host, out_host, ldo);
}
extern "C" int hga_extract_offline_raw_host(hga_handle* h, hga_chanmap* map, const void* x_host, int32_t dtype, const double* gain_host, int64_t ldx,
                                            int32_t n_segments, const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                                            const double* mu_host, const double* sigma_host, double* out_host, int64_t ldo) {
  if (!map) return fail(HGA_ERR_INVALID, "null argument");
  return extract_offline_host_impl(h, map, x_host, dtype, gain_host, ldx, n_segments, seg_start, seg_len, out_row, mu_host, sigma_host, out_host, ldo);
}
