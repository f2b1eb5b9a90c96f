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
#include <string>
#include <vector>

#include "hga_common.cuh"
#include "hga_kernels_exact.cuh"
#include "hga_kernels_fused.cuh"
#include "hga_kernels_pipe2.cuh"
#include "hga_kernels_pre.cuh"

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
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return HGA_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(HGA_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    }
    cap = bytes;
    return HGA_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T>
  T* as() const { return static_cast<T*>(p); }
};
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
// memory (staged into shared memory) and writes the frames to mapped pinned host memory: one launch + one stream
// synchronisation per packet instead of H2D copy + launch + D2H copy + synchronisation.  HGA_B200_STREAM_MAPPED=0 disables.
constexpr int64_t kMappedRows = 64;
static bool stream_flagged() {  // HGA_B200_STREAM_FLAG=0: always synchronise the stream instead of polling completion slots
  static const bool on = [] { const char* e = getenv("HGA_B200_STREAM_FLAG"); return !(e && atoi(e) == 0); }();
  return on;
}
static bool stream_mapped() {
  static const bool on = [] { const char* e = getenv("HGA_B200_STREAM_MAPPED"); return !(e && atoi(e) == 0); }();
  return on;
}

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

// ------------------------------------------------------------------------------------------------ pipe kernel host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return (EncodeTiledFn)p;
  }();
  return fn;
}
static void build_pipe_const(const Cascade& q, int L, PipeConst& pc) {
  memset(&pc, 0, sizeof pc);
  double G = 1.0;
  for (int s = 0; s < q.ns; ++s) {
    G *= q.g[s];
    pc.zu0[s] = q.zi0[s] / G;
    pc.zu1[s] = q.zi1[s] / G;
    auto pack = [](double v) {
      const float f = (float)v;
      uint32_t u;
      memcpy(&u, &f, 4);
      return ((unsigned long long)u << 32) | u;
    };
    pc.na1p[s] = pack(-q.a1[s]);
    pc.na2p[s] = pack(-q.a2[s]);
    pc.b1p[s] = pack(q.b1u[s]);
  }
  pc.pow_scale = G * G / (double)L;
}
// Operands of k_hga_pipe2 (hga_kernels_pipe2.cuh, "section forms").  form56: the band-pass numerators are re-paired into
// [1, 0, -1] sections (possible when its sections are n x [1, 2, 1] and n x [1, -2, 1], which is what scipy designs) and
// the sections are interleaved so that every filter warp of the kernel gets the same mix: with NS sections per warp,
// NS/2 band-pass (A) sections followed by NS/2 notch (U) sections.  Otherwise (or HGA_B200_FORM=64): all sections in
// the U form, in the caller's order.  Section order is free because the cascade is LTI and always starts from zero state.
static bool can_form56(const Cascade& q) {
  if (!is_unit_std(q)) return false;
  int plus = 0, minus = 0;
  for (int s = 0; s < q.ns_a; ++s) {
    if (std::fabs(q.b1u[s] - 2.0) <= 1e-13) ++plus;
    else if (std::fabs(q.b1u[s] + 2.0) <= 1e-13) ++minus;
    else return false;
  }
  if (const char* e = getenv("HGA_B200_FORM")) if (atoi(e) == 64) return false;
  return plus == minus;
}
static void build_pipe2_const(const Cascade& q, int L, int NS, PipeConst& pc) {
  build_pipe_const(q, L, pc);
  const bool f56 = can_form56(q);
  auto pack = [](double v) {
    const float f = (float)v;
    uint32_t u;
    memcpy(&u, &f, 4);
    return ((unsigned long long)u << 32) | u;
  };
  pc.amask = 0;
  for (int i = 0; i < q.ns; ++i) {
    int src = i;
    bool A = false;
    if (f56) {  // slot i of warp i / NS: first half A sections, second half U sections
      const int g = i / NS, k = i % NS, half = NS / 2;
      A = k < half;
      src = A ? g * half + k : q.ns_a + g * half + (k - half);
    }
    pc.c1d[i] = A ? q.a1[src] : -q.a1[src];
    pc.c2d[i] = A ? q.a2[src] : -q.a2[src];
    pc.cbd[i] = A ? 0.0 : q.b1u[src];
    pc.na1p[i] = pack(pc.c1d[i]);
    pc.na2p[i] = pack(pc.c2d[i]);
    pc.b1p[i] = pack(pc.cbd[i]);
    if (A) pc.amask |= 1u << i;
  }
}
// Zero-input response of the caller's initial state (sosfilt_zi of both filters, local/units.py:128-132) through the
// caller's cascade, exact DF2T order, divided by the gain product: what k_hga_pipe2 adds to the rows of a fresh item.
static std::vector<double> zero_input_response(const Cascade& q, int n) {
  std::vector<double> r((size_t)n), z0(q.zi0, q.zi0 + kMaxSec), z1(q.zi1, q.zi1 + kMaxSec);
  double G = 1.0;
  for (int s = 0; s < q.ns; ++s) G *= q.g[s];
  for (int t = 0; t < n; ++t) {
    double x = 0.0;
    for (int s = 0; s < q.ns; ++s) {
      const double y = q.b0[s] * x + z0[s];
      z0[s] = (q.b1[s] * x - q.a1[s] * y) + z1[s];
      z1[s] = q.b2[s] * x - q.a2[s] * y;
      x = y;
    }
    r[t] = x / G;
  }
  return r;
}
// log table of tab_log (hga_kernels_pipe2.cuh): entry i = {rc, -log(rc)} with rc = double(1 / (1 + (i + 1/2) / 128))
static int upload_log_table() {
  double tab[2 * kLogTabN];
  for (int i = 0; i < kLogTabN; ++i) {
    const double rc = (double)(1.0L / (1.0L + ((long double)i + 0.5L) / (long double)kLogTabN));
    tab[2 * i] = rc;
    tab[2 * i + 1] = (double)(-logl((long double)rc));
  }
  CU(cudaMemcpyToSymbol(g_logtab, tab, sizeof tab));
  return HGA_OK;
}
struct PipeShape {
  int ver, S, TB;
};
// Default: k_hga_pipe2 with 4 filter warps x 4 sections, one CTA per SM; 80-row tiles with 2-deep rings in float64 mode
// (K2: 40 rows 183 G, 60 rows 196 G, 80 rows 201 G channel-samples/s: a tile boundary costs ~300 cycles per warp), 40-row
// tiles in float32 mode (its 20 KB x tiles leave no room for more).  HGA_B200_PIPE_V=1 selects the first-generation kernel (k_hga_pipe,
// S in {2, 4}); HGA_B200_PIPE_S / HGA_B200_PIPE_TB pick the other compiled shapes (A/B runs).
static PipeShape pipe_shape(bool f32 = false) {
  PipeShape p{2, 4, f32 ? 40 : 80};
  if (const char* e = getenv("HGA_B200_PIPE_V")) p.ver = atoi(e) == 1 ? 1 : 2;
  if (p.ver == 1) { p.S = 4; p.TB = 40; }
  if (const char* e = getenv("HGA_B200_PIPE_S")) p.S = atoi(e);
  if (const char* e = getenv("HGA_B200_PIPE_TB")) p.TB = atoi(e);
  if (p.ver == 1) {
    if (p.S != 2 && p.S != 4) p.S = 4;
    if (p.TB != 20 || p.S != 4) p.TB = 40;
  } else {
    if (p.S != 4 && p.S != 8) p.S = 4;
    if (p.S == 8) p.TB = 20;
    else if (p.TB != 20 && p.TB != 40 && p.TB != 60 && !((p.TB == 70 || p.TB == 80) && !f32)) p.TB = 40;  // 70 / 80-row tiles: float64 only
  }
  return p;
}
// true the first time the calling instantiation runs on the current device (a process may hold handles on several devices;
// cudaFuncSetAttribute applies to the current device only)
static bool first_use_on_device(std::atomic<uint64_t>& mask) {
  int d = 0;
  cudaGetDevice(&d);
  const uint64_t bit = 1ull << (d & 63);
  return !(mask.fetch_or(bit) & bit);
}
template <typename T, int S, int TB, int HOP>
static int pipe_launch_t(const Cascade& q, const PipeConst& pc, const FusedArgs& a, const CUtensorMap& tm, cudaStream_t st, int* occ) {
  auto kern = k_hga_pipe<T, S, TB, HOP>;
  const int smem = (int)PipeSmem<S, TB, T>::total;
  static std::atomic<uint64_t> attr_done{0};   // function attributes are per device: one bit per ordinal
  if (first_use_on_device(attr_done)) CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (occ) {
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, kern, 64 * S, smem));
    return HGA_OK;
  }
  constexpr int CW = 32 * PipeTraits<T>::CPL;
  const int cg = (a.C + CW - 1) / CW;
  kern<<<(unsigned)((int64_t)a.n_work * cg), 64 * S, smem, st>>>(q, pc, a, tm);
  LAUNCHED();
  return HGA_OK;
}
// framing warps of k_hga_pipe2: 4 (one per SM sub-partition) or 8 (S = 4, TB = 40 only; HGA_B200_PIPE_NF)
static int pipe2_nf(bool f32) {
  int nf = f32 ? 8 : 4;
  if (const char* e = getenv("HGA_B200_PIPE_NF")) nf = atoi(e) == 8 ? 8 : 4;
  return nf;
}
template <typename T, int S, int TB, int HOP, unsigned AM, int kPipe2NF = 4>
static int pipe2_launch_am(const Cascade& q, const PipeConst& pc, const FusedArgs& a, const CUtensorMap& tm, cudaStream_t st, int* occ) {
  // S == 4: one section group per SM sub-partition -> per-group code with constant-bank coefficients; S == 8: shared code.
  constexpr bool kPerG = S == 4;
  auto kern = k_hga_pipe2<T, S, TB, HOP, kPipe2NF, kPerG, AM>;
  const int smem = (int)Pipe2Smem<S, TB, T>::total, threads = 32 * (S + kPipe2NF + 1);
  static std::atomic<uint64_t> attr_done{0};   // function attributes are per device: one bit per ordinal
  if (first_use_on_device(attr_done)) CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (occ) {
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, kern, threads, smem));
    return HGA_OK;
  }
  constexpr int CW = 32 * PipeTraits<T>::CPL;
  const int cg = (a.C + CW - 1) / CW;
  kern<<<(unsigned)((int64_t)a.n_work * cg), threads, smem, st>>>(q, pc, a, tm);
  LAUNCHED();
  return HGA_OK;
}
template <typename T, int S, int TB, int HOP>
static int pipe2_launch_t(const Cascade& q, const PipeConst& pc, const FusedArgs& a, const CUtensorMap& tm, cudaStream_t st, int* occ) {
  constexpr int NS = kMaxSec / S;
  constexpr unsigned kHalf = (1u << (NS / 2)) - 1u;  // form56: the first NS/2 sections of every warp are A sections
  if (S == 4 && (TB == 40 || TB == 60) && pc.amask != 0 && pipe2_nf(PipeTraits<T>::CPL == 2) == 8)
    return pipe2_launch_am<T, S, TB, HOP, kHalf, (S == 4 && (TB == 40 || TB == 60) ? 8 : 4)>(q, pc, a, tm, st, occ);
  if (pc.amask != 0) return pipe2_launch_am<T, S, TB, HOP, kHalf>(q, pc, a, tm, st, occ);
  return pipe2_launch_am<T, S, TB, HOP, 0u>(q, pc, a, tm, st, occ);
}
// occ != nullptr: only query the resident CTAs per SM of the variant that would be launched.
template <typename T>
static int pipe_dispatch(const Cascade& q, const PipeConst& pc, const FusedArgs& a, int hop, const CUtensorMap& tm, cudaStream_t st, int* occ) {
  const PipeShape p = pipe_shape(PipeTraits<T>::CPL == 2);
  if (p.ver == 2) {
    if (hop == 10) {
      if (p.S == 8) return pipe2_launch_t<T, 8, 20, 10>(q, pc, a, tm, st, occ);
      if (p.TB == 20) return pipe2_launch_t<T, 4, 20, 10>(q, pc, a, tm, st, occ);
      if (p.TB == 60) return pipe2_launch_t<T, 4, 60, 10>(q, pc, a, tm, st, occ);
      if constexpr (PipeTraits<T>::CPL == 1) {
        if (p.TB == 70) return pipe2_launch_t<T, 4, 70, 10>(q, pc, a, tm, st, occ);
        if (p.TB == 80) return pipe2_launch_t<T, 4, 80, 10>(q, pc, a, tm, st, occ);
      }
      return pipe2_launch_t<T, 4, 40, 10>(q, pc, a, tm, st, occ);
    }
    if (p.S == 8) return pipe2_launch_t<T, 8, 20, 20>(q, pc, a, tm, st, occ);
    if (p.TB == 20) return pipe2_launch_t<T, 4, 20, 20>(q, pc, a, tm, st, occ);
    if constexpr (PipeTraits<T>::CPL == 1) {
      if (p.TB >= 80) return pipe2_launch_t<T, 4, 80, 20>(q, pc, a, tm, st, occ);
    }
    if (p.TB >= 60) return pipe2_launch_t<T, 4, 60, 20>(q, pc, a, tm, st, occ);
    return pipe2_launch_t<T, 4, 40, 20>(q, pc, a, tm, st, occ);
  }
  if (hop == 10) {
    if (p.S == 2) return pipe_launch_t<T, 2, 40, 10>(q, pc, a, tm, st, occ);
    if (p.TB == 20) return pipe_launch_t<T, 4, 20, 10>(q, pc, a, tm, st, occ);
    return pipe_launch_t<T, 4, 40, 10>(q, pc, a, tm, st, occ);
  }
  if (p.S == 2) return pipe_launch_t<T, 2, 40, 20>(q, pc, a, tm, st, occ);
  if (p.TB == 20) return pipe_launch_t<T, 4, 20, 20>(q, pc, a, tm, st, occ);
  return pipe_launch_t<T, 4, 40, 20>(q, pc, a, tm, st, occ);
}
static bool pipe_supported(const Cascade& q, int hop, int nb) {
  return is_unit_std(q) && (hop == 10 || hop == 20) && nb == kPipeNB && encode_tiled_fn() != nullptr;
}
// x must be 16-byte aligned with an even leading dimension for the TMA descriptor
static bool pipe_layout_ok(const double* x, int64_t ldx) { return ((uintptr_t)x % 16) == 0 && (ldx % 2) == 0; }
struct PipeOperands {  // what the handle keeps for the pipe kernels
  PipeConst pc;    // first-generation kernel (k_hga_pipe): caller's sections and initial state
  PipeConst pc2;   // k_hga_pipe2: its own section order / forms, zero initial state
  const double* zir = nullptr;  // device, n_zir rows
  int n_zir = 0;
};
static int launch_pipe(const Cascade& q, const PipeOperands& po, int precision, const FusedArgs& a_in, cudaStream_t st) {
  const bool v2 = pipe_shape().ver == 2;
  const PipeConst& pc = v2 ? po.pc2 : po.pc;
  FusedArgs a = a_in;
  a.zir = po.zir;
  a.n_zir = po.n_zir;
  CUtensorMap tm;
  const cuuint64_t dims[2] = {(cuuint64_t)a.C, (cuuint64_t)(a.x_row_hi - a.x_row_lo)};
  const cuuint64_t strides[1] = {(cuuint64_t)a.ldx * 8};
  const PipeShape p = pipe_shape(precision == HGA_PREC_F32);
  const int tb_eff = (a.hop == 20 && p.ver == 2 && p.TB == 70) ? 60 : p.TB;  // 70 rows are not a whole number of 20-row hop blocks
  const cuuint32_t box[2] = {precision == HGA_PREC_F32 ? 64u : 32u, (cuuint32_t)tb_eff};
  const cuuint32_t es[2] = {1, 1};
  void* base = (void*)(a.x + a.x_row_lo * a.ldx);
  const CUresult r = encode_tiled_fn()(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(HGA_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  if (getenv("HGA_B200_PROF")) {  // debugging aid: per-warp cycle breakdown, printed to stderr (synchronises)
    const int cw_ = precision == HGA_PREC_F32 ? 64 : 32;
    const int wstride = p.ver == 2 ? 16 : 8;
    const size_t n_cta = (size_t)a.n_work * ((a.C + cw_ - 1) / cw_), n = n_cta * wstride * 6;
    long long* d = nullptr;
    CU(cudaMalloc(&d, n * 8));
    CU(cudaMemset(d, 0, n * 8));
    FusedArgs b = a;
    b.prof = d;
    int rc = precision == HGA_PREC_F32 ? pipe_dispatch<f32x2>(q, pc, b, b.hop, tm, st, nullptr) : pipe_dispatch<double>(q, pc, b, b.hop, tm, st, nullptr);
    CU(cudaStreamSynchronize(st));
    std::vector<long long> hbuf(n);
    CU(cudaMemcpy(hbuf.data(), d, n * 8, cudaMemcpyDeviceToHost));
    cudaFree(d);
    const char* names[6] = {"role", "wait_in", "phaseA", "phaseB", "wait_out", "total"};
    for (int g = 0; g < wstride; ++g) {
      double acc[6] = {0};
      for (size_t cta = 0; cta < n_cta; ++cta)
        for (int i = 0; i < 6; ++i) acc[i] += (double)hbuf[(cta * wstride + g) * 6 + i];
      if (acc[5] == 0) continue;
      fprintf(stderr, "[hga prof] warp %d:", g);
      for (int i = 0; i < 6; ++i) fprintf(stderr, " %s=%.0f", names[i], acc[i] / n_cta);
      fprintf(stderr, "  (cycles per CTA, mean of %zu CTAs)\n", n_cta);
    }
    return rc;
  }
  return precision == HGA_PREC_F32 ? pipe_dispatch<f32x2>(q, pc, a, a.hop, tm, st, nullptr) : pipe_dispatch<double>(q, pc, a, a.hop, tm, st, nullptr);
}
// HGA_B200_FRAMES_ROLL=0 selects the per-window kernel (k_frames_seq) for the regular geometry too (A/B and tests).
static bool frames_roll_on() {
  const char* e = getenv("HGA_B200_FRAMES_ROLL");
  return !(e && e[0] == '0');
}
static int64_t pipe_item_slots(const Cascade& q, const PipeOperands& po, int precision, int hop, int n_sm, int C) {
  const PipeConst& pc = pipe_shape().ver == 2 ? po.pc2 : po.pc;
  int occ = 1;
  FusedArgs a{};
  CUtensorMap tm{};
  if ((precision == HGA_PREC_F32 ? pipe_dispatch<f32x2>(q, pc, a, hop, tm, nullptr, &occ) : pipe_dispatch<double>(q, pc, a, hop, tm, nullptr, &occ)) != HGA_OK) occ = 1;
  if (const char* e = getenv("HGA_B200_PIPE_OCC")) occ = std::max(1, std::min(occ, atoi(e)));
  (void)C;
  return std::max<int64_t>(1, (int64_t)n_sm * std::max(1, occ));  // resident CTAs of the whole device
}

// ------------------------------------------------------------------------------------------------ the handle
struct hga_handle {
  hga_config cfg;
  Cascade q;
  PipeOperands po;
  DevBuf zir;
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
  DevBuf xdev, outdev, mudev, sigdev, rawdev;
  PinBuf pin_in, pin_out, pin_flag;   // pin_flag: per-warp completion slots of the zero-copy packet path
  unsigned long long push_seq = 0;
  bool push_inflight = false;         // a host push returned on its completion slots; its kernel may still be storing state
  std::vector<double> mu_cache, sig_cache;
  bool musig_valid = false;
  cudaStream_t s_own = nullptr, s_copy = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::vector<cudaEvent_t> slab_ev;
  int timing = 0;
  bool timed = false;
};

// Host pushes that poll completion slots return before their kernel has stored the new remainder / filter state on the
// handle's own stream; anything that touches that state from another stream (or frees it) waits for it first.
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
  if (cfg->precision != HGA_PREC_EXACT && cfg->precision != HGA_PREC_F64 && cfg->precision != HGA_PREC_F32) return fail(HGA_ERR_INVALID, "bad precision %d", cfg->precision);
  if (hga_device_count() < 1) return fail(HGA_ERR_NO_DEVICE, "no CUDA device visible: libhga_b200 has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= hga_device_count()) return fail(HGA_ERR_INVALID, "bad device %d", cfg->device);
  CU(cudaSetDevice(cfg->device));
  hga_handle* h = new hga_handle();
  h->cfg = *cfg;
  int rc = build_cascade(h->q, cfg->sos_bp, cfg->n_sections_bp, cfg->zi_bp, cfg->sos_notch, cfg->n_sections_notch, cfg->zi_notch);
  if (rc) { delete h; return rc; }
  build_pipe_const(h->q, cfg->frame_len, h->po.pc);
  build_pipe2_const(h->q, cfg->frame_len, kMaxSec / pipe_shape().S, h->po.pc2);
  if ((rc = upload_log_table())) { delete h; return rc; }
  cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, cfg->device);
  h->overlap = cfg->frame_len - cfg->hop;
  h->nb = (cfg->frame_len % cfg->hop == 0 && cfg->frame_len / cfg->hop <= kMaxBlocksPerFrame) ? cfg->frame_len / cfg->hop : 0;
  const double r = max_pole_radius(h->q);
  if (r >= 1.0) { delete h; return fail(HGA_ERR_INVALID, "unstable filter: pole radius %g", r); }
  if (cfg->warmup > 0) h->warmup = cfg->warmup;
  else {
    const double tol = cfg->warmup_tol > 0 ? cfg->warmup_tol : (cfg->precision == HGA_PREC_F32 ? 1e-7 : 1e-12);
    h->warmup = r > 0 ? (int)std::ceil(std::log(tol) / std::log(r)) : 1;
  }
  h->warmup = std::max(h->warmup, h->overlap);
  h->warmup = (h->warmup + cfg->hop - 1) / cfg->hop * cfg->hop;
  if (h->q.ns == kMaxSec) {  // k_hga_pipe2 candidates: zero-input response of the initial state, down to 1e-15 of its start
    int n = r > 0 ? (int)std::min(1e6, std::ceil(std::log(1e-15) / std::log(r))) : 1;
    n = (std::max(n, 1) + cfg->hop - 1) / cfg->hop * cfg->hop;
    const std::vector<double> zr = zero_input_response(h->q, n);
    if ((rc = h->zir.reserve((size_t)n * 8))) { delete h; return rc; }
    CU(cudaMemcpy(h->zir.p, zr.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
    h->po.zir = h->zir.as<double>();
    h->po.n_zir = n;
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
  if ((rc = tile_zi(h, h->s_own))) { hga_destroy(h); return rc; }
  *out = h;
  return HGA_OK;
}

extern "C" int hga_destroy(hga_handle* h) {
  if (!h) return HGA_OK;
  cudaSetDevice(h->cfg.device);
  drain_push(h);
  for (DevBuf* b : {&h->state, &h->framebuf, &h->state2, &h->framebuf2, &h->work, &h->segs, &h->ybuf, &h->stats_part, &h->seg_state, &h->seg_rem, &h->zir,
                    &h->xdev, &h->outdev, &h->mudev, &h->sigdev, &h->rawdev})
    b->release();
  h->pin_in.release();
  h->pin_out.release();
  h->pin_flag.release();
  for (cudaEvent_t e : h->slab_ev) cudaEventDestroy(e);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
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
extern "C" int hga_warmup(const hga_handle* h) { return h ? h->warmup : 0; }
extern "C" int hga_set_timing(hga_handle* h, int32_t enable) {
  if (!h) return fail(HGA_ERR_INVALID, "null handle");
  h->timing = enable;
  h->timed = false;
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

// Which fused kernel: 0 = thread-per-(item, channel) sequential cascade (f64 only), 1 = section-pipelined (f64 / f32).
static int choose_variant(const hga_handle* h, const double* x, int64_t ldx, bool plain) {
  const bool can_pipe = plain && pipe_supported(h->q, h->cfg.hop, h->nb) && pipe_layout_ok(x, ldx);
  if (h->cfg.precision == HGA_PREC_F32) return can_pipe ? 1 : -3;
  const char* env = getenv("HGA_B200_VARIANT");
  if (env && *env) return (atoi(env) == 1 && can_pipe) ? 1 : 0;
  return can_pipe ? 1 : 0;
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
  const bool fused = (h->cfg.precision == HGA_PREC_F64 || h->cfg.precision == HGA_PREC_F32) && h->nb > 0;
  // plan (cached on the segment table)
  std::vector<int64_t> key;
  key.reserve(3 * (size_t)n_seg + 1);
  key.insert(key.end(), seg_start, seg_start + n_seg);
  key.insert(key.end(), seg_len, seg_len + n_seg);
  key.insert(key.end(), out_row, out_row + n_seg);
  const bool plain = !state_in && !rem_in && !state_out && !rem_out;
  const int variant = fused ? choose_variant(h, x, ldx, plain) : -2;
  if (variant == -3) return fail(HGA_ERR_UNSUPPORTED, "f32 precision needs the standard 8+8-section unit-gain cascade, hop 10 or 20, "
                                                      "16-byte-aligned x with an even leading dimension, and no carried state");
  if (key != h->plan_key || variant != h->plan_variant) {
    const int cg = (C + 31) / 32;
    int64_t slots = 0;
    int64_t groups = cg;
    if (variant == 0) slots = std::max<int64_t>(1, (int64_t)h->n_sm * 12);
    else if (variant == 1) {
      slots = pipe_item_slots(h->q, h->po, h->cfg.precision, h->cfg.hop, h->n_sm, C);
      groups = h->cfg.precision == HGA_PREC_F32 ? (C + 63) / 64 : cg;
    }
    plan_chunks(h, n_seg, seg_start, seg_len, out_row, slots, groups, fused);
    if ((rc = upload_plan(h, st))) return rc;
    h->plan_key = key;
    h->plan_variant = variant;
  }
  if (fused) {
    int64_t xlo = INT64_MAX, xhi = 0;
    for (int i = 0; i < n_seg; ++i) { xlo = std::min(xlo, seg_start[i]); xhi = std::max(xhi, seg_start[i] + seg_len[i]); }
    FusedArgs a{x, ldx, h->work.as<Work>(), (int)h->work_host.size(), C, L, hop, h->nb, h->cfg.eps, mu, sigma, out, ldo, state_in, rem_in, state_out, rem_out, xlo, xhi, nullptr};
    if (h->timing) CU(cudaEventRecord(h->ev0, st));
    if (variant == 1) rc = launch_pipe(h->q, h->po, h->cfg.precision, a, st);
    else if (is_unit_std(h->q)) rc = launch_fused_seq<kUnitStd>(h, a, st);
    else if (h->q.unit) rc = launch_fused_seq<kUnit>(h, a, st);
    else rc = launch_fused_seq<kFused>(h, a, st);
    if (rc) return rc;
    if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; }
    return HGA_OK;
  }
  // exact: filtered rows to scratch (same row indexing as x, compacted to the span used), then row-order windows
  int64_t lo = INT64_MAX, hi = 0, maxF = 0;
  for (int i = 0; i < n_seg; ++i) {
    lo = std::min(lo, seg_start[i]);
    hi = std::max(hi, seg_start[i] + seg_len[i]);
    maxF = std::max(maxF, (seg_len[i] - L) / hop + 1);
  }
  if ((rc = h->ybuf.reserve((size_t)(hi - lo) * C * 8))) return rc;
  double* y = h->ybuf.as<double>() - lo * C;  // y row r <-> x row r
  const int cg = (C + 31) / 32;
  const unsigned grid = (unsigned)(((int64_t)n_seg * cg + 3) / 4);
  if (h->timing) CU(cudaEventRecord(h->ev0, st));
  k_iir_seq<kExact><<<grid, 128, 0, st>>>(h->q, x, ldx, y, C, h->segs.as<Segment>(), Segment{}, n_seg, C, nullptr, state_out);
  LAUNCHED();
  if (h->timing) { CU(cudaEventRecord(h->ev1, st)); h->timed = true; }
  if (n_seg > 65535) return fail(HGA_ERR_UNSUPPORTED, "more than 65535 segments per call");
  if (L == 5 * hop && frames_roll_on()) {
    const int R = roll_run_length(maxF * n_seg, cg, h->n_sm);
    dim3 g2((unsigned)cg, (unsigned)((maxF + (int64_t)R * kRollWarps - 1) / ((int64_t)R * kRollWarps)), (unsigned)n_seg);
    k_frames_roll<5><<<g2, 32 * kRollWarps, 0, st>>>(y, C, h->segs.as<Segment>(), Segment{}, n_seg, C, hop, R, h->cfg.eps, mu, sigma, out, ldo);
  } else {
    dim3 g2((unsigned)cg, (unsigned)((maxF + kFramesPerBlock - 1) / kFramesPerBlock), (unsigned)n_seg);
    k_frames_seq<<<g2, 256, 0, st>>>(y, C, 0, h->segs.as<Segment>(), Segment{}, n_seg, C, maxF, L, hop, nullptr, nullptr, h->cfg.eps, mu, sigma, out, ldo);
  }
  LAUNCHED();
  if (rem_out) {
    for (int i = 0; i < n_seg; ++i)
      CU(cudaMemcpy2DAsync(rem_out + (size_t)i * h->overlap * C, (size_t)C * 8, y + (seg_start[i] + seg_len[i] - h->overlap) * C,
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
    std::swap(h->state, h->state2);
    std::swap(h->framebuf, h->framebuf2);
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
  static const bool fast_on = [] { const char* e = getenv("HGA_B200_STREAM_FAST"); return !(e && atoi(e) == 0); }();
  if (fast_on && g.rows_total + n <= kStreamFastRows) {
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
  return stream_push_dev(h, x_dev, n, ldx, mu_dev, sigma_dev, out_dev, ldo, n_frames, (cudaStream_t)stream);
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
  if (stream_mapped() && n <= kMappedRows && in_map && out_map) {  // zero-copy: see kMappedRows
    unsigned long long* slots = stream_flagged() ? h->pin_flag.dev<unsigned long long>() : nullptr;
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
  if (stream_mapped() && n <= kMappedRows && in_map && out_map) {
    unsigned long long* slots = stream_flagged() ? h->pin_flag.dev<unsigned long long>() : nullptr;
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
  if (!start_dev && zero_rows == 0 && frame_len == 5 * hop && frames_roll_on()) {
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
  static const bool no_quad = [] { const char* e = getenv("HGA_B200_NO_ZSCORE4"); return e && *e && *e != '0'; }();
  if (!no_quad && f_dev != out_dev && (C & 3) == 0 && (ld & 3) == 0 && (ldo & 3) == 0 &&
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
  double* part = nullptr;
  CU(cudaMallocAsync((void**)&part, (size_t)nby * 3 * C * 8, st));
  k_stats_partial<<<dim3((unsigned)cg, (unsigned)nby), 256, 0, st>>>(feats_dev, F, C, ld, rpb, part);
  LAUNCHED();
  k_stats_merge<<<(C + 127) / 128, 128, 0, st>>>(acc_dev, part, (int)nby, C);
  LAUNCHED();
  CU(cudaFreeAsync(part, st));
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

// Identity map over compact rows: the flat k_rawscale stream (HGA_B200_NO_RAWSCALE=1 keeps the row-wise kernel, for A/B runs).
static bool rawscale_applies(const hga_chanmap* m, const void* x, int32_t dtype, const double* gain, int64_t ldx, const double* out, int64_t ldo) {
  static const bool off = [] { const char* e = getenv("HGA_B200_NO_RAWSCALE"); return e && *e && *e != '0'; }();
  if (off || !m->identity || ldx != m->c_in || ldo != m->c_out || (m->c_in & 3)) return false;
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

// ------------------------------------------------------------------------------------------------ offline, host buffers
// Upload x in time slabs on the copy stream; after each slab, launch the work items whose rows are resident; download
// each group's frame rows on the handle's own stream while the next slab uploads.
//
// Raw ingest (map != nullptr): the host rows are amplifier counts of `dtype`, c_in columns wide; each slab is uploaded at its
// raw element size and expanded on the copy stream by the channel-map kernel (signal * gain, then the folded
// pre-transform chain) into the float64 rows the extraction kernels read.
static size_t raw_elem_size(int32_t dtype) { return dtype == HGA_RAW_F64 ? 8 : dtype == HGA_RAW_I16 ? 2 : 4; }

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
  const double* x_host = static_cast<const double*>(x_host_v);   // the plain float64 path (map == nullptr)
  int64_t lo = INT64_MAX, hi = 0, olo = INT64_MAX, ohi = 0;
  for (int i = 0; i < n_segments; ++i) {
    if (seg_len[i] < L) return fail(HGA_ERR_INVALID, "segment %d has %lld rows < frame_len %d", i, (long long)seg_len[i], L);
    lo = std::min(lo, seg_start[i]);
    hi = std::max(hi, seg_start[i] + seg_len[i]);
    olo = std::min(olo, out_row[i]);
    ohi = std::max(ohi, out_row[i] + (seg_len[i] - L) / hop + 1);
  }
  const size_t row_bytes = (size_t)C * 8;
  if ((rc = h->xdev.reserve((size_t)(hi - lo) * row_bytes)) || (rc = h->outdev.reserve((size_t)(ohi - olo) * row_bytes))) return rc;
  cudaStream_t sc = h->s_copy, sk = h->s_own;
  const double *mu_d, *sig_d;
  if ((rc = sync_musigma(h, mu_host, sigma_host, sk, &mu_d, &sig_d))) return rc;
  double* xd = h->xdev.as<double>() - lo * C;     // device row r <-> host row r
  double* od = h->outdev.as<double>() - olo * C;  // device out row r <-> host out row r
  const bool fused = (h->cfg.precision == HGA_PREC_F64 || h->cfg.precision == HGA_PREC_F32) && h->nb > 0;
  const double* gain_d = nullptr;
  if (map) {
    if ((rc = h->rawdev.reserve((size_t)(hi - lo) * raw_row + (gain_host ? (size_t)c_in * 8 : 0) + 16))) return rc;
    if (gain_host) {
      const size_t off = ((size_t)(hi - lo) * raw_row + 15) / 16 * 16;
      gain_d = reinterpret_cast<const double*>(static_cast<unsigned char*>(h->rawdev.p) + off);
      CU(cudaMemcpyAsync(const_cast<double*>(gain_d), gain_host, (size_t)c_in * 8, cudaMemcpyHostToDevice, sc));
    }
  }
  // rows [r0, r1) of the host recording -> float64 device rows, enqueued on stream st
  auto upload = [&](int64_t r0, int64_t r1, cudaStream_t st) -> int {
    if (!map) {
      CU(cudaMemcpy2DAsync(xd + r0 * C, row_bytes, x_host + r0 * ldx, (size_t)ldx * 8, row_bytes, (size_t)(r1 - r0), cudaMemcpyHostToDevice, st));
      return HGA_OK;
    }
    unsigned char* rd = static_cast<unsigned char*>(h->rawdev.p) + (size_t)(r0 - lo) * raw_row;
    if (ldx == c_in)   // contiguous rows: one linear copy (a 2-D copy of 256 B rows runs at about half the link rate)
      CU(cudaMemcpyAsync(rd, x_raw + (size_t)r0 * raw_row, (size_t)(r1 - r0) * raw_row, cudaMemcpyHostToDevice, st));
    else
      CU(cudaMemcpy2DAsync(rd, raw_row, x_raw + (size_t)r0 * ldx * esz, (size_t)ldx * esz, raw_row, (size_t)(r1 - r0), cudaMemcpyHostToDevice, st));
    return hga_chanmap_apply_raw(map, rd, dtype, gain_d, r1 - r0, c_in, xd + r0 * C, C, st);
  };

  if (!fused) {
    if (map) CU(cudaStreamSynchronize(sc));   // the gain vector went up on the copy stream
    if ((rc = upload(lo, hi, sk))) return rc;
    if ((rc = extract_offline_dev(h, xd, C, n_segments, seg_start, seg_len, out_row, mu_d, sig_d, od, C, sk, nullptr, nullptr))) return rc;
    CU(cudaMemcpy2DAsync(out_host + olo * ldo, (size_t)ldo * 8, od + olo * C, row_bytes, row_bytes, (size_t)(ohi - olo), cudaMemcpyDeviceToHost, sk));
    CU(cudaStreamSynchronize(sk));
    return HGA_OK;
  }
  // plan once for the whole table (extract_offline_dev would do the same; here we launch sub-ranges of the work list)
  {
    std::vector<int64_t> key;
    key.insert(key.end(), seg_start, seg_start + n_segments);
    key.insert(key.end(), seg_len, seg_len + n_segments);
    key.insert(key.end(), out_row, out_row + n_segments);
    const int variant = choose_variant(h, xd + lo * C, C, true);
    if (variant == -3) return fail(HGA_ERR_UNSUPPORTED, "f32 precision needs the standard 8+8-section unit-gain cascade and hop 10 or 20");
    if (key != h->plan_key || variant != h->plan_variant) {
      const int cg = (C + 31) / 32;
      const int64_t slots = variant == 1 ? pipe_item_slots(h->q, h->po, h->cfg.precision, h->cfg.hop, h->n_sm, C) : std::max<int64_t>(1, (int64_t)h->n_sm * 12);
      const int64_t groups = (variant == 1 && h->cfg.precision == HGA_PREC_F32) ? (C + 63) / 64 : cg;
      plan_chunks(h, n_segments, seg_start, seg_len, out_row, slots, groups, true);
      if ((rc = upload_plan(h, sk))) return rc;
      h->plan_key = key;
      h->plan_variant = variant;
    }
  }
  const int variant = h->plan_variant;
  const std::vector<Work>& W = h->work_host;
  const int64_t total_bytes = (hi - lo) * (int64_t)row_bytes;
  // Each slab's launch runs its share of the items at the full per-item duration, so S slabs cost about S x (whole-job kernel
  // time) on the kernel stream; that hides behind the upload only while S <~ (upload time) / (kernel time) = about 3.8 x the
  // element size at 52 GB/s and 200 G channel-samples/s.  float64: 12 slabs; int16 counts: 5.
  const int64_t max_slabs = map ? std::max<int64_t>(2, std::min<int64_t>(12, (int64_t)(2.7 * (double)esz))) : 12;
  int64_t slab_bytes = std::min<int64_t>(std::max<int64_t>(total_bytes / max_slabs, 16LL << 20), (512LL << 20) * 8 / (int64_t)esz);
  const int64_t slab_rows = std::max<int64_t>(hop, slab_bytes / (int64_t)row_bytes / hop * hop);
  const size_t n_slabs = (size_t)((hi - lo + slab_rows - 1) / slab_rows);
  while (h->slab_ev.size() < 2 * n_slabs + 2) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->slab_ev.push_back(e);
  }
  size_t next = 0, ev = 0;
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
    FusedArgs a{xd, C, h->work.as<Work>() + next, (int)(end - next), C, L, hop, h->nb, h->cfg.eps, mu_d, sig_d, od, C, nullptr, nullptr, nullptr, nullptr, lo, hi, nullptr};
    if (variant == 1) rc = launch_pipe(h->q, h->po, h->cfg.precision, a, sk);
    else if (is_unit_std(h->q)) rc = launch_fused_seq<kUnitStd>(h, a, sk);
    else if (h->q.unit) rc = launch_fused_seq<kUnit>(h, a, sk);
    else rc = launch_fused_seq<kFused>(h, a, sk);
    if (rc) return rc;
    // frame rows this group produced
    int64_t f0 = INT64_MAX, f1 = 0;
    for (size_t i = next; i < end; ++i) {
      const Work& w = W[i];
      const int64_t a0 = w.out_row0 + w.t_acc / hop;                          // first frame: block t_acc/hop + nb-1 -> w = t_acc/hop
      const int64_t a1 = w.out_row0 + std::max<int64_t>(0, w.t_end / hop - (h->nb - 1));  // one past the last
      f0 = std::min(f0, a0);
      f1 = std::max(f1, a1);
    }
    if (f1 > f0)
      CU(cudaMemcpy2DAsync(out_host + f0 * ldo, (size_t)ldo * 8, od + f0 * C, row_bytes, row_bytes, (size_t)(f1 - f0), cudaMemcpyDeviceToHost, sk));
    next = end;
  }
  CU(cudaStreamSynchronize(sk));
  CU(cudaStreamSynchronize(sc));
  return HGA_OK;
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
