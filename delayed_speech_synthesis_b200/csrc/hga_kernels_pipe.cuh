// hga_kernels_pipe.cuh — K-hga-pipe: the section-pipelined form of the fused HGA kernel.
//
// Same function as k_hga_fused_seq (band-pass + notch cascade -> hop-block power sums -> log -> z-score for fresh
// extractors, /root/reference/local/units.py:151-160, extensions/hga/hga_optimized.pyx:27-47,104-107,
// prepare_corpus.py:226-227) but the 16 biquad sections of a channel are spread over the S warps of a CTA instead
// of living in one thread:
//
//   TMA (x tile, 32 channels x TB rows, f64) -> warp 0: sections [0, NS) -> smem link -> warp 1: sections [NS, 2NS)
//     -> ... -> warp S-1 -> smem y tile -> all warps: hop-block power sums, log, z-score -> HBM
//
// A lane owns one channel, so every shared-memory access is a conflict-free 256 B (f64) / 128 B (f32) row and the
// filter coefficients stay warp-uniform constant-bank operands.  Warp g works on tile n-g at step n (one bar.sync per
// step), so a (32-channel, time-chunk) work item occupies S warps rather than one: S times fewer time-chunks are
// needed to fill the machine, and the per-chunk warm-up (see hga_kernels_fused.cuh) shrinks accordingly.
//
// Gains: the design puts the filter gain in section 0 of each filter (b = g*[1, b1u, 1]).  The cascade here runs
// without them (3 FMA + 1 ADD per section) on the linearly scaled signal y/G, G = g_bp * g_notch; initial states
// are divided by the gain product in front of them and the window power is multiplied by G^2 before the log.
#pragma once
#include <cuda.h>

#include "hga_async.cuh"
#include "hga_kernels_fused.cuh"

namespace hga {

struct PipeConst {  // compile-time-indexable, warp-uniform operands (constant bank)
  double zu0[kMaxSec], zu1[kMaxSec];  // sosfilt_zi divided by the gains in front of each section
  double pow_scale;                   // G^2 / L
  unsigned long long na1p[kMaxSec], na2p[kMaxSec], b1p[kMaxSec];  // (-a1,-a1), (-a2,-a2), (b1u,b1u) as float32 pairs
  // k_hga_pipe2 only (build_pipe2_const): the same three operands as doubles, in the kernel's own section order and with
  // the signs its section forms use (hga_kernels_pipe2.cuh, "section forms"); na1p / na2p / b1p then hold their float32 pairs.
  double c1d[kMaxSec], c2d[kMaxSec], cbd[kMaxSec];
  unsigned amask;  // bit s set: section s has numerator [1, 0, -1] (3-slot form)
  int pad_;
};

// Per-arithmetic traits: channels per lane, coefficient operands (warp-uniform constant-bank reads; a1 / a2 are stored
// negated for the packed type, negated through the operand modifier otherwise), conversions at the tile boundaries.
template <typename T> struct PipeTraits;
template <> struct PipeTraits<double> {
  static constexpr int CPL = 1;
  static __device__ __forceinline__ double na1(const Cascade& q, const PipeConst&, int s) { return -q.a1[s]; }
  static __device__ __forceinline__ double na2(const Cascade& q, const PipeConst&, int s) { return -q.a2[s]; }
  static __device__ __forceinline__ double b1(const Cascade& q, const PipeConst&, int s) { return q.b1u[s]; }
  static __device__ __forceinline__ double zero() { return 0.0; }
  static __device__ __forceinline__ double splat(double v) { return v; }
  static __device__ __forceinline__ double load_x(const double* p) { return p[0]; }
  static __device__ __forceinline__ void split(double y, double (&o)[1]) { o[0] = y; }
  static __device__ __forceinline__ double load_part(const double* p, int) { return p[0]; }
};
template <> struct PipeTraits<f32x2> {
  static constexpr int CPL = 2;
  static __device__ __forceinline__ f32x2 na1(const Cascade&, const PipeConst& p, int s) { return f32x2{p.na1p[s]}; }
  static __device__ __forceinline__ f32x2 na2(const Cascade&, const PipeConst& p, int s) { return f32x2{p.na2p[s]}; }
  static __device__ __forceinline__ f32x2 b1(const Cascade&, const PipeConst& p, int s) { return f32x2{p.b1p[s]}; }
  static __device__ __forceinline__ f32x2 zero() { return f32x2{0ull}; }
  static __device__ __forceinline__ f32x2 splat(double v) { return pack2((float)v, (float)v); }
  static __device__ __forceinline__ f32x2 load_x(const double* p) { return pack2((float)p[0], (float)p[32]); }
  static __device__ __forceinline__ void split(f32x2 y, double (&o)[2]) {
    float a, b;
    unpack2(y, a, b);
    o[0] = (double)a;
    o[1] = (double)b;
  }
  // channel half `part` (0: lane, 1: lane + 32) of a packed pair in shared memory, widened
  static __device__ __forceinline__ double load_part(const f32x2* p, int part) { return (double)reinterpret_cast<const float*>(p)[part]; }
};
// tile readers: x tiles hold 32 * CPL doubles per row, links hold 32 T per row
template <typename T>
__device__ __forceinline__ T pipe_load(const double* in, int r) { return PipeTraits<T>::load_x(in + r * 32 * PipeTraits<T>::CPL); }
template <typename T>
__device__ __forceinline__ T pipe_load(const f32x2* in, int r) { return in[r * 32]; }

// One wavefront step: section s in [LO, HI] consumes the output its predecessor produced in the PREVIOUS step (section 0
// consumes x0), so the NS section updates of a step are mutually independent and the four instruction groups below
// (adds, numerator FMAs, z0 FMAs, z1 FMAs) keep every dependent pair >= NS issue slots apart: the FP64 pipe's 8-cycle
// latency at one issue per 2 cycles is covered by a single warp when NS >= 4.  Returns the output of section NS-1.
template <typename T, int G, int NS, int LO, int HI>
__device__ __forceinline__ T wave_step(const Cascade& q, const PipeConst& pc, T x0, T (&yp)[NS], T (&z0)[NS], T (&z1)[NS]) {
  using K = PipeTraits<T>;
  T xin[NS], yn[NS], t[NS];
#pragma unroll
  for (int s = LO; s <= HI; ++s) xin[s] = s == 0 ? x0 : yp[s - 1];
#pragma unroll
  for (int s = LO; s <= HI; ++s) yn[s] = xin[s] + z0[s];
#pragma unroll
  for (int s = LO; s <= HI; ++s) t[s] = fma(K::b1(q, pc, G * NS + s), xin[s], z1[s]);
#pragma unroll
  for (int s = LO; s <= HI; ++s) z0[s] = fma(K::na1(q, pc, G * NS + s), yn[s], t[s]);
#pragma unroll
  for (int s = LO; s <= HI; ++s) z1[s] = fma(K::na2(q, pc, G * NS + s), yn[s], xin[s]);
#pragma unroll
  for (int s = LO; s <= HI; ++s) yp[s] = yn[s];
  return yn[NS - 1];
}
template <typename T, int G, int NS, int K0, int K1>  // prologue steps K0..K1-1: sections 0..k active
struct WaveFill {
  static __device__ __forceinline__ void run(const Cascade& q, const PipeConst& pc, const T* v, T (&yp)[NS], T (&z0)[NS], T (&z1)[NS]) {
    wave_step<T, G, NS, 0, K0>(q, pc, v[K0], yp, z0, z1);
    WaveFill<T, G, NS, K0 + 1, K1>::run(q, pc, v, yp, z0, z1);
  }
};
template <typename T, int G, int NS, int K1>
struct WaveFill<T, G, NS, K1, K1> {
  static __device__ __forceinline__ void run(const Cascade&, const PipeConst&, const T*, T (&)[NS], T (&)[NS], T (&)[NS]) {}
};
template <typename T, int G, int NS, int J, int J1>  // epilogue steps: sections J..NS-1 active, one output each
struct WaveDrain {
  static __device__ __forceinline__ void run(const Cascade& q, const PipeConst& pc, T* o, T (&yp)[NS], T (&z0)[NS], T (&z1)[NS]) {
    o[(J - 1) * 32] = wave_step<T, G, NS, J, NS - 1>(q, pc, PipeTraits<T>::zero(), yp, z0, z1);
    WaveDrain<T, G, NS, J + 1, J1>::run(q, pc, o, yp, z0, z1);
  }
};
template <typename T, int G, int NS, int J1>
struct WaveDrain<T, G, NS, J1, J1> {
  static __device__ __forceinline__ void run(const Cascade&, const PipeConst&, T*, T (&)[NS], T (&)[NS], T (&)[NS]) {}
};

// `rows` rows of a tile through sections [G*NS, (G+1)*NS): in (row-major, 32 lanes per row) -> out.  Full tiles run
// the skewed wavefront (fill NS-1 steps, TB-NS+1 steady steps, drain NS-1 steps); ragged tiles go row by row.
template <typename T, typename TIN, int G, int NS, int TB>
__device__ __forceinline__ void pipe_rows(const Cascade& q, const PipeConst& pc, const TIN* __restrict__ in, T* __restrict__ out,
                                          int rows, T (&z0)[NS], T (&z1)[NS]) {
  using K = PipeTraits<T>;
  if (rows == TB) {
    T yp[NS];
    {
      T v[NS > 1 ? NS - 1 : 1];
#pragma unroll
      for (int i = 0; i < NS - 1; ++i) v[i] = pipe_load<T>(in, i);
      WaveFill<T, G, NS, 0, NS - 1>::run(q, pc, v, yp, z0, z1);
    }
    // steady state: all NS sections busy.  Small loop body (2 * U wavefront steps ~ 2.5 KB of SASS) so that the loop
    // stays in the sub-partition's L0 instruction cache (a 12-step body made the float32x2 variant fetch-bound:
    // ncu stall_no_instruction 1.1 per issue), with the next U input rows loaded while the current U are filtered.
    constexpr int U = 4;
    constexpr int NSTEADY = TB - NS + 1;
    constexpr int NPAIR = NSTEADY / (2 * U);
    int k = NS - 1;
    T va[U], vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) va[u] = pipe_load<T>(in, k + u);
#pragma unroll 1
    for (int it = 0; it < NPAIR; ++it, k += 2 * U) {
#pragma unroll
      for (int u = 0; u < U; ++u) vb[u] = pipe_load<T>(in, k + U + u);
#pragma unroll
      for (int u = 0; u < U; ++u) out[(k + u - (NS - 1)) * 32] = wave_step<T, G, NS, 0, NS - 1>(q, pc, va[u], yp, z0, z1);
      if (it + 1 < NPAIR || NSTEADY - NPAIR * 2 * U >= U) {
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = pipe_load<T>(in, k + 2 * U + u);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) out[(k + U + u - (NS - 1)) * 32] = wave_step<T, G, NS, 0, NS - 1>(q, pc, vb[u], yp, z0, z1);
    }
    if (NSTEADY - NPAIR * 2 * U >= U) {  // compile-time: one more batch of U, already in va
#pragma unroll
      for (int u = 0; u < U; ++u) out[(k + u - (NS - 1)) * 32] = wave_step<T, G, NS, 0, NS - 1>(q, pc, va[u], yp, z0, z1);
      k += U;
    }
#pragma unroll
    for (int r = 0; r < NSTEADY % U; ++r, ++k)
      out[(k - (NS - 1)) * 32] = wave_step<T, G, NS, 0, NS - 1>(q, pc, pipe_load<T>(in, k), yp, z0, z1);
    WaveDrain<T, G, NS, 1, NS>::run(q, pc, out + (TB - NS + 1) * 32, yp, z0, z1);
    return;
  }
  for (int r = 0; r < rows; ++r) {
    T x = pipe_load<T>(in, r);
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const T y = x + z0[s];
      z0[s] = fma(K::na1(q, pc, G * NS + s), y, fma(K::b1(q, pc, G * NS + s), x, z1[s]));
      z1[s] = fma(K::na2(q, pc, G * NS + s), y, x);
      x = y;
    }
    out[r * 32] = x;
  }
}

// Branch-free natural log for finite x > 0 (here x >= eps = 0.01): x = m * 2^e with m in [sqrt(1/2), sqrt(2)),
// log m = 2 atanh(s), s = (m - 1) / (m + 1), odd series to s^23 (|s| <= 0.1716), reciprocal by two Newton steps on
// rcp.approx, e * ln2 in two parts.  <= 2 ulp against glibc over [1e-2, 1e15] (tools/check_fast_log.py).  No special
// cases means several logs of one thread interleave, which the framing warps rely on.
__device__ __forceinline__ double fast_log(double x) {
  const long long bits = __double_as_longlong(x);
  int e = (int)((bits >> 52) & 0x7ff) - 1023;
  double m = __longlong_as_double((bits & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);  // [1, 2)
  const bool big = m > 1.4142135623730951;
  m = big ? m * 0.5 : m;
  e = big ? e + 1 : e;
  const double d = m + 1.0;
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  r = fma(r, fma(-d, r, 1.0), r);
  r = fma(r, fma(-d, r, 1.0), r);
  const double num = m - 1.0;
  double sv = num * r;
  sv = fma(fma(-d, sv, num), r, sv);  // one correction step: s = num / d to within an ulp
  const double z = sv * sv;
  double qv = 1.0 / 23.0;
  qv = fma(z, qv, 1.0 / 21.0);
  qv = fma(z, qv, 1.0 / 19.0);
  qv = fma(z, qv, 1.0 / 17.0);
  qv = fma(z, qv, 1.0 / 15.0);
  qv = fma(z, qv, 1.0 / 13.0);
  qv = fma(z, qv, 1.0 / 11.0);
  qv = fma(z, qv, 1.0 / 9.0);
  qv = fma(z, qv, 1.0 / 7.0);
  qv = fma(z, qv, 1.0 / 5.0);
  qv = fma(z, qv, 1.0 / 3.0);
  qv *= z;
  const double ed = (double)e, s2 = sv + sv;
  return fma(ed, 0x1.62e42fee00000p-1, s2) + fma(s2, qv, ed * 0x1.a39ef35793c76p-33);
}

constexpr int kFrameBatch = 4;  // tiles per log / z-score / store batch of the framing warps
constexpr int kPipeNB = 5;      // hop blocks per frame the pipe kernel is built for (L = 5 * hop at 1 and 2 kHz)

template <int S, int TB, typename T>
struct PipeSmem {
  static constexpr int CPL = PipeTraits<T>::CPL;
  static constexpr int NX = 2;        // TMA ring depth (tile n+1 is in flight while tile n is filtered)
  static constexpr int kBsRing = TB >= 40 ? 32 : 16;  // hop-block sums kept (>= (kFrameBatch + 1) * TB / hop + nb - 1)
  static constexpr size_t x_bytes = (size_t)NX * TB * 32 * CPL * 8;  // f64 input tiles
  static constexpr size_t link_bytes = (size_t)S * 2 * TB * 32 * sizeof(T);
  static constexpr size_t bs_bytes = (size_t)kBsRing * 32 * CPL * 8;
  static constexpr size_t total = x_bytes + link_bytes + bs_bytes + 64;
};

// One CTA = one (work item, group of 32 * CPL channels).  Warps [0, S) are the filter warps (one section group each);
// warps [S, 2S) are the framing warps: they issue the TMA loads, reduce the y tiles to hop-block power sums and turn
// closed frames into log / z-score features, so that their latency-bound chains (LDS -> FP64 add chain -> log) overlap
// with the pipe-bound filter warps instead of extending every step.  HOP = hop size (compile time: 10 or 20).
template <typename T, int S, int TB, int HOP>
__global__ void __launch_bounds__(64 * S, TB >= 40 ? 1 : 3) k_hga_pipe(const __grid_constant__ Cascade q, const __grid_constant__ PipeConst pc,
                                                      const __grid_constant__ FusedArgs a, const __grid_constant__ CUtensorMap tmap) {
  constexpr int NS = kMaxSec / S;
  using SM = PipeSmem<S, TB, T>;
  using TR = PipeTraits<T>;
  constexpr int CPL = TR::CPL, CW = 32 * CPL;  // channels per lane / per CTA
  constexpr int NX = SM::NX;
  constexpr int NBT = TB / HOP;  // hop blocks per tile
  constexpr uint32_t kTileBytes = TB * CW * 8;
  static_assert(TB % HOP == 0, "tile must hold whole hop blocks");
  static_assert((kFrameBatch + 1) * NBT + kPipeNB - 1 <= SM::kBsRing, "block-sum ring too small");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* base = smem_raw;
  double* xring = (double*)base;                                    // [NX][TB][CW]
  T* links = (T*)(base + SM::x_bytes);                              // [S][2][TB][32]; link S-1 is the y tile
  double* bs = (double*)(base + SM::x_bytes + SM::link_bytes);      // [kBsRing][CW]
  uint64_t* full = (uint64_t*)(base + SM::x_bytes + SM::link_bytes + SM::bs_bytes);  // [NX]

  const int cg = (a.C + CW - 1) / CW;
  const int item = blockIdx.x / cg, c0 = (blockIdx.x % cg) * CW;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const Work wk = a.work[item];
  const int64_t span = wk.t_end - wk.t_begin;
  const int ntiles = (int)((span + TB - 1) / TB);
  const int row_origin = (int)(wk.row0 + wk.t_begin - a.x_row_lo);  // TMA row coordinate of the item's first row

  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int i = 0; i < NX - 1 && i < ntiles; ++i) {
      mbar_expect_tx(&full[i], kTileBytes);
      tma_load_2d(xring + (size_t)i * TB * CW, &tmap, c0, row_origin + i * TB, &full[i]);
    }
  }
  __syncthreads();

  long long pr[6] = {0, 0, 0, 0, 0, 0};  // role, tma wait, phase A, phase B, barrier, total
  const bool prof = a.prof != nullptr;
  long long tk = prof ? clock64() : 0, tstart = tk;
#define HGA_TICK(slot)                  \
  if (prof) {                           \
    const long long now_ = clock64();   \
    pr[slot] += now_ - tk;              \
    tk = now_;                          \
  }

  if (g < S) {
    // ------------------------------------------------------------------ filter warps
    const bool fresh = wk.t_begin == 0;
    T z0[NS], z1[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) {  // g * NS + s is warp-uniform: a constant-bank lookup, once per kernel
      z0[s] = fresh ? TR::splat(pc.zu0[g * NS + s]) : TR::zero();
      z1[s] = fresh ? TR::splat(pc.zu1[g * NS + s]) : TR::zero();
    }
    for (int n = 0; n < ntiles + S + 1; ++n) {
      const int tile = n - g;
      if (tile >= 0 && tile < ntiles) {
        const int rows = (int)min((int64_t)TB, span - (int64_t)tile * TB);
        T* out = links + ((size_t)(g * 2 + (tile & 1)) * TB) * 32 + lane;
        if (g == 0) {
          mbar_wait(&full[tile % NX], (tile / NX) & 1);
          HGA_TICK(1)
          pipe_rows<T, double, 0, NS, TB>(q, pc, xring + (size_t)(tile % NX) * TB * CW + lane, out, rows, z0, z1);
        } else {
          const T* in = links + ((size_t)((g - 1) * 2 + (tile & 1)) * TB) * 32 + lane;
          if (S >= 2 && g == 1) pipe_rows<T, T, (S >= 2 ? 1 : 0), NS, TB>(q, pc, in, out, rows, z0, z1);
          if (S >= 4 && g == 2) pipe_rows<T, T, (S >= 4 ? 2 : 0), NS, TB>(q, pc, in, out, rows, z0, z1);
          if (S >= 4 && g == 3) pipe_rows<T, T, (S >= 4 ? 3 : 0), NS, TB>(q, pc, in, out, rows, z0, z1);
        }
      }
      HGA_TICK(0)
      asm volatile("bar.sync 0;" ::: "memory");
      HGA_TICK(4)
    }
  } else {
    // ------------------------------------------------------------------ framing warps
    const int w = g - S;
    const bool zs = a.mu != nullptr;
    double mu[CPL], inv_sigma[CPL];
#pragma unroll
    for (int p = 0; p < CPL; ++p) {
      const int c = c0 + lane + 32 * p;
      mu[p] = (zs && c < a.C) ? a.mu[c] : 0.0;
      inv_sigma[p] = (zs && c < a.C) ? 1.0 / a.sigma[c] : 1.0;
    }
    const int64_t blk_begin = wk.t_begin / HOP;
    double* op = a.out + wk.out_row0 * a.ldo + c0 + lane;
    for (int n = 0; n < ntiles + S + 1; ++n) {
      // TMA producer: the slot of tile n + NX - 1 was last read by filter warp 0 in step n - 1
      if (w == 0 && lane == 0) {
        const int pre = n + NX - 1;
        if (pre < ntiles) {
          uint64_t* b = &full[pre % NX];
          mbar_expect_tx(b, kTileBytes);
          tma_load_2d(xring + (size_t)(pre % NX) * TB * CW, &tmap, c0, row_origin + pre * TB, b);
        }
      }
      HGA_TICK(1)
      // phase A: hop-block power sums of y tile mA (written by filter warp S-1 in step n-1)
      const int mA = n - S, mB = n - S - 1;
      if (mA >= 0 && mA < ntiles) {
        const T* yt = links + ((size_t)((S - 1) * 2 + (mA & 1)) * TB) * 32 + lane;
        const int64_t t0 = wk.t_begin + (int64_t)mA * TB;
#pragma unroll
        for (int j0 = 0; j0 < NBT; j0 += S) {
          const int j = j0 + w;
          const int64_t tb = t0 + (int64_t)j * HOP;
          if (j < NBT && tb >= wk.t_acc && tb + HOP <= wk.t_end) {
            T v[HOP];
#pragma unroll
            for (int i = 0; i < HOP; ++i) v[i] = yt[(j * HOP + i) * 32];
            double acc0[CPL], acc1[CPL];
#pragma unroll
            for (int p = 0; p < CPL; ++p) acc0[p] = acc1[p] = 0.0;
#pragma unroll
            for (int i = 0; i < HOP; i += 2) {
              double u0[CPL], u1[CPL];
              TR::split(v[i], u0);
              TR::split(v[i + 1], u1);
#pragma unroll
              for (int p = 0; p < CPL; ++p) {
                acc0[p] = fma(u0[p], u0[p], acc0[p]);
                acc1[p] = fma(u1[p], u1[p], acc1[p]);
              }
            }
#pragma unroll
            for (int p = 0; p < CPL; ++p)
              bs[((blk_begin + (int64_t)mA * NBT + j) & (SM::kBsRing - 1)) * CW + lane + 32 * p] = acc0[p] + acc1[p];
          }
        }
      }
      HGA_TICK(2)
      // phase B, every kFrameBatch tiles: log / z-score / store of the frames closed by the blocks of tiles [mB0, mB]
      // (their sums were written in step n-1 or earlier); PER * CPL independent logs in flight per thread
      if (mB >= 0 && mB < ntiles && (mB % kFrameBatch == kFrameBatch - 1 || mB == ntiles - 1)) {
        const int mB0 = mB - mB % kFrameBatch;
        const int nblk = (mB - mB0 + 1) * NBT;
        const int64_t blk0 = blk_begin + (int64_t)mB0 * NBT;
        constexpr int PER = (kFrameBatch * NBT + S - 1) / S;
        double sum[PER][CPL];
        bool ok[PER];
#pragma unroll
        for (int u = 0; u < PER; ++u) {
          const int jb = u * S + w;
          const int64_t blk = blk0 + jb, tb = blk * HOP;
          ok[u] = jb < nblk && tb >= wk.t_acc + (int64_t)(kPipeNB - 1) * HOP && tb + HOP <= wk.t_end;
#pragma unroll
          for (int p = 0; p < CPL; ++p) {
            double v[kPipeNB];
#pragma unroll
            for (int k = 0; k < kPipeNB; ++k) v[k] = bs[((blk - (kPipeNB - 1) + k) & (SM::kBsRing - 1)) * CW + lane + 32 * p];
            double sacc = v[0];
#pragma unroll
            for (int k = 1; k < kPipeNB; ++k) sacc += v[k];  // oldest block first, like the sequential kernel
            sum[u][p] = ok[u] ? sacc : 1.0;
          }
        }
#pragma unroll
        for (int u = 0; u < PER; ++u) {
#pragma unroll
          for (int p = 0; p < CPL; ++p) {
            double f = fast_log(fma(sum[u][p], pc.pow_scale, a.eps));
            if (zs) f = (f - mu[p]) * inv_sigma[p];
            if (ok[u] && c0 + lane + 32 * p < a.C) op[(blk0 + u * S + w - (kPipeNB - 1)) * a.ldo + 32 * p] = f;
          }
        }
      }
      HGA_TICK(3)
      asm volatile("bar.sync 0;" ::: "memory");
      HGA_TICK(4)
    }
  }
  if (prof && lane == 0) {
    pr[5] = clock64() - tstart;
    for (int i = 0; i < 6; ++i) a.prof[((size_t)blockIdx.x * 8 + g) * 6 + i] = pr[i];
  }
#undef HGA_TICK
}

}  // namespace hga
