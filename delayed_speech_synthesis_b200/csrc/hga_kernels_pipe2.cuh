// hga_kernels_pipe2.cuh — K-hga-pipe2: section-pipelined fused HGA kernel, second form.
//
// Same function as k_hga_pipe (band-pass + notch cascade -> hop-block power sums -> log -> z-score for fresh extractors,
// /root/reference/local/units.py:151-160, extensions/hga/hga_optimized.pyx:27-47,104-107, prepare_corpus.py:226-227)
// and the same arithmetic per section (wave_step), with three structural changes:
//
//  * the skewed wavefront runs as ONE CONTINUOUS STREAM per warp: section s of a warp always works on the row its
//    predecessor finished one step earlier, across tile boundaries, so a tile costs exactly TB wavefront steps (no
//    per-tile fill / drain).  Warp g therefore emits its stream NS-1 rows late; tiles of link g hold the stream shifted
//    by d(g+1) = E + (g+1)(NS-1) rows, and the y tiles by DP = E + S(NS-1) rows, where E pads DP to a multiple of the
//    hop so that hop blocks stay aligned with tiles.  Only the first tile of a FRESH item (initial state = sosfilt_zi)
//    needs the section-by-section start; items that start from zero state simply see E more rows of warm-up.
//  * warps synchronise pairwise through mbarriers (full / empty per link buffer) instead of one CTA-wide bar.sync per
//    step: a warp waits only for the tile it consumes and the buffer it overwrites, so jitter in one stage (TMA
//    latency, the framing warps' log batches) is absorbed by the buffers instead of stalling all stages.
//  * the float32 mode's float64 -> float32 conversions of the x tiles (F2F is a slow unit on this part) are done by the
//    framing warps, one tile per warp in rotation, so that they are spread over the four SM sub-partitions and stay off
//    the filter warps; the hop-block power sums of that mode stay in float32 (only block sums are widened).
//  * the log is table-driven (tab_log: ~10 FP64 instructions instead of ~45).
//
// Measured shapes (profiles/r01_pipe2_sweep.txt): S = 4 filter warps x NS = 4 sections with the section group as a
// compile-time constant (coefficients are constant-bank operands; warp g and its code copy own SM sub-partition g) is
// the fastest.  S = 8 x NS = 2 halves the time-chunks but needs ONE copy of the code for all filter warps (two warps of
// a sub-partition running different copies of the loop thrash its instruction buffer: 38.6 -> 56.6 cycles per row,
// tools/microbench/cascade_rates), hence register coefficients, and FFMA2 / DFMA with three register operands run
// 6 - 20 % slower; it is kept compiled for A/B runs (HGA_B200_PIPE_S=8).
#pragma once
#include "hga_kernels_pipe.cuh"

namespace hga {

template <int S, int TB, typename T>
struct Pipe2Smem {
  static constexpr int CPL = PipeTraits<T>::CPL;
  // Ring depths shrink as the tiles grow (227 KB of shared memory): a tile boundary costs ~300 cycles per filter warp
  // (two mbarrier waits, two arrivals, the load prologue, hand-off jitter), so long tiles with shallow rings beat short
  // tiles with deep rings (K2 float64: 40 rows 183 G, 60 rows 196 G, 80 rows 201 G channel-samples/s).
  static constexpr int NX = PipeTraits<T>::CPL == 1 ? (TB >= 80 ? 2 : TB >= 70 ? 3 : 4) : (TB > 40 ? 2 : 3);  // TMA ring depth
#ifndef HGA_PIPE2_NB64
#define HGA_PIPE2_NB64 3
#endif
  static constexpr int NB = (PipeTraits<T>::CPL == 1 && TB <= 40) ? HGA_PIPE2_NB64 : 2;  // buffers per internal link
  static constexpr int NBY = TB <= 40 ? 4 : (TB >= 80 || PipeTraits<T>::CPL == 2) ? 2 : 3;   // y tiles (filter warp S-1 -> framing warps)
  static constexpr int kBsRing = TB / 10 >= 4 ? 32 : 16;  // hop-block sums kept (framing warps may be two tiles apart)
  static constexpr int NBS = 4;   // block-sum hand-off barriers (phase A of tile m -> phase B of tile m)
  static constexpr int NBC = NX;  // float32 mode: converted x tiles (framing warps -> filter warp 0)
  static constexpr size_t x_bytes = (size_t)NX * TB * 32 * CPL * 8;
  static constexpr size_t c_bytes = CPL == 2 ? (size_t)NBC * TB * 32 * sizeof(T) : 0;
  static constexpr int NL = S - 1;  // links between filter warps
  static constexpr size_t link_bytes = (size_t)NL * NB * TB * 32 * sizeof(T);
  static constexpr size_t y_bytes = (size_t)NBY * TB * 32 * sizeof(T);
  static constexpr size_t bs_bytes = (size_t)kBsRing * 32 * CPL * (CPL == 2 ? 4 : 8);  // float32 mode keeps its block sums in float32
  static constexpr size_t tab_bytes = 256 * 8;  // log table (kLogTabN x {1/c, -log(1/c)})
  static constexpr size_t c_off = x_bytes, link_off = c_off + c_bytes, y_off = link_off + link_bytes, bs_off = y_off + y_bytes,
                          tab_off = bs_off + bs_bytes, bar_off = tab_off + tab_bytes;
  // mbarriers: full_x[NX] empty_x[NX] full_c[NBC] empty_c[NBC] full_l[NL][NB] empty_l[NL][NB] full_y[NBY] empty_y[NBY] bs_full[NBS]
  static constexpr int b_full_x = 0, b_empty_x = NX, b_full_c = 2 * NX, b_empty_c = b_full_c + NBC, b_full_l = b_empty_c + NBC,
                       b_empty_l = b_full_l + NL * NB, b_full_y = b_empty_l + NL * NB, b_empty_y = b_full_y + NBY, b_bs = b_empty_y + NBY, n_bars = b_bs + NBS;
  static constexpr size_t total = bar_off + (size_t)n_bars * 8;
};

// Section forms.  Every section runs on the gain-free signal (see hga_kernels_pipe.cuh); its three operands c1, c2, cb
// come from PipeConst (c1d / c2d / cbd as doubles, na1p / na2p / b1p as float32 pairs), filled by build_pipe2_const:
//   U  numerator [1, b1, 1], transposed direct form II:   y = x + z0;  z0' = (-a1) y + (b1 x + z1);  z1' = (-a2) y + x
//        operands c1 = -a1, c2 = -a2, cb = b1: 1 ADD + 3 FMA
//   A  numerator [1, 0, -1] (a zero at DC and one at Nyquist), same form on the NEGATED state v = -z0, w = -z1:
//        y = x - v;  v' = a1 y + w;  w' = a2 y + x        operands c1 = +a1, c2 = +a2: 1 ADD + 2 FMA
// The reference's band-pass has numerator (1 + z^-1)^8 (1 - z^-1)^8 = (1 - z^-2)^8 (four sections [1, 2, 1], four
// [1, -2, 1]); paired as eight A sections it is the same transfer function with 8 FP slots fewer
// per channel-sample (56 instead of 64).  Re-pairing changes what the section states mean, so the kernel never loads the
// reference's initial state: every item starts from ZERO state and, for items that begin a segment, the zero-input
// response of sosfilt_zi (one sequence for all channels, FusedArgs::zir) is added to the filtered rows (linearity).
// AM = bit mask of the A sections among the NS sections of a filter warp (the same for every warp, so that all filter
// warps do the same amount of work: the host interleaves the sections [A, A, U, U] per warp).
template <typename T> struct P2Op;
template <> struct P2Op<double> {
  static __device__ __forceinline__ double c1(const PipeConst& p, int s) { return p.c1d[s]; }
  static __device__ __forceinline__ double c2(const PipeConst& p, int s) { return p.c2d[s]; }
  static __device__ __forceinline__ double cb(const PipeConst& p, int s) { return p.cbd[s]; }
};
template <> struct P2Op<f32x2> {
  static __device__ __forceinline__ f32x2 c1(const PipeConst& p, int s) { return f32x2{p.na1p[s]}; }
  static __device__ __forceinline__ f32x2 c2(const PipeConst& p, int s) { return f32x2{p.na2p[s]}; }
  static __device__ __forceinline__ f32x2 cb(const PipeConst& p, int s) { return f32x2{p.b1p[s]}; }
};
// Coefficient source of a filter warp.  G >= 0: the section group is a compile-time constant, coefficients are
// constant-bank / uniform-register operands (fastest arithmetic: FFMA2 with three register operands runs ~20 % slower),
// but every group is its own copy of the code, so it is only used when each SM sub-partition runs ONE group (S == 4:
// warp g sits on sub-partition g; two warps of a sub-partition running different copies of the loop thrash its
// instruction buffer, tools/microbench/cascade_rates).  G == -1: run-time group, coefficients in registers, one copy.
template <typename T, int NS, int G>
struct Coef {
  const PipeConst& pc;
  __device__ __forceinline__ Coef(const Cascade&, const PipeConst& pc_, int) : pc(pc_) {}
  __device__ __forceinline__ T na1(int s) const { return P2Op<T>::c1(pc, G * NS + s); }
  __device__ __forceinline__ T na2(int s) const { return P2Op<T>::c2(pc, G * NS + s); }
  __device__ __forceinline__ T b1(int s) const { return P2Op<T>::cb(pc, G * NS + s); }
};
template <typename T, int NS>
struct Coef<T, NS, -1> {
  T c1[NS], c2[NS], cb[NS];
  __device__ __forceinline__ Coef(const Cascade&, const PipeConst& pc, int g) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      c1[s] = P2Op<T>::c1(pc, g * NS + s);
      c2[s] = P2Op<T>::c2(pc, g * NS + s);
      cb[s] = P2Op<T>::cb(pc, g * NS + s);
    }
  }
  __device__ __forceinline__ T na1(int s) const { return c1[s]; }
  __device__ __forceinline__ T na2(int s) const { return c2[s]; }
  __device__ __forceinline__ T b1(int s) const { return cb[s]; }
};

// One wavefront step (see wave_step in hga_kernels_pipe.cuh): section s consumes the output its predecessor produced in
// the PREVIOUS step, so the NS section updates of a step are mutually independent.
template <typename T, int NS, unsigned AM, typename CF>
__device__ __forceinline__ T wave_step_r(const CF& c, T x0, T (&yp)[NS], T (&z0)[NS], T (&z1)[NS]) {
  T xin[NS], yn[NS], t[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) xin[s] = s == 0 ? x0 : yp[s - 1];
#pragma unroll
  for (int s = 0; s < NS; ++s) yn[s] = ((AM >> s) & 1u) ? xin[s] - z0[s] : xin[s] + z0[s];
#pragma unroll
  for (int s = 0; s < NS; ++s) t[s] = ((AM >> s) & 1u) ? z1[s] : fma(c.b1(s), xin[s], z1[s]);
#pragma unroll
  for (int s = 0; s < NS; ++s) z0[s] = fma(c.na1(s), yn[s], t[s]);
#pragma unroll
  for (int s = 0; s < NS; ++s) z1[s] = fma(c.na2(s), yn[s], xin[s]);
#pragma unroll
  for (int s = 0; s < NS; ++s) yp[s] = yn[s];
  return yp[NS - 1];
}

// One full tile through the sections of a warp: in-tile row i -> out-tile row i (the out tile is the stream shifted by
// NS-1 more rows than the in tile).  Loop body = 2U rows, next U input rows loaded while the current U are filtered.
template <typename T, int NS, int TB, unsigned AM = 0u, typename CF>
__device__ __forceinline__ void pipe2_tile(const CF& c, const T* __restrict__ in, T* __restrict__ out,
                                           T (&yp)[NS], T (&z0)[NS], T (&z1)[NS]) {
  constexpr int U = (TB % 10 == 0) ? 5 : 4;
  static_assert(TB % (2 * U) == 0, "tile rows must be a multiple of the loop body");
  T va[U], vb[U];
#pragma unroll
  for (int u = 0; u < U; ++u) va[u] = in[u * 32];
#pragma unroll 1
  for (int k = 0; k < TB; k += 2 * U) {
#pragma unroll
    for (int u = 0; u < U; ++u) vb[u] = in[(k + U + u) * 32];
#pragma unroll
    for (int u = 0; u < U; ++u) out[(k + u) * 32] = wave_step_r<T, NS, AM>(c, va[u], yp, z0, z1);
    if (k + 2 * U < TB) {
#pragma unroll
      for (int u = 0; u < U; ++u) va[u] = in[(k + 2 * U + u) * 32];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) out[(k + U + u) * 32] = wave_step_r<T, NS, AM>(c, vb[u], yp, z0, z1);
  }
}

// Filter warp g of a CTA (g is a run-time value: one copy of this code serves all filter warps): sections
// [g*NS, (g+1)*NS) over every tile of the work item.
template <typename T, int S, int TB, int E, int G, unsigned AM>
__device__ __forceinline__ void pipe2_filter_warp(const Cascade& q, const PipeConst& pc, const CUtensorMap* tmap, unsigned char* base, int ntiles,
                                                  bool fresh, int g_rt, int c0, int row_first, int lane, long long* pr, bool prof) {
  const int g = G >= 0 ? G : g_rt;
  constexpr int NS = kMaxSec / S;
  using SM = Pipe2Smem<S, TB, T>;
  using TR = PipeTraits<T>;
  constexpr int CPL = TR::CPL, NX = SM::NX, NB = SM::NB, NBY = SM::NBY, NBC = SM::NBC;
  constexpr bool CONV = CPL == 2;  // float32 mode: the framing warps convert the x tiles, warp 0 reads the converted ring
  T* xring = (T*)base;             // read directly only in float64 mode (T == double, 32 columns per row like the links)
  T* cring = (T*)(base + SM::c_off);
  T* links = (T*)(base + SM::link_off);
  T* ylink = (T*)(base + SM::y_off);
  uint64_t* bars = (uint64_t*)(base + SM::bar_off);
  const bool first = g == 0, last = g == S - 1;

  // ring bases, barriers and depths of this warp's input and output
  const T* in_base = (first ? (CONV ? cring : xring) : links + (size_t)(g - 1) * NB * TB * 32) + lane;
  T* out_base = (last ? ylink : links + (size_t)g * NB * TB * 32) + lane;
  uint64_t* in_full0 = bars + (first ? (CONV ? SM::b_full_c : SM::b_full_x) : SM::b_full_l + (g - 1) * NB);
  uint64_t* in_empty0 = bars + (first ? (CONV ? SM::b_empty_c : SM::b_empty_x) : SM::b_empty_l + (g - 1) * NB);
  uint64_t* out_full0 = bars + (last ? SM::b_full_y : SM::b_full_l + g * NB);
  uint64_t* out_empty0 = bars + (last ? SM::b_empty_y : SM::b_empty_l + g * NB);
  const int nin = first ? (CONV ? NBC : NX) : NB, nout = last ? NBY : NB;

  const Coef<T, NS, G> c(q, pc, g);
  T z0[NS], z1[NS], yp[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    z0[s] = TR::zero();  // every item starts from zero state (fresh items: + FusedArgs::zir in the framing warps)
    z1[s] = TR::zero();
    yp[s] = TR::zero();
  }
  long long tk = prof ? clock64() : 0;
#define HGA_TICK2(slot)               \
  if (prof) {                         \
    const long long now_ = clock64(); \
    pr[slot] += now_ - tk;            \
    tk = now_;                        \
  }
  int si = 0, pi = 0;  // input slot / phase
  int so = 0, po = 0;  // output slot / phase
  for (int m = 0; m < ntiles; ++m) {
    mbar_wait(in_full0 + si, pi);
    HGA_TICK2(1)
    mbar_wait(out_empty0 + so, po ^ 1);
    HGA_TICK2(4)
    const T* in = in_base + (size_t)si * TB * 32;
    T* out = out_base + (size_t)so * TB * 32;
    if (m == 0 && fresh && first) {
      // the E rows in front of a segment are not part of it (they may hold the previous segment): the stream starts at 0
      T* inw = const_cast<T*>(in);
#pragma unroll
      for (int i = 0; i < E; ++i) inw[i * 32] = TR::zero();  // both modes: this warp's input rows hold 32 T
      __syncwarp();
    }
    pipe2_tile<T, NS, TB, AM>(c, in, out, yp, z0, z1);
    __syncwarp();
    if (lane == 0) {
      mbar_arrive(out_full0 + so);
      mbar_arrive(in_empty0 + si);  // float64 mode, warp 0: releases the TMA slot to the producer warp
    }
    if (++si == nin) { si = 0; pi ^= 1; }
    if (++so == nout) { so = 0; po ^= 1; }
    HGA_TICK2(0)
  }
#undef HGA_TICK2
}

// TB rows of 64 float64 -> packed float32 pairs, in batches of 10 rows: all loads of a batch are issued before the first
// conversion and all conversions before the first store (the row-by-row form serialises on load -> F2F -> store latency
// because the compiler must assume the tiles alias: 66 cycles per row, tools/microbench/cascade_rates).
template <int TB>
__device__ __forceinline__ void convert_rows(const double* __restrict__ in, f32x2* __restrict__ out) {
  constexpr int R = 10;
  static_assert(TB % R == 0, "tile rows must be a multiple of the conversion batch");
#pragma unroll 1
  for (int r0 = 0; r0 < TB; r0 += R) {
    double a[R], b[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      a[r] = in[(r0 + r) * 64];
      b[r] = in[(r0 + r) * 64 + 32];
    }
    f32x2 v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) v[r] = pack2((float)a[r], (float)b[r]);
#pragma unroll
    for (int r = 0; r < R; ++r) out[(r0 + r) * 32] = v[r];
  }
}

// float32 mode: one x tile (float64, 64 channels per row) -> packed float32 pairs (channels c, c + 32 per lane) in the
// converted ring.  Run by the framing warps (tile j by warp j % NF) so that the slow F2F conversions are spread over
// the four sub-partitions and stay off the filter warps; the converting warp also refills the TMA slot it has read.
template <typename T, int S, int TB>
__device__ __forceinline__ void pipe2_convert_tile(const CUtensorMap* tmap, unsigned char* base, int j, int ntiles, int c0, int row_first, int lane) {
  using SM = Pipe2Smem<S, TB, T>;
  using TR = PipeTraits<T>;
  constexpr int CPL = TR::CPL, CW = 32 * CPL, NX = SM::NX, NBC = SM::NBC;
  constexpr uint32_t kTileBytes = TB * CW * 8;
  double* xring = (double*)base;
  T* cring = (T*)(base + SM::c_off);
  uint64_t* bars = (uint64_t*)(base + SM::bar_off);
  const int sx = j % NX, sc = j % NBC;
  // Order matters: once the converted slot of tile j - NBC has been released, tile j - NX (same x slot, NBC == NX) has
  // been converted and the TMA of tile j issued, so the parity wait on full_x cannot alias an older phase.
  static_assert(NBC == NX, "converted ring and TMA ring must have the same depth");
  mbar_wait(bars + SM::b_empty_c + sc, ((j / NBC) & 1) ^ 1);
  mbar_wait(bars + SM::b_full_x + sx, (j / NX) & 1);
  const double* in = xring + (size_t)sx * TB * CW + lane;
  T* out = cring + (size_t)sc * TB * 32 + lane;
  convert_rows<TB>(in, reinterpret_cast<f32x2*>(out));
  __syncwarp();
  if (lane == 0) {
    mbar_arrive(bars + SM::b_full_c + sc);
    if (j + NX < ntiles) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_expect_tx(bars + SM::b_full_x + sx, kTileBytes);
      tma_load_2d(xring + (size_t)sx * TB * CW, tmap, c0, row_first + (j + NX) * TB, bars + SM::b_full_x + sx);
    }
  }
}

// One CTA = one (work item, group of 32 * CPL channels): warps [0, S) filter, warps [S, S + NF) frame.
template <typename T, int S, int TB, int HOP, int NF, bool PERG, unsigned AM>
__global__ void __launch_bounds__(32 * (S + NF + 1), 1) k_hga_pipe2(const __grid_constant__ Cascade q, const __grid_constant__ PipeConst pc,
                                                                 const __grid_constant__ FusedArgs a, const __grid_constant__ CUtensorMap tmap) {
  constexpr int NS = kMaxSec / S;
  using SM = Pipe2Smem<S, TB, T>;
  using TR = PipeTraits<T>;
  constexpr int CPL = TR::CPL, CW = 32 * CPL, NX = SM::NX, NB = SM::NB, NBY = SM::NBY;
  constexpr int NBT = TB / HOP;
  constexpr int LAG = S * (NS - 1);
  constexpr int DP = (LAG + HOP - 1) / HOP * HOP;  // total lag of the y tiles, a whole number of hop blocks
  constexpr int E = DP - LAG;                      // rows the x tiles start early
  constexpr uint32_t kTileBytes = TB * CW * 8;
  static_assert(TB % HOP == 0 && DP <= TB, "tile must hold whole hop blocks and the whole lag");
  static_assert(kPipeNB - 1 + 2 * NBT <= SM::kBsRing, "block-sum ring too small");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* base = smem_raw;
  double* xring = (double*)base;
  T* ylink = (T*)(base + SM::y_off);
  double* bs = (double*)(base + SM::bs_off);
  double* logtab = (double*)(base + SM::tab_off);
  uint64_t* bars = (uint64_t*)(base + SM::bar_off);
  uint64_t* full_x = bars + SM::b_full_x;
  uint64_t* full_y = bars + SM::b_full_y;
  uint64_t* empty_y = bars + SM::b_empty_y;
  // float32 mode: the framing warps convert x tile m + LEAD before they frame y tile m.  The converter of tile j waits on
  // the (j / NBC - 1)-th completion of empty_c[j % NBC] by PARITY, which is only sound if completion j / NBC - 2 (tile
  // j - 2 NBC consumed) is certain to have happened: at iteration m only tiles <= m - 1 are known to be through the
  // filter warps, so LEAD <= 2 NBC - 1 (a 2-deep converted ring with LEAD = 5 let a converter run one phase ahead of
  // its barrier at start-up, overwrite an unconsumed tile and hang the CTA).
  constexpr int LEAD = (S + 1) < (2 * SM::NBC - 1) ? (S + 1) : (2 * SM::NBC - 1);

  const int cg = (a.C + CW - 1) / CW;
  const int item = blockIdx.x / cg, c0 = (blockIdx.x % cg) * CW;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const Work wk = a.work[item];
  const int64_t span = wk.t_end - wk.t_begin;
  const int ntiles = (int)((span + DP + TB - 1) / TB);
  const int row_first = (int)(wk.row0 + wk.t_begin - a.x_row_lo) - E;  // TMA row of tile 0, position 0 (may be < 0)

  for (int i = threadIdx.x; i < 2 * kLogTabN; i += blockDim.x) logtab[i] = g_logtab[i];
  if (threadIdx.x == 0) {
    for (int i = 0; i < SM::n_bars; ++i) mbar_init(&bars[i], i >= SM::b_empty_y ? NF : 1);  // empty_y and bs_full collect one arrival per framing warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int i = 0; i < NX && i < ntiles; ++i) {
      mbar_expect_tx(&full_x[i], kTileBytes);
      tma_load_2d(xring + (size_t)i * TB * CW, &tmap, c0, row_first + i * TB, &full_x[i]);
    }
  }
  __syncthreads();

  long long pr[6] = {0, 0, 0, 0, 0, 0};  // work, wait in, phase A, phase B, wait out / barrier, total
  const bool prof = a.prof != nullptr;
  const long long tstart = prof ? clock64() : 0;
  const bool fresh = wk.t_begin == 0;

  if (g < S) {
    if (PERG) {
#define HGA_FW(G_) \
  if (g == G_) pipe2_filter_warp<T, S, TB, E, (G_ < S ? G_ : 0), AM>(q, pc, &tmap, base, ntiles, fresh, g, c0, row_first, lane, pr, prof);
      HGA_FW(0) HGA_FW(1) HGA_FW(2) HGA_FW(3)
      if (S > 4) { HGA_FW(4) HGA_FW(5) HGA_FW(6) HGA_FW(7) }
#undef HGA_FW
    } else {
      pipe2_filter_warp<T, S, TB, E, -1, AM>(q, pc, &tmap, base, ntiles, fresh, g, c0, row_first, lane, pr, prof);
    }
  } else if (g == S + NF) {
    // ------------------------------------------------------------------ TMA producer warp (float64 mode; in float32 mode
    // the converting framing warps are the consumers of the x ring and refill it themselves)
    if (CPL == 1 && lane == 0) {
      uint64_t* empty_x = bars + SM::b_empty_x;
      for (int m = NX; m < ntiles; ++m) {
        const int sx = m % NX;
        mbar_wait(&empty_x[sx], ((m / NX) - 1) & 1);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&full_x[sx], kTileBytes);
        tma_load_2d(xring + (size_t)sx * TB * CW, &tmap, c0, row_first + m * TB, &full_x[sx]);
      }
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
    constexpr int NU = NBT * CPL;  // (hop block, channel half) units per tile
    long long tk = prof ? clock64() : 0;
#define HGA_TICK2(slot)               \
  if (prof) {                         \
    const long long now_ = clock64(); \
    pr[slot] += now_ - tk;            \
    tk = now_;                        \
  }
    uint64_t* bs_full = bars + SM::b_bs;
    constexpr int NBS = SM::NBS;
    int sy = 0, py = 0;
    // Iteration m: [float32 mode: convert x tile m + LEAD if it is this warp's] -> phase A of y tile m (block sums) ->
    // phase B of y tile m - 1 (frames).  Phase B runs one tile late so that a warp delayed by a conversion does not
    // stall the others: they only meet through the bs_full barriers.
    for (int m = (CPL == 2 ? -LEAD : 0); m <= ntiles; ++m) {
      if (CPL == 2) {
        const int j = m + LEAD;
        if (j < ntiles && (j % NF) == w) pipe2_convert_tile<T, S, TB>(&tmap, base, j, ntiles, c0, row_first, lane);
        HGA_TICK2(0)
        if (m < 0) continue;
      }
      if (m < ntiles) {
      mbar_wait(&full_y[sy], py);
      HGA_TICK2(1)
      const T* yt = ylink + (size_t)sy * TB * 32 + lane;
      const int64_t brel0 = (int64_t)m * NBT - DP / HOP;  // item-relative index of the tile's first hop block
      // phase A: hop-block power sums; unit u = (block j, half p) goes to framing warp u % NF
#pragma unroll
      for (int u0 = 0; u0 < NU; u0 += NF) {
        const int u = u0 + w;
        if (u < NU) {
          const int j = u / CPL, p = u % CPL;
          const int64_t brel = brel0 + j, tb = wk.t_begin + brel * HOP;
          if (brel >= 0 && tb >= wk.t_acc && tb + HOP <= wk.t_end) {
            double bsum;
            const bool add_zir = fresh && tb < a.n_zir;  // n_zir is a multiple of HOP: the whole block is inside
            if (CPL == 1) {
              double v[HOP];
#pragma unroll
              for (int i = 0; i < HOP; ++i) v[i] = TR::load_part(yt + (j * HOP + i) * 32, p);
              if (add_zir) {
#pragma unroll
                for (int i = 0; i < HOP; ++i) v[i] += a.zir[tb + i];
              }
              double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
              for (int i = 0; i < HOP; i += 2) {
                acc0 = fma(v[i], v[i], acc0);
                acc1 = fma(v[i + 1], v[i + 1], acc1);
              }
              bsum = acc0 + acc1;
            } else {
              // float32 mode: squares and the 10 / 20-term block sum stay in float32 (relative error ~1e-7, far inside the
              // mode's tolerance); float <-> double conversions are slow on this part, so only the block sum is widened
              const float* yf = reinterpret_cast<const float*>(yt + j * HOP * 32) + p;
              float v[HOP];
#pragma unroll
              for (int i = 0; i < HOP; ++i) v[i] = yf[i * 64];
              if (add_zir) {
#pragma unroll
                for (int i = 0; i < HOP; ++i) v[i] += (float)a.zir[tb + i];
              }
              float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
              for (int i = 0; i < HOP; i += 2) {
                acc0 = fmaf(v[i], v[i], acc0);
                acc1 = fmaf(v[i + 1], v[i + 1], acc1);
              }
              reinterpret_cast<float*>(bs)[((blk_begin + brel) & (SM::kBsRing - 1)) * CW + lane + 32 * p] = acc0 + acc1;
              bsum = 0.0;
            }
            if (CPL == 1) bs[((blk_begin + brel) & (SM::kBsRing - 1)) * CW + lane + 32 * p] = bsum;
          }
        }
      }
      HGA_TICK2(2)
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&empty_y[sy]);
        mbar_arrive(&bs_full[m % NBS]);
      }
      if (++sy == NBY) { sy = 0; py ^= 1; }
      }
      if (m == 0) continue;
      const int mb = m - 1;  // tile whose frames are emitted now
      mbar_wait(&bs_full[mb % NBS], (mb / NBS) & 1);
      HGA_TICK2(4)
      const int64_t brel0b = (int64_t)mb * NBT - DP / HOP;
      // phase B: frames closed by this tile's blocks; unit u goes to framing warp (u + NU) % NF
#pragma unroll
      for (int u0 = 0; u0 < NU; u0 += NF) {
        const int u = u0 + (w + NF - NU % NF) % NF;
        if (u < NU) {
          const int j = u / CPL, p = u % CPL;
          const int64_t brel = brel0b + j, tb = wk.t_begin + brel * HOP, blk = blk_begin + brel;
          if (brel >= 0 && tb >= wk.t_acc + (int64_t)(kPipeNB - 1) * HOP && tb + HOP <= wk.t_end) {
            double vv[kPipeNB];
#pragma unroll
            for (int k = 0; k < kPipeNB; ++k) {
              const int bi = ((blk - (kPipeNB - 1) + k) & (SM::kBsRing - 1)) * CW + lane + 32 * p;
              vv[k] = CPL == 1 ? bs[bi] : (double)reinterpret_cast<const float*>(bs)[bi];  // widened here: window sums, log, z-score are float64
            }
            double sacc = vv[0];
#pragma unroll
            for (int k = 1; k < kPipeNB; ++k) sacc += vv[k];  // oldest block first, like the sequential kernel
            double f = tab_log(fma(sacc, pc.pow_scale, a.eps), logtab);
            const double mu_p = CPL == 1 ? mu[0] : (p ? mu[CPL - 1] : mu[0]);
            const double is_p = CPL == 1 ? inv_sigma[0] : (p ? inv_sigma[CPL - 1] : inv_sigma[0]);
            if (zs) f = (f - mu_p) * is_p;
            if (c0 + lane + 32 * p < a.C) op[(blk - (kPipeNB - 1)) * a.ldo + 32 * p] = f;
          }
        }
      }
      HGA_TICK2(3)
    }
#undef HGA_TICK2
  }
  if (prof && lane == 0) {
    pr[5] = clock64() - tstart;
    for (int i = 0; i < 6; ++i) a.prof[((size_t)blockIdx.x * 16 + g) * 6 + i] = pr[i];
  }
}

}  // namespace hga
