// hga_kernels_exact.cuh — reference-order ("exact") kernels: sequential IIR cascade, row-order window sums, z-score,
// the single-launch streaming push, and the Chan-merge baseline statistics.
#pragma once
#include "hga_common.cuh"

namespace hga {

struct Segment {       // one independent stream (trial / session): rows [row0, row0 + T) of x
  int64_t row0, T, out_row0;
};

// ---------------------------------------------------------------------------------------------------------------
// K-iir (exact): thread per (segment, channel), sequential in time.  scipy.signal.sosfilt x2 (units.py:151-152).
// state_in == nullptr: every segment starts from sosfilt_zi (fresh extractor).  state layout (ns, 2, C).
// ---------------------------------------------------------------------------------------------------------------
template <int ARITH>
__global__ void __launch_bounds__(128) k_iir_seq(const __grid_constant__ Cascade q, const double* __restrict__ x,
                                                 int64_t ldx, double* __restrict__ y, int64_t ldy,
                                                 const Segment* __restrict__ segs, Segment one, int n_segs, int C,
                                                 const double* __restrict__ state_in, double* __restrict__ state_out,
                                                 const Segment* __restrict__ ysegs = nullptr) {
  const int cg = (C + 31) / 32;
  const int warp = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  if (warp >= n_segs * cg) return;
  const int seg = warp / cg, c = (warp % cg) * 32 + (threadIdx.x & 31);
  if (c >= C) return;
  const Segment sg = segs ? segs[seg] : one;
  // ysegs: where each segment's filtered rows go (segments may overlap in x: every one gets its own rows of y)
  const int64_t y_row0 = ysegs ? ysegs[seg].row0 : sg.row0;
  double z0[kMaxSec], z1[kMaxSec];
#pragma unroll
  for (int s = 0; s < kMaxSec; ++s) {
    z0[s] = state_in ? (s < q.ns ? state_in[(int64_t)(2 * s) * C + c] : 0.0) : q.zi0[s];
    z1[s] = state_in ? (s < q.ns ? state_in[(int64_t)(2 * s + 1) * C + c] : 0.0) : q.zi1[s];
  }
  const double* xp = x + sg.row0 * ldx + c;
  double* yp = y + y_row0 * ldy + c;
  constexpr int U = 8;
  int64_t t = 0;
  for (; t + U <= sg.T; t += U) {
    double v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(xp + (t + u) * ldx);
#pragma unroll
    for (int u = 0; u < U; ++u) yp[(t + u) * ldy] = cascade<ARITH, 0, kMaxSec>(q, v[u], z0, z1);
  }
  for (; t < sg.T; ++t) yp[t * ldy] = cascade<ARITH, 0, kMaxSec>(q, __ldg(xp + t * ldx), z0, z1);
  if (state_out) {
#pragma unroll
    for (int s = 0; s < kMaxSec; ++s)
      if (s < q.ns) {
        state_out[(int64_t)(2 * s) * C + c] = z0[s];
        state_out[(int64_t)(2 * s + 1) * C + c] = z1[s];
      }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// K-frame (exact): compute_log_power_features (hga_optimized.pyx:8-47) + optional z-score (common.py:375-376).
// Thread per (window, channel); the window's squares are added in row order, unfused, like the reference loop.
// Rows [0, zero_rows) of the logical matrix are zeros (warm-start padding, pyx:111-122); row r >= zero_rows is
// data[(r - zero_rows) * ld + c].  Per-segment batching: window w of segment s reads rows relative to seg.row0.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kFramesPerBlock = 8;   // 32 (a forward sweep of 4 windows per warp for L1 reuse) measured slower: 3.26 vs 1.72 ms on K2

// Regular windows with L = NB * hop (the path's 50 ms / 10 ms geometry, NB = 5), no leading zero rows: a thread walks down
// its channel over a run of R consecutive windows and keeps the NB overlapping windows' sums live in registers.  Every
// sample is loaded and squared once and added to all NB sums -- each sum still receives its rows in row order starting from
// 0.0, so the result is bit-identical to the per-window loop (k_frames_seq) and to the reference's `array_sum_and_power`.
// Slot k of acc belongs to the window whose first hop-block b has b % NB == k; it is cleared when that block starts and
// emitted NB - 1 blocks later.  Loads: 1 + (NB - 1) / R per sample instead of NB.
constexpr int kRollWarps = 8;
template <int NB>
__global__ void __launch_bounds__(32 * kRollWarps) k_frames_roll(const double* __restrict__ data, int64_t ld, const Segment* __restrict__ segs,
                                                                 Segment one, int n_segs, int C, int hop, int R, double eps,
                                                                 const double* __restrict__ mu, const double* __restrict__ sigma,
                                                                 double* __restrict__ out, int64_t ldo) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int seg = blockIdx.z;
  if (c >= C || seg >= n_segs) return;
  const Segment sg = segs ? segs[seg] : one;
  const int L = NB * hop;
  if (sg.T < L) return;
  const int64_t F = (sg.T - L) / hop + 1;
  const int64_t w0 = ((int64_t)blockIdx.y * kRollWarps + (threadIdx.x >> 5)) * R;
  if (w0 >= F) return;
  const int64_t n = min((int64_t)R, F - w0), nblk = n + NB - 1;
  const double* dp = data + (sg.row0 + w0 * hop) * ld + c;
  double* op = out + (sg.out_row0 + w0) * ldo + c;
  const bool zs = mu != nullptr;
  const double m = zs ? mu[c] : 0.0, sd = zs ? sigma[c] : 1.0;
  double acc[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) acc[k] = 0.0;
  for (int64_t b0 = 0; b0 < nblk; b0 += NB) {
#pragma unroll
    for (int k = 0; k < NB; ++k) {
      const int64_t b = b0 + k;
      if (b < nblk) {
        acc[k] = 0.0;
        const double* bp = dp + b * hop * ld;
        for (int j0 = 0; j0 < hop; j0 += 10) {
          double v[10];
#pragma unroll
          for (int u = 0; u < 10; ++u) v[u] = j0 + u < hop ? __ldg(bp + (int64_t)(j0 + u) * ld) : 0.0;   // + 0.0 leaves a sum as it is
#pragma unroll
          for (int u = 0; u < 10; ++u) {
            const double sq = __dmul_rn(v[u], v[u]);
#pragma unroll
            for (int q = 0; q < NB; ++q) acc[q] = __dadd_rn(acc[q], sq);
          }
        }
        const int64_t i = b - (NB - 1);
        if (i >= 0) op[i * ldo] = feature(acc[(k + 1) % NB], (double)L, eps, zs, m, sd);
      }
    }
  }
}
// run length: enough warps for two waves of 64 per SM, at most 128 windows (halo (NB - 1) / R <= 3 %), at least 16
inline int roll_run_length(int64_t frames_total, int cg, int n_sm) {
  const int64_t r = frames_total * cg / ((int64_t)n_sm * 128);
  return (int)(r < 16 ? 16 : r > 128 ? 128 : r);
}
__global__ void __launch_bounds__(256) k_frames_seq(const double* __restrict__ data, int64_t ld, int64_t zero_rows,
                                                    const Segment* __restrict__ segs, Segment one, int n_segs, int C,
                                                    int64_t frames_per_seg_max, int L, int hop,
                                                    const int64_t* __restrict__ start, const int64_t* __restrict__ stop,
                                                    double eps, const double* __restrict__ mu,
                                                    const double* __restrict__ sigma, double* __restrict__ out,
                                                    int64_t ldo) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int seg = blockIdx.z;
  if (c >= C || seg >= n_segs) return;
  const Segment sg = segs ? segs[seg] : one;
  // a block covers kFramesPerBlock consecutive windows, 8 at a time (one per warp); consecutive windows share 80 % of
  // their rows, served by L1 / L2
  for (int k = 0; k < kFramesPerBlock / 8; ++k) {
  const int64_t w = (int64_t)blockIdx.y * kFramesPerBlock + k * 8 + threadIdx.x / 32;
  if (w >= frames_per_seg_max) return;
  // frames of this segment: explicit arrays give the count through frames_per_seg_max (single segment only)
  int64_t a, b;
  if (start) {
    a = start[w];
    b = stop[w];
  } else {
    const int64_t rows = sg.T + zero_rows;
    if (rows < L || w > (rows - L) / hop) return;
    a = w * hop;
    b = a + L;
  }
  const double* dp = data + sg.row0 * ld + c;
  double acc = 0.0;
  for (int64_t r = a; r < b; ++r) {
    const double v = r < zero_rows ? 0.0 : dp[(r - zero_rows) * ld];
    acc = __dadd_rn(acc, __dmul_rn(v, v));
  }
  const bool zs = mu != nullptr;
  out[(sg.out_row0 + w) * ldo + c] = feature(acc, (double)(b - a), eps, zs, zs ? mu[c] : 0.0, zs ? sigma[c] : 1.0);
  }
}

// A thread keeps its channel (mu, sigma in registers) and walks rows: no integer division per element, four independent
// row accesses in flight; a warp reads / writes whole 256 B row segments.  Bound: HBM (8 B in + 8 B out per element).
__global__ void __launch_bounds__(256) k_zscore(const double* __restrict__ f, int64_t F, int C, int64_t ld, const double* __restrict__ mu,
                                                const double* __restrict__ sigma, double* __restrict__ out, int64_t ldo) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  if (c >= C) return;
  const double m = mu[c], sd = sigma[c];
  const int64_t step = (int64_t)gridDim.y * 8;
  int64_t r = (int64_t)blockIdx.y * 8 + (threadIdx.x >> 5);
  for (; r + 3 * step < F; r += 4 * step) {
    const double v0 = f[r * ld + c], v1 = f[(r + step) * ld + c], v2 = f[(r + 2 * step) * ld + c], v3 = f[(r + 3 * step) * ld + c];
    out[r * ldo + c] = __ddiv_rn(__dsub_rn(v0, m), sd);
    out[(r + step) * ldo + c] = __ddiv_rn(__dsub_rn(v1, m), sd);
    out[(r + 2 * step) * ldo + c] = __ddiv_rn(__dsub_rn(v2, m), sd);
    out[(r + 3 * step) * ldo + c] = __ddiv_rn(__dsub_rn(v3, m), sd);
  }
  for (; r < F; r += step) out[r * ldo + c] = __ddiv_rn(__dsub_rn(f[r * ld + c], m), sd);
}

// The same map with four adjacent channels per thread: one 256-bit load and one 256-bit store per row (ld / st.global.v4.f64;
// a warp moves 1 KB of whole lines per instruction -- see tools/microbench/write_rates.cu for what the store width buys).
// Needs C, ld, ldo multiples of 4 and 32 B-aligned f / out / mu / sigma; same IEEE operations per element, same bits.
struct ZQuad { double v[4]; };
__device__ __forceinline__ ZQuad zquad_load(const double* p) {
  ZQuad q;
  asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(q.v[0]), "=d"(q.v[1]), "=d"(q.v[2]), "=d"(q.v[3]) : "l"(p));
  return q;
}
__device__ __forceinline__ void zquad_store(double* p, const ZQuad& q) {
  asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(q.v[0]), "d"(q.v[1]), "d"(q.v[2]), "d"(q.v[3]) : "memory");
}
__global__ void __launch_bounds__(256) k_zscore4(const double* __restrict__ f, int64_t F, int C, int64_t ld, const double* __restrict__ mu,
                                                 const double* __restrict__ sigma, double* __restrict__ out, int64_t ldo) {
  const int c = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
  if (c >= C) return;
  const ZQuad m = zquad_load(mu + c), sd = zquad_load(sigma + c);
  auto z = [&](const ZQuad& v) {
    ZQuad o;
#pragma unroll
    for (int k = 0; k < 4; ++k) o.v[k] = __ddiv_rn(__dsub_rn(v.v[k], m.v[k]), sd.v[k]);
    return o;
  };
  const int64_t step = (int64_t)gridDim.y * 8;
  int64_t r = (int64_t)blockIdx.y * 8 + (threadIdx.x >> 5);
  for (; r + 3 * step < F; r += 4 * step) {
    ZQuad v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = zquad_load(f + (r + j * step) * ld + c);
#pragma unroll
    for (int j = 0; j < 4; ++j) zquad_store(out + (r + j * step) * ldo + c, z(v[j]));
  }
  for (; r < F; r += step) zquad_store(out + r * ldo + c, z(zquad_load(f + r * ld + c)));
}

// ---------------------------------------------------------------------------------------------------------------
// K-stream: one launch per push = sosfilt x2 with carried state (units.py:151-152) + WarmStartFrameBuffer.insert
// (pyx:96-131) + compute_log_power_features (pyx:27-47) + optional z-score.  Thread per channel; every thread owns
// its channel's column of the frame buffer `buf` (cap_rows x C), so no inter-thread synchronisation is needed.
//   y_row0     : row of buf where the new filtered rows go (case 1: 0, case 2: L - n, case 3: overlap)
//   zero_rows  : rows [0, zero_rows) of buf are zero-filled first (case 2)
//   rows_total : rows of the logical buffer the windows run over
// After the frames, the last `overlap` rows are moved to the front (remainder, pyx:106,121,130).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) k_stream_push(const __grid_constant__ Cascade q, const double* __restrict__ x,
                                                    int64_t ldx, int n, int C, double* __restrict__ state,
                                                    double* __restrict__ buf, int y_row0, int zero_rows,
                                                    int rows_total, int overlap, int n_frames, int L, int hop,
                                                    double eps, const double* __restrict__ mu,
                                                    const double* __restrict__ sigma, double* __restrict__ out,
                                                    int64_t ldo, const float* __restrict__ xf = nullptr,
                                                    float* __restrict__ out_f32 = nullptr, int staged = 0) {
  // xf != nullptr: the input is the payload of a wire packet, float32 [C][n] channel-major (local/units.py:78-82:
  // frombuffer(float32).reshape(n_channels, n_samples), transposed, upcast to float64) read in place of x.
  // out_f32 != nullptr: the frames are also written as float32 (n_frames, C) for the nVAD's `.float()` (local/units.py:433).
  // staged != 0: x / xf is host memory mapped into the device's address space (zero-copy packet path): the block's 32
  // channels x n samples are first pulled into shared memory with coalesced reads, all of them in flight at once, so the
  // packet crosses PCIe in one round trip instead of one dependent read per sample.  Dynamic shared memory: n * 32 * 8 B.
  extern __shared__ __align__(16) unsigned char stage_raw[];
  const int c = blockIdx.x * 32 + threadIdx.x;
  if (staged) {
    if (xf != nullptr) {
      float* sf = reinterpret_cast<float*>(stage_raw);
      const int c0 = blockIdx.x * 32, cnt = min(32, C - c0) * n;
      const float* src = xf + (int64_t)c0 * n;
      for (int i = threadIdx.x; i < cnt; i += 32) sf[i] = src[i];
    } else if (c < C) {
      double* sd = reinterpret_cast<double*>(stage_raw);
      for (int t = 0; t < n; ++t) sd[t * 32 + threadIdx.x] = x[(int64_t)t * ldx + c];
    }
    __syncwarp();
  }
  if (c >= C) return;
  double z0[kMaxSec], z1[kMaxSec];
#pragma unroll
  for (int s = 0; s < kMaxSec; ++s) {
    z0[s] = s < q.ns ? state[(int64_t)(2 * s) * C + c] : 0.0;
    z1[s] = s < q.ns ? state[(int64_t)(2 * s + 1) * C + c] : 0.0;
  }
  double* bp = buf + c;
  for (int r = 0; r < zero_rows; ++r) bp[(int64_t)r * C] = 0.0;
  const double* xp = x + c;
  if (staged && xf != nullptr) {
    const float* fp = reinterpret_cast<const float*>(stage_raw) + threadIdx.x * n;
    for (int t = 0; t < n; ++t) bp[(int64_t)(y_row0 + t) * C] = cascade<kExact, 0, kMaxSec>(q, (double)fp[t], z0, z1);
  } else if (staged) {
    const double* sp = reinterpret_cast<const double*>(stage_raw) + threadIdx.x;
    for (int t = 0; t < n; ++t) bp[(int64_t)(y_row0 + t) * C] = cascade<kExact, 0, kMaxSec>(q, sp[t * 32], z0, z1);
  } else if (xf != nullptr) {
    const float* fp = xf + (int64_t)c * n;
    for (int t = 0; t < n; ++t) bp[(int64_t)(y_row0 + t) * C] = cascade<kExact, 0, kMaxSec>(q, (double)fp[t], z0, z1);
  } else {
    for (int t = 0; t < n; ++t)
      bp[(int64_t)(y_row0 + t) * C] = cascade<kExact, 0, kMaxSec>(q, xp[(int64_t)t * ldx], z0, z1);
  }
#pragma unroll
  for (int s = 0; s < kMaxSec; ++s)
    if (s < q.ns) {
      state[(int64_t)(2 * s) * C + c] = z0[s];
      state[(int64_t)(2 * s + 1) * C + c] = z1[s];
    }
  const bool zs = mu != nullptr;
  const double m = zs ? mu[c] : 0.0, sd = zs ? sigma[c] : 1.0;
  for (int w = 0; w < n_frames; ++w) {
    double acc = 0.0;
    const int a = w * hop;
    for (int r = a; r < a + L; ++r) {
      const double v = bp[(int64_t)r * C];
      acc = __dadd_rn(acc, __dmul_rn(v, v));
    }
    const double f = feature(acc, (double)L, eps, zs, m, sd);
    out[(int64_t)w * ldo + c] = f;
    if (out_f32 != nullptr) out_f32[(int64_t)w * C + c] = (float)f;
  }
  const int src = rows_total - overlap;
  if (src > 0) {
    // src > dst: forward is safe.  Batches of 8 rows: the 8 loads of a batch are independent (all issued before the first
    // store), where the row-by-row form is one dependent L2 round trip per row (the ranges overlap, so the compiler must
    // order every load after the previous store): 40 rows cost ~14 us of a 20 us kernel that way.
    int r = 0;
    for (; r + 8 <= overlap; r += 8) {
      double v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = bp[(int64_t)(src + r + k) * C];
#pragma unroll
      for (int k = 0; k < 8; ++k) bp[(int64_t)(r + k) * C] = v[k];
    }
    for (; r < overlap; ++r) bp[(int64_t)r * C] = bp[(int64_t)(src + r) * C];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// K-stream-sys: k_stream_push for small packets (rows_total + n <= kStreamFastRows), the same arithmetic bit for bit,
// laid out for LATENCY instead of one thread per channel:
//   * a warp owns 2 channels; lane = (channel, section): the 16 biquad sections of a channel sit in 16 lanes and hand
//     their output to the next lane with a shuffle (a systolic array: section s works on sample k - s at step k).  A
//     10-sample packet takes 25 steps of ONE section update each instead of 160 dependent section updates per thread
//     (k_stream_push: 47 k cycles = 24 us per packet under ncu, almost all of it dependent-issue stalls of a lone warp);
//   * the warp's slice of the frame buffer and of the packet lives in shared memory for the call; every global access
//     belongs to one of three independent bursts: (1) state, remainder rows and packet in, (2) frames out, (3) new
//     remainder and state out;
//   * frame w of a channel is summed in row order (like hga_optimized.pyx:8-24) by lane w % 16 of that channel.
//   IN = 0: x is float64 (n, ldx) row-major;  IN = 1: x is the float32 [C][n] channel-major payload of a wire packet.
// x / out may be mapped host memory (zero-copy packets).  Dynamic shared memory: kSysWarps * (rows_total + n) * 16 bytes.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kStreamFastRows = 192;
constexpr int kSysWarps = 4;
template <int IN>
__global__ void __launch_bounds__(32 * kSysWarps) k_stream_push_sys(const __grid_constant__ Cascade q, const void* __restrict__ xin, int64_t ldx,
                                                                    int n, int C, double* __restrict__ state, double* __restrict__ buf,
                                                                    int y_row0, int zero_rows, int rows_total, int overlap, int n_frames,
                                                                    int L, int hop, double eps, const double* __restrict__ mu,
                                                                    const double* __restrict__ sigma, double* __restrict__ out,
                                                                    int64_t ldo, float* __restrict__ out_f32,
                                                                    unsigned long long* __restrict__ done_slots = nullptr,
                                                                    unsigned long long done_seq = 0) {
  // done_slots != nullptr (mapped host memory, one slot per warp): once a warp's frames are stored, it publishes
  // done_seq in its slot; the host polls the slots instead of synchronising the stream (the frames are all it needs:
  // remainder and state are device-side and the next push is stream-ordered behind this kernel).
  extern __shared__ __align__(16) double ssm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s = lane & 15, chl = lane >> 4;
  const int cbase = (blockIdx.x * kSysWarps + warp) * 2, c = cbase + chl;
  if (cbase >= C) return;                                   // whole warp out of range
  const bool live = c < C;
  double* fb = ssm + (size_t)warp * (rows_total + n) * 2;   // [rows_total][2] frame rows of the warp's two channels
  double* xs = fb + rows_total * 2;                         // [n][2] the packet's samples
  const int ns = q.ns;
  const bool sec = s < ns;
  // ---- burst 1
  double b0 = 0.0, b1 = 0.0, b2 = 0.0, a1 = 0.0, a2 = 0.0, z0 = 0.0, z1 = 0.0;
  if (sec) {
    b0 = q.b0[s]; b1 = q.b1[s]; b2 = q.b2[s]; a1 = q.a1[s]; a2 = q.a2[s];
    if (live) {
      z0 = state[(int64_t)(2 * s) * C + c];
      z1 = state[(int64_t)(2 * s + 1) * C + c];
    }
  }
  for (int i = lane; i < 2 * n; i += 32) {
    const int ch = i / n, t = i - ch * n, cc = cbase + ch;
    double v = 0.0;
    if (cc < C) v = IN == 1 ? (double)static_cast<const float*>(xin)[(int64_t)cc * n + t] : static_cast<const double*>(xin)[(int64_t)t * ldx + cc];
    xs[t * 2 + ch] = v;
  }
  for (int i = lane; i < 2 * y_row0; i += 32) {
    const int r = i >> 1, cc = cbase + (i & 1);
    fb[i] = (r >= zero_rows && cc < C) ? buf[(int64_t)r * C + cc] : 0.0;
  }
  __syncwarp();
  // ---- the cascade: section s takes sample k - s at step k (reference operation order, every product and sum rounded)
  double ylast = 0.0;
  const int nsteps = n + ns - 1;
  for (int k = 0; k < nsteps; ++k) {
    double xi = __shfl_up_sync(0xffffffffu, ylast, 1, 16);
    const int t = k - s;
    if (s == 0 && k < n) xi = xs[k * 2 + chl];
    if (sec && t >= 0 && t < n) {
      const double y = __dadd_rn(__dmul_rn(b0, xi), z0);
      z0 = __dadd_rn(__dsub_rn(__dmul_rn(b1, xi), __dmul_rn(a1, y)), z1);
      z1 = __dsub_rn(__dmul_rn(b2, xi), __dmul_rn(a2, y));
      ylast = y;
      if (s == ns - 1) fb[(y_row0 + t) * 2 + chl] = y;
    }
  }
  __syncwarp();
  // ---- burst 2: frames
  if (live) {
    const bool zs = mu != nullptr;
    const double m = zs ? mu[c] : 0.0, sd = zs ? sigma[c] : 1.0;
    for (int w = s; w < n_frames; w += 16) {
      double acc = 0.0;
      const int a = w * hop;
      for (int r = a; r < a + L; ++r) {
        const double v = fb[r * 2 + chl];
        acc = __dadd_rn(acc, __dmul_rn(v, v));
      }
      const double f = feature(acc, (double)L, eps, zs, m, sd);
      out[(int64_t)w * ldo + c] = f;
      if (out_f32 != nullptr) out_f32[(int64_t)w * C + c] = (float)f;
    }
  }
  if (done_slots != nullptr) {
    __threadfence_system();   // every lane: its frame stores are visible system-wide before the slot is
    __syncwarp();
    if (lane == 0) {
      *reinterpret_cast<volatile unsigned long long*>(done_slots + (cbase >> 1)) = done_seq;
      __threadfence_system();
    }
  }
  // ---- burst 3: new remainder (the last `overlap` rows) and filter state
  const int src_row = rows_total - overlap;
  if (src_row >= 0)
    for (int i = lane; i < 2 * overlap; i += 32) {
      const int r = i >> 1, ch = i & 1, cc = cbase + ch;
      if (cc < C) buf[(int64_t)r * C + cc] = fb[(src_row + r) * 2 + ch];
    }
  if (live && sec) {
    state[(int64_t)(2 * s) * C + c] = z0;
    state[(int64_t)(2 * s + 1) * C + c] = z1;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// K-packet: payload of a wire packet, float32 [C][n] channel-major -> float64 (n, C) row-major (ZMQConnector.
// interpret_bytes, local/units.py:78-82).  A 32 x 32 tile through shared memory: coalesced reads along n, coalesced
// writes along C.  Bound: HBM / launch latency (a packet is 5 KB).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_packet_rows(const float* __restrict__ p, int C, int n, double* __restrict__ x, int64_t ldx) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int c0 = blockIdx.x * 32, t0 = blockIdx.y * 32;
  for (int k = ty; k < 32; k += 8) {
    const int c = c0 + k, t = t0 + tx;
    tile[k][tx] = (c < C && t < n) ? p[(int64_t)c * n + t] : 0.f;
  }
  __syncthreads();
  for (int k = ty; k < 32; k += 8) {
    const int t = t0 + k, c = c0 + tx;
    if (t < n && c < C) x[(int64_t)t * ldx + c] = (double)tile[tx][k];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// K-stats: per-channel (n, mean, M2) of F feature rows (np.mean / np.std over frames, prepare_corpus.py:198-199).
// Each thread reduces a strided set of rows with shifted sums (shift = its first value), partials are Chan-merged in
// a fixed order: 8 row-lanes per block through shared memory, then blocks by k_stats_merge.  acc layout (3, C).
// ---------------------------------------------------------------------------------------------------------------
struct Moments {
  double n, mean, m2;
};
__device__ __forceinline__ Moments chan_merge(Moments a, Moments b) {
  const double n = a.n + b.n;
  if (n == 0.0) return a;
  const double w = b.n / n, d = b.mean - a.mean;
  return Moments{n, fma(d, w, a.mean), a.m2 + b.m2 + d * d * a.n * w};
}

// Statistics that left the extraction kernels as partials (hga_extract_offline_stats): per channel, the (n, mean, M2) of
// every (item), (fix-up, piece) and (seam) that emitted frames of it.  One warp per channel: lane l Chan-merges partials
// l, l + 32, ... of the concatenated list in index order, a butterfly with the lower lane's moments first merges the 32
// results (fixed shape: deterministic), lane 0 merges the total into acc (acc first).
__global__ void __launch_bounds__(128) k_lane_stats_merge(double* __restrict__ acc, const double* __restrict__ st_main, int n_work, int groups, int cw,
                                                          const double* __restrict__ st_fix, int n_fixpart, const double* __restrict__ st_seam, int n_seam,
                                                          int C, int CP) {
  // one CTA of four warps per channel: thread t merges partials t, t + 128, ... in index order, each warp folds its 32 results
  // with the butterfly, thread 0 merges the four warp results in warp order
  __shared__ Moments wsum[4];
  const int c = blockIdx.x, t = threadIdx.x, lane = t & 31;
  const int g = c / cw, cl = c - g * cw;
  const int total = n_work + n_fixpart + n_seam;
  Moments m{0.0, 0.0, 0.0};
  for (int i = t; i < total; i += 128) {
    const double* p;
    size_t stride;
    if (i < n_work) {
      p = st_main + ((size_t)(i * groups + g) * 3) * cw + cl;
      stride = (size_t)cw;
    } else if (i < n_work + n_fixpart) {
      p = st_fix + (size_t)(i - n_work) * 3 * CP + c;
      stride = (size_t)CP;
    } else {
      p = st_seam + (size_t)(i - n_work - n_fixpart) * 3 * CP + c;
      stride = (size_t)CP;
    }
    const double pn = p[0];
    if (pn > 0.0) m = chan_merge(m, Moments{pn, p[stride], p[2 * stride]});
  }
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Moments other{__shfl_xor_sync(0xffffffffu, m.n, o), __shfl_xor_sync(0xffffffffu, m.mean, o), __shfl_xor_sync(0xffffffffu, m.m2, o)};
    m = (lane & o) ? chan_merge(other, m) : chan_merge(m, other);
  }
  if (lane == 0) wsum[t >> 5] = m;
  __syncthreads();
  if (t == 0) {
    Moments tot = chan_merge(chan_merge(wsum[0], wsum[1]), chan_merge(wsum[2], wsum[3]));
    tot = chan_merge(Moments{acc[c], acc[C + c], acc[2 * (size_t)C + c]}, tot);
    acc[c] = tot.n;
    acc[C + c] = tot.mean;
    acc[2 * (size_t)C + c] = tot.m2;
  }
}

constexpr int kStatsRowsPerBlock = 8;
__global__ void __launch_bounds__(256, 8) k_stats_partial(const double* __restrict__ f, int64_t F, int C, int64_t ld,
                                                       int64_t rows_per_block, double* __restrict__ part) {
  __shared__ Moments sm[kStatsRowsPerBlock][33];
  const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t r1 = min(F, r0 + rows_per_block);
  Moments m{0.0, 0.0, 0.0};
  if (c < C && r0 + ty < r1) {
    const double k = f[(r0 + ty) * ld + c];
    double s1 = 0.0, s2 = 0.0, cnt = 0.0;
    // four loads in flight per thread (the accumulation chain is loop-carried, so the compiler does not hoist them itself:
    // one load in flight per warp gave 2.35 TB/s); the values are accumulated in the same order as row by row
    int64_t r = r0 + ty;
    constexpr int64_t S = kStatsRowsPerBlock;
    for (; r + 3 * S < r1; r += 4 * S) {
      const double v0 = f[r * ld + c], v1 = f[(r + S) * ld + c], v2 = f[(r + 2 * S) * ld + c], v3 = f[(r + 3 * S) * ld + c];
      double d = v0 - k; s1 += d; s2 = fma(d, d, s2);
      d = v1 - k; s1 += d; s2 = fma(d, d, s2);
      d = v2 - k; s1 += d; s2 = fma(d, d, s2);
      d = v3 - k; s1 += d; s2 = fma(d, d, s2);
      cnt += 4.0;
    }
    for (; r < r1; r += S) {
      const double d = f[r * ld + c] - k;
      s1 += d;
      s2 = fma(d, d, s2);
      cnt += 1.0;
    }
    m = Moments{cnt, k + s1 / cnt, s2 - s1 * s1 / cnt};
  }
  sm[ty][lane] = m;
  __syncthreads();
  if (ty == 0 && c < C) {
    Moments a = sm[0][lane];
    for (int j = 1; j < kStatsRowsPerBlock; ++j) a = chan_merge(a, sm[j][lane]);
    double* p = part + (int64_t)blockIdx.y * 3 * C;
    p[c] = a.n;
    p[C + c] = a.mean;
    p[2 * C + c] = a.m2;
  }
}
// acc <- merge(acc, part[0], part[1], ..., part[count-1]) in that order (deterministic).
__global__ void k_stats_merge(double* __restrict__ acc, const double* __restrict__ part, int count, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  Moments a{acc[c], acc[C + c], acc[2 * C + c]};
  for (int j = 0; j < count; ++j) {
    const double* p = part + (int64_t)j * 3 * C;
    a = chan_merge(a, Moments{p[c], p[C + c], p[2 * C + c]});
  }
  acc[c] = a.n;
  acc[C + c] = a.mean;
  acc[2 * C + c] = a.m2;
}
// Day-baseline exchange in one launch: gathered = (n_src, n_days, 3, C) accumulators (rank-major, as all-gathered); for
// every (day, channel) the n_src accumulators are merged in source order (the same on every rank -> bit-identical results)
// and finalised: mu = mean, sigma = sqrt(M2 / n) (np.std, ddof = 0).  Days nobody contributed to get mu = 0, sigma = 1.
__global__ void k_stats_merge_days(const double* __restrict__ gathered, int n_src, int n_days, int C, double* __restrict__ mu,
                                   double* __restrict__ sigma) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, d = blockIdx.y;
  if (c >= C || d >= n_days) return;
  Moments a{0.0, 0.0, 0.0};
  for (int r = 0; r < n_src; ++r) {
    const double* p = gathered + ((int64_t)r * n_days + d) * 3 * C;
    a = chan_merge(a, Moments{p[c], p[C + c], p[2 * C + c]});
  }
  mu[(int64_t)d * C + c] = a.n > 0.0 ? a.mean : 0.0;
  sigma[(int64_t)d * C + c] = a.n > 0.0 ? sqrt(a.m2 / a.n) : 1.0;
}
__global__ void k_stats_finalize(const double* __restrict__ acc, int C, double* __restrict__ mu,
                                 double* __restrict__ sigma) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  mu[c] = acc[C + c];
  sigma[c] = sqrt(acc[2 * C + c] / acc[c]);
}

}  // namespace hga
