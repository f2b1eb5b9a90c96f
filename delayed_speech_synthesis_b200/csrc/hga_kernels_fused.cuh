// hga_kernels_fused.cuh — K-hga (f64): IIR cascade + hop-block power sums + log + z-score in one pass, time-chunked.
//
// Replaces, for one fresh extractor per segment (prepare_corpus.py:47-50), the chain
//   sosfilt(hg) -> sosfilt(fh) -> WarmStartFrameBuffer.insert (case 1) -> compute_log_power_features -> z-score
// (/root/reference/local/units.py:151-160, extensions/hga/hga_optimized.pyx:27-47,104-107, prepare_corpus.py:226-227)
// without materialising the filtered signal.
//
// Time-chunking: a work item covers samples [t_begin, t_end) of a segment.  Items that do not start at the segment
// start begin `warmup` samples early from ZERO state; the cascade's slowest pole (|p| = 0.99758 at 1 kHz) makes the
// missing zero-input response decay below the stated tolerance before the first accumulated sample.
// Window power = sum of L/hop hop-block partial sums (L is a multiple of hop at every configured rate), so chunk
// boundaries need no halo exchange: an item starts accumulating (L/hop - 1) blocks before its first own sample.
#pragma once
#include "hga_common.cuh"

namespace hga {

struct Work {
  int64_t row0;      // first row of the segment in x
  int64_t T;         // segment length
  int64_t t_begin;   // first sample filtered (multiple of hop; 0 = segment start: initial state = zi)
  int64_t t_acc;     // first sample accumulated into hop blocks (multiple of hop)
  int64_t t_end;     // one past the last sample filtered (multiple of hop, or T for the last item of a segment)
  int64_t out_row0;  // output row of the segment's frame 0
  int32_t seg;       // segment index (state_out / remainder_out slot)
  int32_t last;      // 1: last item of its segment (writes final state and remainder rows)
};

constexpr int kMaxBlocksPerFrame = 8;  // L / hop <= 8

struct FrameAcc {
  double ring[kMaxBlocksPerFrame];
  double acc;
  int nblk;
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int i = 0; i < kMaxBlocksPerFrame; ++i) ring[i] = 0.0;
    acc = 0.0;
    nblk = 0;
  }
  // start from the (nb - 1) hop blocks of a (L - hop)-row remainder column (frame-buffer case 3, pyx:123-131)
  __device__ __forceinline__ void preload(const double* col, int64_t ld, int hop, int nb) {
    double dummy;
    for (int b = 0; b < nb - 1; ++b) {
      for (int i = 0; i < hop; ++i) {
        const double v = col[(int64_t)(b * hop + i) * ld];
        acc = fma(v, v, acc);
      }
      close_block(nb, dummy);
    }
  }
  // close the current hop block; returns true (and the window sum, oldest block first) once nb blocks are in
  __device__ __forceinline__ bool close_block(int nb, double& sum) {
#pragma unroll
    for (int i = kMaxBlocksPerFrame - 1; i > 0; --i) ring[i] = ring[i - 1];
    ring[0] = acc;
    acc = 0.0;
    ++nblk;
    double s = 0.0;
#pragma unroll
    for (int i = kMaxBlocksPerFrame - 1; i >= 0; --i)
      if (i < nb) s += ring[i];
    sum = s;
    return nblk >= nb;
  }
};

struct FusedArgs {
  const double* x;
  int64_t ldx;
  const Work* work;
  int n_work;
  int C;
  int L, hop, nb;
  double eps;
  const double* mu;
  const double* sigma;
  double* out;
  int64_t ldo;
  const double* state_in;      // (ns, 2, C) or nullptr: initial state of items with t_begin == 0 (else sosfilt_zi)
  const double* remainder_in;  // (L - hop, C) or nullptr: filtered rows preceding row 0 (frame-buffer case 3)
  double* state_out;      // (n_segs, ns, 2, C) or nullptr
  double* remainder_out;  // (n_segs, L - hop, C) or nullptr: last L-hop filtered rows of each segment
  int64_t x_row_lo, x_row_hi;  // rows [x_row_lo, x_row_hi) of x are addressable (TMA extents of the pipe kernel)
};

// Thread per (work item, channel); a warp spans 32 adjacent channels so every row access is one 256-byte segment.
template <int ARITH, int HOP>
__global__ void __launch_bounds__(128, 3) k_hga_fused_seq(const __grid_constant__ Cascade q,
                                                          const __grid_constant__ FusedArgs a) {
  const int cg = (a.C + 31) / 32;
  const int warp = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (warp >= a.n_work * cg) return;
  const int c = (warp % cg) * 32 + (threadIdx.x & 31);
  if (c >= a.C) return;
  const Work wk = a.work[warp / cg];
  const int hop = HOP > 0 ? HOP : a.hop;

  double z0[kMaxSec], z1[kMaxSec];
  const bool fresh = wk.t_begin == 0;
#pragma unroll
  for (int s = 0; s < kMaxSec; ++s) {
    z0[s] = fresh ? q.zi0[s] : 0.0;
    z1[s] = fresh ? q.zi1[s] : 0.0;
  }
  if (fresh && a.state_in) {
#pragma unroll
    for (int s = 0; s < kMaxSec; ++s)
      if (s < q.ns) {
        z0[s] = a.state_in[(int64_t)(2 * s) * a.C + c];
        z1[s] = a.state_in[(int64_t)(2 * s + 1) * a.C + c];
      }
  }
  FrameAcc fa;
  fa.init();
  // only the item that accumulates from row 0 continues the carried frame buffer; later items that merely START at
  // row 0 (their warm-up reaches back to it) begin accumulating at t_acc > 0 with an empty ring
  if (fresh && a.remainder_in && wk.t_acc == 0) fa.preload(a.remainder_in + c, a.C, hop, a.nb);
  const bool zs = a.mu != nullptr;
  const double mu = zs ? a.mu[c] : 0.0, sigma = zs ? a.sigma[c] : 1.0;
  const double* xp = a.x + wk.row0 * a.ldx + c;
  double* op = a.out + wk.out_row0 * a.ldo + c;
  const int ov = a.L - hop;
  const int64_t rem_t0 = (wk.last && a.remainder_out) ? wk.T - ov : INT64_MAX;
  double* rem = a.remainder_out ? a.remainder_out + ((int64_t)wk.seg * ov) * a.C + c : nullptr;

  int64_t t = wk.t_begin;
  // ---- warm-up: filter only ----
  for (; t + hop <= wk.t_acc; t += hop) {
    if (HOP > 0) {
      double v[HOP > 0 ? HOP : 1];
#pragma unroll
      for (int i = 0; i < HOP; ++i) v[i] = __ldg(xp + (t + i) * a.ldx);
#pragma unroll
      for (int i = 0; i < HOP; ++i) cascade<ARITH, 0, kMaxSec>(q, v[i], z0, z1);
    } else {
      for (int i = 0; i < hop; ++i) cascade<ARITH, 0, kMaxSec>(q, __ldg(xp + (t + i) * a.ldx), z0, z1);
    }
  }
  // ---- own blocks: filter, accumulate, emit ----
  for (; t + hop <= wk.t_end; t += hop) {
    const bool store_rem = t + hop > rem_t0;
    if (HOP > 0) {
      double v[HOP > 0 ? HOP : 1];
#pragma unroll
      for (int i = 0; i < HOP; ++i) v[i] = __ldg(xp + (t + i) * a.ldx);
#pragma unroll
      for (int i = 0; i < HOP; ++i) {
        const double y = cascade<ARITH, 0, kMaxSec>(q, v[i], z0, z1);
        fa.acc = fma(y, y, fa.acc);
        if (store_rem && t + i >= rem_t0) rem[(t + i - rem_t0) * a.C] = y;
      }
    } else {
      for (int i = 0; i < hop; ++i) {
        const double y = cascade<ARITH, 0, kMaxSec>(q, __ldg(xp + (t + i) * a.ldx), z0, z1);
        fa.acc = fma(y, y, fa.acc);
        if (store_rem && t + i >= rem_t0) rem[(t + i - rem_t0) * a.C] = y;
      }
    }
    double sum;
    if (fa.close_block(a.nb, sum)) {
      const int64_t w = t / hop - (a.nb - 1);
      op[w * a.ldo] = feature(sum, (double)a.L, a.eps, zs, mu, sigma);
    }
  }
  // ---- ragged tail of the segment (fewer than hop samples): state / remainder only ----
  if (wk.last) {
    for (; t < wk.t_end; ++t) {
      const double y = cascade<ARITH, 0, kMaxSec>(q, __ldg(xp + t * a.ldx), z0, z1);
      if (t >= rem_t0) rem[(t - rem_t0) * a.C] = y;
    }
    if (a.state_out) {
      double* so = a.state_out + (int64_t)wk.seg * q.ns * 2 * a.C + c;
#pragma unroll
      for (int s = 0; s < kMaxSec; ++s)
        if (s < q.ns) {
          so[(int64_t)(2 * s) * a.C] = z0[s];
          so[(int64_t)(2 * s + 1) * a.C] = z1[s];
        }
    }
  }
}

}  // namespace hga
