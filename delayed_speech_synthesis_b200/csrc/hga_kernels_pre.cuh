// hga_kernels_pre.cuh — K-chanmap: the channel-axis transforms that sit either side of the HGA path (SURVEY.md §8f
// rows 1 and 2), as ONE row-wise kernel:
//
//     out[t, j] = alpha[j] * x[t, sel[j]]  +  beta[j] * mean_{c in group grp[j]} x[t, c]
//
//   * column gather     SelectElectrodesFromBothGrids / SelectElectrodesOverSpeechAreas   (alpha 1, beta 0)
//                       /root/reference/local/common.py:16-58  (`data[:, mapping]`)
//   * common average    CommonAverageReferencing.__call__, local/common.py:336-343: per grid, the mean over the grid's
//     referencing       non-excluded channels at each time sample is subtracted from every channel of the grid
//                       (alpha 1, beta -1, group = the grid's computation mask)
//   * bad-channel patch BadChannelCorrection.__call__, local/common.py:279-284: a bad channel's column is replaced by
//                       the mean of its non-bad neighbours (alpha 0, beta 1, group = the neighbours)
//   and any composition gather -> CAR -> gather (prepare_corpus.py:152-160, decode_online.py:69-82), which the host
//   folds into one (sel, grp, group-table) triple expressed in RAW input columns.
//
// Bound: HBM.  Algorithmic bytes per input row: 8 * C_in read + 8 * C_out written.  One warp owns two rows at a time:
// the rows are staged in shared memory with coalesced 256 B loads (both rows' loads in flight together), group means are
// lane-strided sums + a shuffle butterfly (fixed order -> deterministic) times a per-group reciprocal count, outputs are
// written with coalesced 256 B stores.  The index / weight tables live in shared memory (read once per CTA).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hga {

struct ChanMapArgs {
  const void* x;         // [T, ldx] of IN (double for the transforms; float / int32 / int16 for raw amplifier counts)
  const double* gain;    // [C_in] per-column factor applied at the load (GAIN variants), else unused
  int64_t ldx, T;
  int C_in, C_out, n_groups, n_idx;
  const int32_t* sel;    // [C_out] raw column of the direct term (-1: none)
  const int32_t* grp;    // [C_out] group of the mean term (-1: none)
  const double* alpha;   // [C_out]
  const double* beta;    // [C_out]
  const int32_t* gptr;   // [n_groups + 1] CSR offsets into gidx
  const int32_t* gidx;   // raw columns of each group
  double* out;
  int64_t ldo;
};

constexpr int kChanMapWarps = 8;     // warps per CTA
constexpr int kChanMapRows = 2;      // rows a warp works on at a time (their loads are in flight together)
constexpr int kChanMapMaxGroups = 64;

// Shared memory: the tables (read once per CTA instead of once per row), then per warp kChanMapRows x (row | group means).
__host__ __device__ inline size_t chanmap_table_bytes(int C_out, int n_groups, int n_idx) {
  size_t b = (size_t)C_out * (8 + 8) + (size_t)kChanMapMaxGroups * 8;                  // alpha, beta, 1 / count
  b += ((size_t)2 * C_out + (n_groups + 1) + n_idx) * 4;                               // sel, grp, gptr, gidx
  return (b + 15) / 16 * 16;
}
__host__ __device__ inline size_t chanmap_smem_bytes(int C_in, int C_out, int n_groups, int n_idx) {
  return chanmap_table_bytes(C_out, n_groups, n_idx) + (size_t)kChanMapWarps * kChanMapRows * (C_in + n_groups) * 8;
}

// IN / GAIN: the load stage of the raw-ingest variant, `BCI2000MatFile.signals()` = `signal * gain`
// (/root/reference/local/common.py:456-459): double(x[t, c]) * gain[c], one rounding like numpy's promoted product.
template <typename IN, bool GAIN>
__global__ void __launch_bounds__(32 * kChanMapWarps, 8) k_chanmap(const __grid_constant__ ChanMapArgs a) {
  extern __shared__ __align__(16) unsigned char smem_cm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double* s_alpha = reinterpret_cast<double*>(smem_cm);
  double* s_beta = s_alpha + a.C_out;
  double* s_ginv = s_beta + a.C_out;
  int32_t* s_sel = reinterpret_cast<int32_t*>(s_ginv + kChanMapMaxGroups);
  int32_t* s_grp = s_sel + a.C_out;
  int32_t* s_gptr = s_grp + a.C_out;
  int32_t* s_gidx = s_gptr + a.n_groups + 1;
  for (int j = threadIdx.x; j < a.C_out; j += blockDim.x) {
    s_alpha[j] = a.alpha[j]; s_beta[j] = a.beta[j]; s_sel[j] = a.sel[j]; s_grp[j] = a.grp[j];
  }
  for (int g = threadIdx.x; g <= a.n_groups; g += blockDim.x) s_gptr[g] = a.n_groups > 0 ? a.gptr[g] : 0;
  for (int g = threadIdx.x; g < a.n_groups; g += blockDim.x) s_ginv[g] = 1.0 / (double)(a.gptr[g + 1] - a.gptr[g]);
  for (int k = threadIdx.x; k < a.n_idx; k += blockDim.x) s_gidx[k] = a.gidx[k];
  __syncthreads();
  const int rstride = a.C_in + a.n_groups;   // a row of x followed by its group means
  double* rows = reinterpret_cast<double*>(smem_cm + chanmap_table_bytes(a.C_out, a.n_groups, a.n_idx)) + (size_t)w * kChanMapRows * rstride;
  const int64_t step = (int64_t)gridDim.x * kChanMapWarps * kChanMapRows;
  for (int64_t t0 = ((int64_t)blockIdx.x * kChanMapWarps + w) * kChanMapRows; t0 < a.T; t0 += step) {
    const int nr = (int)min((int64_t)kChanMapRows, a.T - t0);
#pragma unroll
    for (int r = 0; r < kChanMapRows; ++r)
      if (r < nr) {
        const IN* xr = static_cast<const IN*>(a.x) + (t0 + r) * a.ldx;
        double* row = rows + r * rstride;
        if constexpr (GAIN) {
          for (int c = lane; c < a.C_in; c += 32) row[c] = (double)xr[c] * __ldg(a.gain + c);
        } else {
          for (int c = lane; c < a.C_in; c += 32) row[c] = (double)xr[c];
        }
      }
    __syncwarp();
    // group means: the n_groups x kChanMapRows sums are reduced side by side, each by a segment of W lanes (W = the
    // largest power of two with 32 / W >= number of sums, e.g. 8 lanes for CAR over two grids): lane-strided partial
    // sums, then log2(W) shuffle rounds inside the segment -- 3 rounds for everything instead of 5 rounds per sum.
    {
      const int jobs = a.n_groups * kChanMapRows;
      int W = 32;
      while (W > 1 && 32 / W < jobs) W >>= 1;
      const int per_pass = 32 / W, seg = lane / W, sl = lane - seg * W;
      for (int base = 0; base < jobs; base += per_pass) {
        const int job = base + seg;
        const bool on = job < jobs;
        const int g = on ? job / kChanMapRows : 0, r = on ? job - g * kChanMapRows : 0;
        double sum = 0.0;
        if (on) {
          const int k1 = s_gptr[g + 1];
          const double* row = rows + r * rstride;
          for (int k = s_gptr[g] + sl; k < k1; k += W) sum += row[s_gidx[k]];
        }
        for (int o = W >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (on && sl == 0) rows[r * rstride + a.C_in + g] = sum * s_ginv[g];
      }
    }
    __syncwarp();
    for (int j = lane; j < a.C_out; j += 32) {
      const int sj = s_sel[j], gj = s_grp[j];
      const double al = s_alpha[j], be = s_beta[j];
#pragma unroll
      for (int r = 0; r < kChanMapRows; ++r)
        if (r < nr) {
          const double* row = rows + r * rstride;
          double v = sj >= 0 ? al * row[sj] : 0.0;
          if (gj >= 0) v = fma(be, row[a.C_in + gj], v);
          a.out[(t0 + r) * a.ldo + j] = v;
        }
    }
    __syncwarp();  // the row buffers are reused by the next iteration
  }
}

// K-rawscale: the identity channel map (raw ingest with no selection / re-referencing folded in), i.e. only
// `BCI2000MatFile.signals()` = `signal * gain` (/root/reference/local/common.py:456-459): out[t, c] = double(x[t, c]) * gain[c].
// A flat stream over T * C elements in groups of four: one 8 / 16 / 32 B load and one 32 B store per group, four groups
// per thread in flight.  The launch makes the total thread count a multiple of C / 4, so a thread's four columns (and its
// four gain factors, kept in registers) never change.  Same product, same single rounding as k_chanmap<IN, GAIN>.
// Bound: HBM, sizeof(IN) + 8 bytes per element.
constexpr int kRawScaleThreads = 256;
constexpr int kRawScaleUnroll = 4;
struct RawQuadF64 { double2 a, b; };
template <typename IN> struct RawQuad;
template <> struct RawQuad<int16_t> { using type = short4; };
template <> struct RawQuad<int32_t> { using type = int4; };
template <> struct RawQuad<float> { using type = float4; };
template <> struct RawQuad<double> { using type = RawQuadF64; };
__device__ __forceinline__ short4 rawquad_load(const short4* p) { return __ldg(p); }
__device__ __forceinline__ int4 rawquad_load(const int4* p) { return __ldg(p); }
__device__ __forceinline__ float4 rawquad_load(const float4* p) { return __ldg(p); }
__device__ __forceinline__ RawQuadF64 rawquad_load(const RawQuadF64* p) {
  const double2* q = reinterpret_cast<const double2*>(p);
  return RawQuadF64{__ldg(q), __ldg(q + 1)};
}
template <typename Q>
__device__ __forceinline__ void rawquad_cvt(const Q& r, double (&d)[4]) {
  d[0] = (double)r.x; d[1] = (double)r.y; d[2] = (double)r.z; d[3] = (double)r.w;
}
__device__ __forceinline__ void rawquad_cvt(const RawQuadF64& r, double (&d)[4]) {
  d[0] = r.a.x; d[1] = r.a.y; d[2] = r.b.x; d[3] = r.b.y;
}
template <typename IN, bool GAIN>
__global__ void __launch_bounds__(kRawScaleThreads, 4) k_rawscale(const IN* __restrict__ x, const double* __restrict__ gain,
                                                                  double* __restrict__ out, int64_t n_quads, int quads_per_row) {
  using Q = typename RawQuad<IN>::type;
  const Q* xq = reinterpret_cast<const Q*>(x);
  const int64_t S = (int64_t)gridDim.x * kRawScaleThreads, tid = (int64_t)blockIdx.x * kRawScaleThreads + threadIdx.x;
  double g[4] = {1.0, 1.0, 1.0, 1.0};
  if constexpr (GAIN) {
    const double2* gq = reinterpret_cast<const double2*>(gain) + 2 * (tid % quads_per_row);   // S % quads_per_row == 0
    const double2 g0 = __ldg(gq), g1 = __ldg(gq + 1);
    g[0] = g0.x; g[1] = g0.y; g[2] = g1.x; g[3] = g1.y;
  }
  // one 256-bit store per quad (st.global.v4.f64, sm_100): a warp writes 1 KB of whole lines per instruction.  Measured
  // (tools/microbench/write_rates.cu, 3.69 GB of output): 2 x 16 B stores per quad 3.4 TB/s of writes, 32 B stores 6.3 TB/s.
  auto emit = [&](int64_t i, const Q& r) {
    double d[4];
    rawquad_cvt(r, d);
    if constexpr (GAIN) {
      d[0] *= g[0]; d[1] *= g[1]; d[2] *= g[2]; d[3] *= g[3];
    }
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(out + 4 * i), "d"(d[0]), "d"(d[1]), "d"(d[2]), "d"(d[3]) : "memory");
  };
  int64_t i = tid;
  for (; i + (kRawScaleUnroll - 1) * S < n_quads; i += kRawScaleUnroll * S) {
    Q r[kRawScaleUnroll];
#pragma unroll
    for (int j = 0; j < kRawScaleUnroll; ++j) r[j] = rawquad_load(xq + i + j * S);
#pragma unroll
    for (int j = 0; j < kRawScaleUnroll; ++j) emit(i + j * S, r[j]);
  }
  for (; i < n_quads; i += S) emit(i, rawquad_load(xq + i));
}

}  // namespace hga
