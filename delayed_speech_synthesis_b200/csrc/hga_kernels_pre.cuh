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
// Bound: HBM.  Algorithmic bytes per input row: 8 * C_in read + 8 * C_out written.  One warp owns a row: the row is
// staged in shared memory with coalesced 256 B loads, group means are lane-strided sums + a shuffle butterfly (fixed
// order -> deterministic), outputs are written with coalesced 256 B stores.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hga {

struct ChanMapArgs {
  const double* x;
  int64_t ldx, T;
  int C_in, C_out, n_groups;
  const int32_t* sel;    // [C_out] raw column of the direct term (-1: none)
  const int32_t* grp;    // [C_out] group of the mean term (-1: none)
  const double* alpha;   // [C_out]
  const double* beta;    // [C_out]
  const int32_t* gptr;   // [n_groups + 1] CSR offsets into gidx
  const int32_t* gidx;   // raw columns of each group
  double* out;
  int64_t ldo;
};

constexpr int kChanMapWarps = 8;     // rows in flight per CTA
constexpr int kChanMapMaxGroups = 32;

__global__ void __launch_bounds__(32 * kChanMapWarps) k_chanmap(const __grid_constant__ ChanMapArgs a) {
  extern __shared__ __align__(16) double smem_cm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double* row = smem_cm + (size_t)w * (a.C_in + kChanMapMaxGroups);
  double* gmean = row + a.C_in;
  for (int64_t t = (int64_t)blockIdx.x * kChanMapWarps + w; t < a.T; t += (int64_t)gridDim.x * kChanMapWarps) {
    const double* xr = a.x + t * a.ldx;
    for (int c = lane; c < a.C_in; c += 32) row[c] = xr[c];
    __syncwarp();
    for (int g = 0; g < a.n_groups; ++g) {
      const int k0 = a.gptr[g], k1 = a.gptr[g + 1];
      double s = 0.0;
      for (int k = k0 + lane; k < k1; k += 32) s += row[a.gidx[k]];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) gmean[g] = s / (double)(k1 - k0);
    }
    __syncwarp();
    double* orow = a.out + t * a.ldo;
    for (int j = lane; j < a.C_out; j += 32) {
      const int sj = a.sel[j], gj = a.grp[j];
      double v = sj >= 0 ? a.alpha[j] * row[sj] : 0.0;
      if (gj >= 0) v = fma(a.beta[j], gmean[gj], v);
      orow[j] = v;
    }
    __syncwarp();  // the row buffer is reused by the next iteration
  }
}

}  // namespace hga
