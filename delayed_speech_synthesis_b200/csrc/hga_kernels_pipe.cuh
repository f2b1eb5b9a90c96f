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

#include "hga_kernels_fused.cuh"

namespace hga {

struct PipeConst {  // compile-time-indexable, warp-uniform operands (constant bank)
  double zu0[kMaxSec], zu1[kMaxSec];  // sosfilt_zi divided by the gains in front of each section
  double pow_scale;                   // G^2 / L
  float a1f[kMaxSec], a2f[kMaxSec], b1uf[kMaxSec];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA: one (32 channels x TB rows) f64 box of x -> shared memory, completion on an mbarrier.  Out-of-range columns
// (C not a multiple of 32) and rows past the end of x are zero-filled by the hardware.
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int row, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(c0), "r"(row), "r"(smem_u32(bar))
      : "memory");
}

template <typename T> struct PipeCoef;
template <> struct PipeCoef<double> {
  static __device__ __forceinline__ double a1(const Cascade& q, const PipeConst&, int s) { return q.a1[s]; }
  static __device__ __forceinline__ double a2(const Cascade& q, const PipeConst&, int s) { return q.a2[s]; }
  static __device__ __forceinline__ double b1(const Cascade& q, const PipeConst&, int s) { return q.b1u[s]; }
};
template <> struct PipeCoef<float> {
  static __device__ __forceinline__ float a1(const Cascade&, const PipeConst& p, int s) { return p.a1f[s]; }
  static __device__ __forceinline__ float a2(const Cascade&, const PipeConst& p, int s) { return p.a2f[s]; }
  static __device__ __forceinline__ float b1(const Cascade&, const PipeConst& p, int s) { return p.b1uf[s]; }
};

// `rows` rows of a tile through sections [G*NS, (G+1)*NS): in (row-major, 32 lanes per row) -> out.
template <typename T, typename TIN, int G, int NS, int RB>
__device__ __forceinline__ void pipe_rows(const Cascade& q, const PipeConst& pc, const TIN* __restrict__ in, T* __restrict__ out,
                                          int rows, T (&z0)[NS], T (&z1)[NS]) {
  using K = PipeCoef<T>;
  int r = 0;
  for (; r + RB <= rows; r += RB) {
    T v[RB];
#pragma unroll
    for (int i = 0; i < RB; ++i) v[i] = (T)in[(r + i) * 32];
#pragma unroll
    for (int i = 0; i < RB; ++i) {
      T x = v[i];
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const T y = x + z0[s];
        z0[s] = fma(-K::a1(q, pc, G * NS + s), y, fma(K::b1(q, pc, G * NS + s), x, z1[s]));
        z1[s] = fma(-K::a2(q, pc, G * NS + s), y, x);
        x = y;
      }
      v[i] = x;
    }
#pragma unroll
    for (int i = 0; i < RB; ++i) out[(r + i) * 32] = v[i];
  }
  for (; r < rows; ++r) {
    T x = (T)in[r * 32];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const T y = x + z0[s];
      z0[s] = fma(-K::a1(q, pc, G * NS + s), y, fma(K::b1(q, pc, G * NS + s), x, z1[s]));
      z1[s] = fma(-K::a2(q, pc, G * NS + s), y, x);
      x = y;
    }
    out[r * 32] = x;
  }
}

template <int S, int TB, typename T>
struct PipeSmem {
  static constexpr int NX = TB >= 40 ? 2 : 3;                  // TMA ring depth
  static constexpr int kBsRing = 16;                           // hop-block sums kept (>= nb - 1 + 2 * TB / hop)
  static constexpr size_t x_bytes = (size_t)NX * TB * 32 * 8;  // f64 input tiles
  static constexpr size_t link_bytes = (size_t)S * 2 * TB * 32 * sizeof(T);
  static constexpr size_t bs_bytes = (size_t)kBsRing * 32 * 8;
  static constexpr size_t total = 128 + x_bytes + link_bytes + bs_bytes + 64;
};

// One CTA = one (work item, 32-channel group); S warps = S section groups.
template <typename T, int S, int TB>
__global__ void __launch_bounds__(32 * S) k_hga_pipe(const __grid_constant__ Cascade q, const __grid_constant__ PipeConst pc,
                                                      const __grid_constant__ FusedArgs a, const __grid_constant__ CUtensorMap tmap) {
  constexpr int NS = kMaxSec / S;
  constexpr int RB = 10;
  using SM = PipeSmem<S, TB, T>;
  constexpr int NX = SM::NX;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* base = (unsigned char*)(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
  double* xring = (double*)base;                                    // [NX][TB][32]
  T* links = (T*)(base + SM::x_bytes);                              // [S][2][TB][32]; link S-1 is the y tile
  double* bs = (double*)(base + SM::x_bytes + SM::link_bytes);      // [kBsRing][32]
  uint64_t* full = (uint64_t*)(base + SM::x_bytes + SM::link_bytes + SM::bs_bytes);  // [NX]

  const int cg = (a.C + 31) / 32;
  const int item = blockIdx.x / cg, c0 = (blockIdx.x % cg) * 32;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const Work wk = a.work[item];
  const int hop = a.hop, nb = a.nb;
  const int64_t span = wk.t_end - wk.t_begin;
  const int ntiles = (int)((span + TB - 1) / TB);
  const bool fresh = wk.t_begin == 0;
  const int c = c0 + lane;
  const bool zs = a.mu != nullptr;
  const double mu = (zs && c < a.C) ? a.mu[c] : 0.0;
  const double inv_sigma = (zs && c < a.C) ? 1.0 / a.sigma[c] : 1.0;

  T z0[NS], z1[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    // section index g*NS+s is warp-uniform but not compile-time: read the initial state through a small switch-free
    // constant-bank lookup (once per kernel)
    z0[s] = fresh ? (T)pc.zu0[g * NS + s] : (T)0;
    z1[s] = fresh ? (T)pc.zu1[g * NS + s] : (T)0;
  }

  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int i = 0; i < NX - 1 && i < ntiles; ++i) {
      mbar_expect_tx(&full[i], TB * 32 * 8);
      tma_load_2d(xring + (size_t)i * TB * 32, &tmap, c0, (int)(wk.row0 + wk.t_begin + (int64_t)i * TB - a.x_row_lo), &full[i]);
    }
  }
  __syncthreads();

  const int nbt = TB / hop;  // hop blocks per tile
  const int64_t blk_begin = wk.t_begin / hop;
  double* op = a.out + wk.out_row0 * a.ldo + c;

  for (int n = 0; n < ntiles + S + 1; ++n) {
    const int tile = n - g;
    if (tile >= 0 && tile < ntiles) {
      const int rows = (int)min((int64_t)TB, span - (int64_t)tile * TB);
      T* out = links + ((size_t)(g * 2 + (tile & 1)) * TB) * 32 + lane;
      if (g == 0) {
        const int pre = tile + NX - 1;
        if (lane == 0 && pre < ntiles) {
          uint64_t* b = &full[pre % NX];
          mbar_expect_tx(b, TB * 32 * 8);
          tma_load_2d(xring + (size_t)(pre % NX) * TB * 32, &tmap, c0, (int)(wk.row0 + wk.t_begin + (int64_t)pre * TB - a.x_row_lo), b);
        }
        mbar_wait(&full[tile % NX], (tile / NX) & 1);
        pipe_rows<T, double, 0, NS, RB>(q, pc, xring + (size_t)(tile % NX) * TB * 32 + lane, out, rows, z0, z1);
      } else {
        const T* in = links + ((size_t)((g - 1) * 2 + (tile & 1)) * TB) * 32 + lane;
        if (S >= 2 && g == 1) pipe_rows<T, T, (S >= 2 ? 1 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 4 && g == 2) pipe_rows<T, T, (S >= 4 ? 2 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 4 && g == 3) pipe_rows<T, T, (S >= 4 ? 3 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 8 && g == 4) pipe_rows<T, T, (S >= 8 ? 4 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 8 && g == 5) pipe_rows<T, T, (S >= 8 ? 5 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 8 && g == 6) pipe_rows<T, T, (S >= 8 ? 6 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
        if (S >= 8 && g == 7) pipe_rows<T, T, (S >= 8 ? 7 : 0), NS, RB>(q, pc, in, out, rows, z0, z1);
      }
    }
    // ---- framing, all warps: phase A = hop-block power sums of y tile mA (written by warp S-1 in step n-1);
    //      phase B = frames closed by the blocks of tile mA-1 (sums written in step n-1).
    const int mA = n - S, mB = n - S - 1;
    if (mA >= 0 && mA < ntiles) {
      const T* yt = links + ((size_t)((S - 1) * 2 + (mA & 1)) * TB) * 32 + lane;
      const int64_t t0 = wk.t_begin + (int64_t)mA * TB;
      for (int j = g; j < nbt; j += S) {
        const int64_t tb = t0 + (int64_t)j * hop;
        if (tb >= wk.t_acc && tb + hop <= wk.t_end) {
          double acc = 0.0;
          for (int i = 0; i < hop; ++i) {
            const double v = (double)yt[(j * hop + i) * 32];
            acc = fma(v, v, acc);
          }
          bs[((blk_begin + (int64_t)mA * nbt + j) & (SM::kBsRing - 1)) * 32 + lane] = acc;
        }
      }
    }
    if (mB >= 0 && mB < ntiles && c < a.C) {
      const int64_t t0 = wk.t_begin + (int64_t)mB * TB;
      for (int j = g; j < nbt; j += S) {
        const int64_t tb = t0 + (int64_t)j * hop;
        if (tb >= wk.t_acc + (int64_t)(nb - 1) * hop && tb + hop <= wk.t_end) {
          const int64_t blk = blk_begin + (int64_t)mB * nbt + j;
          double sum = 0.0;
          for (int k = nb - 1; k >= 0; --k) sum += bs[((blk - k) & (SM::kBsRing - 1)) * 32 + lane];
          double f = log(fma(sum, pc.pow_scale, a.eps));
          if (zs) f = (f - mu) * inv_sigma;
          op[(blk - (nb - 1)) * a.ldo] = f;
        }
      }
    }
    asm volatile("bar.sync 0;" ::: "memory");
  }
}

}  // namespace hga
