// hga_common.cuh — parameter blocks and the biquad-cascade arithmetic shared by every HGA kernel (sm_100a).
//
// Arithmetic follows scipy.signal.sosfilt's transposed-direct-form-II recurrence as called by the reference at
// /root/reference/local/units.py:151-152:
//     y = b0*x + z0 ;  z0' = (b1*x - a1*y) + z1 ;  z1' = b2*x - a2*y ;  x <- y
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hga {

constexpr int kMaxSec = 16;  // band-pass (<= 8) + notch (<= 8) sections in one cascade

// Cascade coefficients, passed as a __grid_constant__ kernel parameter so that they are read from the constant bank.
struct Cascade {
  int ns;      // total sections
  int ns_a;    // sections of the first filter (its state block is reported separately)
  int unit;    // 1: every section has b0 == b2 and the "unit" form below is valid
  int pad;
  double b0[kMaxSec], b1[kMaxSec], b2[kMaxSec], a1[kMaxSec], a2[kMaxSec];
  // unit form: section s computes on xs = g[s] * x with numerator [1, b1u, 1]; g == 1 except where the design put
  // the filter gain (section 0 of each filter).  has_gain bit s set <=> g[s] != 1.
  double g[kMaxSec], b1u[kMaxSec];
  unsigned has_gain;
  double zi0[kMaxSec], zi1[kMaxSec];  // sosfilt_zi, untiled
};

enum Arith { kExact = 0, kFused = 1, kUnit = 2, kUnitStd = 3 };  // kUnitStd: kUnit with ns == 16, gains at sections 0 and 8

// One sample through section S.  kExact: the reference's operation order with every product and sum rounded
// separately (bit-identical to scipy's C loop compiled without FMA contraction).  kFused: same dataflow with FMAs
// (5 issue slots).  kUnit: numerator [1, b1u, 1] after an optional input gain (4 slots; 5 where the gain sits).
template <int ARITH>
__device__ __forceinline__ double biquad(const Cascade& q, int s, double x, double& z0, double& z1) {
  double y;
  if (ARITH == kExact) {
    y = __dadd_rn(__dmul_rn(q.b0[s], x), z0);
    z0 = __dadd_rn(__dsub_rn(__dmul_rn(q.b1[s], x), __dmul_rn(q.a1[s], y)), z1);
    z1 = __dsub_rn(__dmul_rn(q.b2[s], x), __dmul_rn(q.a2[s], y));
  } else if (ARITH == kFused) {
    y = fma(q.b0[s], x, z0);
    z0 = fma(-q.a1[s], y, fma(q.b1[s], x, z1));
    z1 = fma(-q.a2[s], y, q.b2[s] * x);
  } else {
    if (ARITH == kUnitStd ? (s == 0 || s == 8) : (((q.has_gain >> s) & 1u) != 0)) x *= q.g[s];
    y = x + z0;
    z0 = fma(-q.a1[s], y, fma(q.b1u[s], x, z1));
    z1 = fma(-q.a2[s], y, x);
  }
  return y;
}

// Sections [S0, S1) of the cascade, states in registers (fully unrolled: constant indices only).
template <int ARITH, int S0, int S1>
__device__ __forceinline__ double cascade(const Cascade& q, double x, double (&z0)[kMaxSec], double (&z1)[kMaxSec]) {
#pragma unroll
  for (int s = S0; s < S1; ++s) {
    if (ARITH == kUnitStd || s < q.ns) x = biquad<ARITH>(q, s, x, z0[s], z1[s]);
  }
  return x;
}

// log-power feature from a window power sum (hga_optimized.pyx:22,46) and the optional z-score (common.py:376).
__device__ __forceinline__ double feature(double sum, double n, double eps, bool zs, double mu, double sigma) {
  double f = log(__dadd_rn(__ddiv_rn(sum, n), eps));
  if (zs) f = __ddiv_rn(__dsub_rn(f, mu), sigma);
  return f;
}

}  // namespace hga
