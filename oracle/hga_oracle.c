/* CPU ORACLE (plain C) for the HGA front-end hot path — TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 * It restates, in scalar C with no FMA contraction (-ffp-contract=off), what the reference computes:
 *   - scipy.signal.sosfilt's transposed-direct-form-II cascade, called twice by
 *     /root/reference/local/units.py:151-152 (band-pass, then notch) with state carried (zi in, zf out);
 *   - extensions/hga/hga_optimized.pyx:8-22,27-47  (sequential sum of squares / window length, + 0.01, natural log);
 *   - local/common.py:375-376 (z-score).
 * tests/test_oracle.py asserts it bit-identical to scipy.signal.sosfilt and to the compiled reference extension.
 *
 * Parallelism: OpenMP over channels (channels never interact).  Build: see oracle/c_oracle.py.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define HGA_MAX_SECTIONS 64

int hga_oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

#define CB 8 /* channels per block: one cache line of float64, inner loops vectorise across it */

/* One cascade (n_sections biquads) over x (T, ldx) -> y (T, ldy), channels [0, C).  state: (n_sections, 2, C),
 * updated in place exactly as `y, zf = sosfilt(sos, x, axis=0, zi=state)` does.  x and y may alias.
 * Recurrence (scipy/signal/_sosfilt.pyx, float64 path):
 *     y = b0*x + z0;  z0 = (b1*x - a1*y) + z1;  z1 = b2*x - a2*y;  x = y                                        */
void hga_oracle_sosfilt(const double* x, long T, long C, long ldx, const double* sos, int n_sections, double* state,
                        double* y, long ldy, int n_threads) {
  if (n_sections > HGA_MAX_SECTIONS) return;
  const long n_blocks = (C + CB - 1) / CB;
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
#endif
  for (long blk = 0; blk < n_blocks; ++blk) {
    const long c0 = blk * CB;
    const int nc = (int)((C - c0) < CB ? (C - c0) : CB);
    double z0[HGA_MAX_SECTIONS][CB], z1[HGA_MAX_SECTIONS][CB], u[CB];
    for (int s = 0; s < n_sections; ++s)
      for (int k = 0; k < CB; ++k) {
        z0[s][k] = k < nc ? state[((long)s * 2 + 0) * C + c0 + k] : 0.0;
        z1[s][k] = k < nc ? state[((long)s * 2 + 1) * C + c0 + k] : 0.0;
      }
    for (long t = 0; t < T; ++t) {
      for (int k = 0; k < CB; ++k) u[k] = k < nc ? x[t * ldx + c0 + k] : 0.0;
      for (int s = 0; s < n_sections; ++s) {
        const double b0 = sos[6 * s], b1 = sos[6 * s + 1], b2 = sos[6 * s + 2], a1 = sos[6 * s + 4], a2 = sos[6 * s + 5];
        for (int k = 0; k < CB; ++k) {
          double v = b0 * u[k] + z0[s][k];
          z0[s][k] = (b1 * u[k] - a1 * v) + z1[s][k];
          z1[s][k] = b2 * u[k] - a2 * v;
          u[k] = v;
        }
      }
      for (int k = 0; k < nc; ++k) y[t * ldy + c0 + k] = u[k];
    }
    for (int s = 0; s < n_sections; ++s)
      for (int k = 0; k < nc; ++k) {
        state[((long)s * 2 + 0) * C + c0 + k] = z0[s][k];
        state[((long)s * 2 + 1) * C + c0 + k] = z1[s][k];
      }
  }
}

/* Windowed log power: out[w, c] = log( sum_{r in [start[w], stop[w])} data[r, c]^2 / (stop-start) + eps ).
 * Sequential accumulation order, as hga_optimized.pyx:19-22.  mu/sigma optional (NULL): z-score (f - mu)/sigma.   */
void hga_oracle_logpower(const double* data, long C, long ld, const long* start, const long* stop, long F, double eps,
                         const double* mu, const double* sigma, double* out, int n_threads) {
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(n_threads)
#endif
  for (long w = 0; w < F; ++w) {
    double* o = out + w * C;
    for (long c = 0; c < C; ++c) o[c] = 0.0;
    for (long r = start[w]; r < stop[w]; ++r) {
      const double* d = data + r * ld;
      for (long c = 0; c < C; ++c) o[c] += d[c] * d[c];
    }
    const double n = (double)(stop[w] - start[w]);
    for (long c = 0; c < C; ++c) {
      double f = log(o[c] / n + eps);
      if (mu) f = (f - mu[c]) / sigma[c];
      o[c] = f;
    }
  }
}

/* Fused single-stream offline extraction for the CPU baseline: per channel block, both cascades + framing on a
 * regular (L, hop) grid + optional z-score, without materialising the filtered signal (keeps an L-row ring of
 * squares per block).  Identical arithmetic and accumulation order to sosfilt x2 -> compute_log_power_features ->
 * z-score for a fresh extractor whose single call has T >= L (frame-buffer case 1, hga_optimized.pyx:104-107).
 * x: (T, ldx); sos_a/sos_b: (na,6)/(nb,6); zi_a/zi_b: (na,2)/(nb,2) untiled; out: (F, C), F = (T-L)/hop + 1.       */
long hga_oracle_extract_fused(const double* x, long T, long C, long ldx, const double* sos_a, int na,
                              const double* zi_a, const double* sos_b, int nb, const double* zi_b, long L, long hop,
                              double eps, const double* mu, const double* sigma, double* out, int n_threads) {
  if (T < L || na + nb > HGA_MAX_SECTIONS || L > 1024) return 0;
  const long F = (T - L) / hop + 1;
  const long n_blocks = (C + CB - 1) / CB;
  const int ns = na + nb;
  double q[HGA_MAX_SECTIONS][6], zi[HGA_MAX_SECTIONS][2];
  for (int s = 0; s < na; ++s) { memcpy(q[s], sos_a + 6 * s, 48); zi[s][0] = zi_a[2 * s]; zi[s][1] = zi_a[2 * s + 1]; }
  for (int s = 0; s < nb; ++s) { memcpy(q[na + s], sos_b + 6 * s, 48); zi[na + s][0] = zi_b[2 * s]; zi[na + s][1] = zi_b[2 * s + 1]; }
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
#endif
  for (long blk = 0; blk < n_blocks; ++blk) {
    const long c0 = blk * CB;
    const int nc = (int)((C - c0) < CB ? (C - c0) : CB);
    double z0[HGA_MAX_SECTIONS][CB], z1[HGA_MAX_SECTIONS][CB], u[CB], acc[CB];
    double (*ring)[CB] = malloc(sizeof(double[CB]) * (size_t)L);
    for (int s = 0; s < ns; ++s)
      for (int k = 0; k < CB; ++k) { z0[s][k] = zi[s][0]; z1[s][k] = zi[s][1]; }
    long w = 0, slot = 0, next_end = L - 1;
    for (long t = 0; t < T; ++t) {
      for (int k = 0; k < CB; ++k) u[k] = k < nc ? x[t * ldx + c0 + k] : 0.0;
      for (int s = 0; s < ns; ++s) {
        const double b0 = q[s][0], b1 = q[s][1], b2 = q[s][2], a1 = q[s][4], a2 = q[s][5];
        for (int k = 0; k < CB; ++k) {
          double v = b0 * u[k] + z0[s][k];
          z0[s][k] = (b1 * u[k] - a1 * v) + z1[s][k];
          z1[s][k] = b2 * u[k] - a2 * v;
          u[k] = v;
        }
      }
      for (int k = 0; k < CB; ++k) ring[slot][k] = u[k] * u[k];
      slot = slot + 1 == L ? 0 : slot + 1;              /* slot now points at the oldest row = first row of window */
      if (t == next_end && w < F) {                     /* window w = rows [w*hop, w*hop+L) is complete */
        for (int k = 0; k < CB; ++k) acc[k] = 0.0;
        long r = slot;
        for (long i = 0; i < L; ++i) {
          for (int k = 0; k < CB; ++k) acc[k] += ring[r][k];
          r = r + 1 == L ? 0 : r + 1;
        }
        for (int k = 0; k < nc; ++k) {
          double f = log(acc[k] / (double)L + eps);
          if (mu) f = (f - mu[c0 + k]) / sigma[c0 + k];
          out[w * C + c0 + k] = f;
        }
        ++w;
        next_end += hop;
      }
    }
    free(ring);
  }
  return F;
}
