// lane_bench.cu — stand-alone harness for the lane kernels (csrc/hga_kernels_lane.cuh): parity against a scalar CPU
// restatement of the reference's DF2T cascade + framing (float64, reference section order), and CUDA-event timing of
// the kernel shapes on K2-sized input.  Development tool; the product path is libhga_b200.so.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fopenmp -I../../delayed_speech_synthesis_b200/csrc \
//        -o lane_bench lane_bench.cu -lcuda
//   ./lane_bench [check|time|all]
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "hga_lane_host.cuh"

using namespace hga;

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      fprintf(stderr, "%s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                                \
    }                                                                                         \
  } while (0)

static LaneEncodeFn encode_fn() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr));
  return (LaneEncodeFn)p;
}

struct Coefs {
  double sos_bp[48], zi_bp[16], sos_n[48], zi_n[16];
};
static Coefs load_coefs(int fs, const char* dir) {
  Coefs c;
  const std::string path = std::string(dir) + "/lane_coefs_" + std::to_string(fs) + ".bin";
  FILE* f = fopen(path.c_str(), "rb");
  if (!f || fread(&c, sizeof c, 1, f) != 1) {
    fprintf(stderr, "cannot read %s (run tools/microbench/lane_coefs.py)\n", path.c_str());
    exit(2);
  }
  fclose(f);
  return c;
}
static void build_cascade(const Coefs& c, Cascade& q) {
  memset(&q, 0, sizeof q);
  q.ns = 16;
  q.ns_a = 8;
  q.unit = 1;
  for (int s = 0; s < 16; ++s) {
    const double* r = s < 8 ? c.sos_bp + 6 * s : c.sos_n + 6 * (s - 8);
    const double* z = s < 8 ? c.zi_bp + 2 * s : c.zi_n + 2 * (s - 8);
    q.b0[s] = r[0]; q.b1[s] = r[1]; q.b2[s] = r[2]; q.a1[s] = r[4]; q.a2[s] = r[5];
    q.zi0[s] = z[0]; q.zi1[s] = z[1];
    q.g[s] = r[0];
    q.b1u[s] = r[1] / r[0];
    if (r[0] != 1.0) q.has_gain |= 1u << s;
  }
}

// counter-based generator shared by host and device: white noise + per-channel DC offset + 60 / 120 Hz lines
__host__ __device__ inline double synth(int64_t t, int c, int fs) {
  unsigned long long z = (unsigned long long)t * 0x9E3779B97F4A7C15ull + (unsigned long long)c * 0xBF58476D1CE4E5B9ull + 0x94D049BB133111EBull;
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
  const double u = (double)(z >> 11) * (1.0 / 9007199254740992.0) - 0.5;   // uniform, variance 1/12
  const double ph = 6.283185307179586 * (double)(t % fs) / (double)fs;
  return 30.0 * 3.4641016151377544 * u + 20.0 * sin(60.0 * ph) + 5.0 * sin(120.0 * ph + 0.3) + 100.0 * (double)((c * 37) % 11 - 5);
}
__global__ void k_fill(double* x, int64_t T, int C, int64_t ldx, int fs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T * C) return;
  const int64_t t = i / C;
  const int c = (int)(i % C);
  x[t * ldx + c] = synth(t, c, fs);
}

// CPU reference: the reference's arithmetic (scipy DF2T, caller's section order, sosfilt_zi start), float64.
static void cpu_reference(const Cascade& q, int fs, int64_t T, int C, int L, int hop, const double* mu, const double* sigma, std::vector<double>& out) {
  const int64_t F = (T - L) / hop + 1;
  out.assign((size_t)F * C, 0.0);
#pragma omp parallel for schedule(dynamic, 1)
  for (int c = 0; c < C; ++c) {
    double z0[16], z1[16];
    for (int s = 0; s < 16; ++s) { z0[s] = q.zi0[s]; z1[s] = q.zi1[s]; }
    std::vector<double> bsum((size_t)(T / hop));
    double acc = 0.0;
    for (int64_t t = 0; t < (T / hop) * hop; ++t) {
      double x = synth(t, c, fs);
      for (int s = 0; s < 16; ++s) {
        const double y = q.b0[s] * x + z0[s];
        z0[s] = (q.b1[s] * x - q.a1[s] * y) + z1[s];
        z1[s] = q.b2[s] * x - q.a2[s] * y;
        x = y;
      }
      acc += x * x;
      if ((t + 1) % hop == 0) { bsum[t / hop] = acc; acc = 0.0; }
    }
    for (int64_t w = 0; w < F; ++w) {
      double s = 0.0;
      for (int j = 0; j < L / hop; ++j) s += bsum[w + j];
      double f = log(s / L + 0.01);
      if (mu) f = (f - mu[c]) / sigma[c];
      out[(size_t)w * C + c] = f;
    }
  }
}

struct DevPlan {
  LaneDev d;
  std::vector<void*> owned;
  template <class T>
  T* up(const std::vector<T>& v) {
    void* p = nullptr;
    CK(cudaMalloc(&p, std::max<size_t>(16, v.size() * sizeof(T))));
    if (!v.empty()) CK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    owned.push_back(p);
    return (T*)p;
  }
  void* raw(size_t bytes) {
    void* p = nullptr;
    CK(cudaMalloc(&p, std::max<size_t>(16, bytes)));
    owned.push_back(p);
    return p;
  }
  void release() {
    for (void* p : owned) cudaFree(p);
    owned.clear();
  }
};

static double g_piece_ops = 14000.0;   // `fix` mode sweeps it
static void upload_plan(const LaneConst& k, const LanePlan& p, int C, int hop, size_t ysize, DevPlan& dp, bool verbose, double tol = 1e-11) {
  dp.d.work = dp.up(p.work);
  dp.d.fix = dp.up(p.fix);
  dp.d.segs = dp.up(p.segs);
  dp.d.order = dp.up(p.order);
  dp.d.counter = (int*)dp.raw(4);
  const size_t CP = (size_t)p.groups * p.cw;
  dp.d.ybuf = dp.raw((size_t)p.y_rows * CP * ysize);
  dp.d.bs = (double*)dp.raw((size_t)p.bs_rows * CP * 8);
  dp.d.E = (double*)dp.raw(p.fix.size() * kLaneState * CP * 8);
  dp.d.S0 = (double*)dp.raw(p.fix.size() * kLaneState * CP * 8);
  LaneFixPlan fp;
  lane_build_fix(k, hop, fp, g_piece_ops);
  dp.d.subs = dp.up(fp.subs);
  dp.d.n_sub = (int)fp.subs.size();
  dp.d.P = dp.up(fp.P);
  std::vector<double> M(kLaneState * kLaneState);
  const double mx = lane_transition(k, p.Lc, M.data());
  dp.d.M = mx > 1e-3 * tol ? dp.up(M) : nullptr;   // the carry is only propagated across a chunk when it survives it
  if (dp.d.M) {
    std::vector<double> MT(M.size());
    for (int i = 0; i < kLaneState; ++i)
      for (int j = 0; j < kLaneState; ++j) MT[j * kLaneState + i] = M[i * kLaneState + j];
    dp.d.MT = dp.up(MT);
    dp.d.carry_window = lane_carry_window(M.data(), 1e-3 * tol);
    if (verbose) printf("  carry window %d\n", dp.d.carry_window);
  }
  if (verbose)
    printf("  plan: Lc=%lld chunks=%zu fixups=%zu head rows=%lld (%.1f MB) carry rows=%d pieces=%d |A^Lc|max=%.2e%s\n", (long long)p.Lc, p.work.size(),
           p.fix.size(), (long long)p.y_rows, (double)p.y_rows * CP * ysize / 1e6, lane_carry_rows(k), dp.d.n_sub, mx, dp.d.M ? "" : " (underflows: no propagation)");
}

// the rows the last tensor map was built over (LaneDev::x0 / ldx: the float32 offset estimate of the staged kernel)
static const double* g_x0 = nullptr;
static int64_t g_ldx = 0;
template <class P, int HOP, int TB, int NW, int NSLOT, int U>
struct OneStage {
  using Pol = P;
  static constexpr int hop = HOP, tb = TB, workers_per_sm = NW, prof_warps = NW;
  template <bool PROF>
  static int launch(const LaneConst& k, const LanePlan& p, const LaneDev& d, const CUtensorMap& tm, int C, const double* mu, const double* sigma, double* out, int n_sm, long long* prof) {
    return lane_launch_t<P, HOP, TB, NW, NSLOT, U, PROF>(k, p, d, tm, C, mu, sigma, out, C, n_sm, 0, prof);
  }
};
template <class P, int HOP, int TB, int NI, int NS, int NSLOT, int NL, int U>
struct Staged {
  using Pol = P;
  static constexpr int hop = HOP, tb = TB, workers_per_sm = NI, prof_warps = NI * NS;
  template <bool PROF>
  static int launch(const LaneConst& k, const LanePlan& p, const LaneDev& d, const CUtensorMap& tm, int C, const double* mu, const double* sigma, double* out, int n_sm, long long* prof) {
    LaneDev dx = d;
    dx.x0 = g_x0;
    dx.ldx = g_ldx;
    return lanes_launch_t<P, HOP, TB, NI, NS, NSLOT, NL, U, PROF>(k, p, dx, tm, C, mu, sigma, out, C, n_sm, 0, prof);
  }
};

template <class V>
static float run_variant(const char* name, const LaneConst& k, const LanePlan& p, const DevPlan& dp, const double* x, int64_t ldx, int64_t T, int C, const double* mu,
                         const double* sigma, double* out, int n_sm, int reps, bool prof) {
  using P = typename V::Pol;
  constexpr int HOP = V::hop, TB = V::tb, NW = V::prof_warps;
  CUtensorMap tm;
  g_x0 = x;
  g_ldx = ldx;
  if (lane_tensor_map(encode_fn(), x, ldx, C, T, 32 * P::CPL, TB, &tm) != CUDA_SUCCESS) { fprintf(stderr, "tensor map failed\n"); exit(2); }
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  long long* dprof = nullptr;
  const size_t nprof = (size_t)n_sm * NW * 3;
  if (prof) { CK(cudaMalloc(&dprof, nprof * 8)); CK(cudaMemset(dprof, 0, nprof * 8)); }
  for (int i = 0; i < 2; ++i)
    if ((prof ? V::template launch<true>(k, p, dp.d, tm, C, mu, sigma, out, n_sm, dprof)
              : V::template launch<false>(k, p, dp.d, tm, C, mu, sigma, out, n_sm, nullptr)) < 0) { fprintf(stderr, "%s: launch failed: %s\n", name, cudaGetErrorString(cudaGetLastError())); exit(2); }
  CK(cudaDeviceSynchronize());
  float best = 1e30f, sum = 0.f;
  for (int i = 0; i < reps; ++i) {
    CK(cudaEventRecord(e0));
    V::template launch<false>(k, p, dp.d, tm, C, mu, sigma, out, n_sm, nullptr);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    best = std::min(best, ms);
    sum += ms;
  }
  const double cs = (double)T * C;
  const double bytes = cs * (8.0 + 8.0 / HOP);
  printf("%-34s T=%-8lld C=%-5d  mean %.4f ms  best %.4f ms  %.1f G ch-samples/s  frac(6552.3 GB/s)=%.3f\n", name, (long long)T, C, sum / reps, best,
         cs / (sum / reps) * 1e-6, bytes / (sum / reps * 1e-3) / 6552.3e9);
  if (prof) {
    std::vector<long long> h(nprof);
    CK(cudaMemcpy(h.data(), dprof, nprof * 8, cudaMemcpyDeviceToHost));
    double busy = 0, wait = 0, tot = 0, maxtot = 0;
    int n = 0;
    for (size_t i = 0; i < nprof / 3; ++i)
      if (h[3 * i + 2] > 0) { busy += h[3 * i]; wait += h[3 * i + 1]; tot += h[3 * i + 2]; maxtot = std::max(maxtot, (double)h[3 * i + 2]); ++n; }
    if (n) printf("    per-warp cycles (mean of %d): busy %.0f  tma-wait %.0f  total %.0f  (max total %.0f)\n", n, busy / n, wait / n, tot / n, maxtot);
    if (NW == 4) {   // section-pipelined shapes (2 items x 2 stages): warp w of a CTA runs stage (w / 2 odd ? 1 - w % 2 : w % 2)
      double sb[2] = {0, 0}, sw[2] = {0, 0};
      int sn[2] = {0, 0};
      for (size_t i = 0; i < nprof / 3; ++i)
        if (h[3 * i + 2] > 0) {
          const int w = (int)(i % NW), st = ((w / 2) & 1) ? 1 - (w % 2) : (w % 2);
          sb[st] += h[3 * i];
          sw[st] += h[3 * i + 1];
          ++sn[st];
        }
      for (int st = 0; st < 2; ++st)
        if (sn[st]) printf("    stage %d: busy %.0f  waiting %.0f\n", st, sb[st] / sn[st], sw[st] / sn[st]);
    }
    cudaFree(dprof);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return sum / reps;
}

template <class V>
static double check_variant(const char* name, const Coefs& co, int fs, int64_t T, int C, int64_t force_Lc, double tol_carry, const std::vector<double>& ref,
                            const double* x, const double* mu, const double* sigma, int n_sm) {
  using P = typename V::Pol;
  constexpr int HOP = V::hop, TB = V::tb;
  Cascade q;
  build_cascade(co, q);
  LaneConst k;
  lane_build_const(q, 5 * HOP, HOP, 0.01, tol_carry, k);
  LanePlan p;
  const int64_t s0 = 0, o0 = 0;
  lane_plan(k, HOP, C, 32 * P::CPL, 1, &s0, &T, &o0, 0, (int64_t)n_sm * V::workers_per_sm, true, p, force_Lc);
  DevPlan dp;
  upload_plan(k, p, C, HOP, sizeof(typename P::TY), dp, true, tol_carry);
  const int64_t F = (T - 5 * HOP) / HOP + 1;
  double* out;
  CK(cudaMalloc(&out, (size_t)F * C * 8));
  CK(cudaMemset(out, 0xff, (size_t)F * C * 8));  // NaN pattern: every frame must be written
  CUtensorMap tm;
  g_x0 = x;
  g_ldx = C;
  lane_tensor_map(encode_fn(), x, C, C, T, 32 * P::CPL, TB, &tm);
  if (V::template launch<false>(k, p, dp.d, tm, C, mu, sigma, out, n_sm, nullptr) < 0) { fprintf(stderr, "launch failed\n"); exit(2); }
  CK(cudaDeviceSynchronize());
  std::vector<double> got((size_t)F * C);
  CK(cudaMemcpy(got.data(), out, got.size() * 8, cudaMemcpyDeviceToHost));
  double err = 0.0;
  int64_t bad = 0, worst = -1;
  for (size_t i = 0; i < got.size(); ++i) {
    const double d = fabs(got[i] - ref[i]);
    if (!(d == d)) { ++bad; continue; }
    if (d > err) { err = d; worst = (int64_t)i; }
  }
  printf("%-34s T=%-8lld C=%-5d Lc=%-7lld max|feature - cpu| = %.3e (frame %lld ch %lld)  unwritten/NaN = %lld\n", name, (long long)T, C, (long long)p.Lc, err,
         (long long)(worst / C), (long long)(worst % C), (long long)bad);
  cudaFree(out);
  dp.release();
  return bad ? 1e30 : err;
}

int main(int argc, char** argv) {
  const std::string mode = argc > 1 ? argv[1] : "all";
  const char* dir = argc > 2 ? argv[2] : "tools/microbench";
  int dev = 0, n_sm = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  {
    double tab[2 * kLogTabN];
    fill_log_table(tab);
    CK(cudaMemcpyToSymbol(g_logtab, tab, sizeof tab));
  }
  printf("device SMs %d\n", n_sm);
  int fails = 0;
  if (mode == "check" || mode == "all") {
    for (int fs : {1000, 2000}) {
      const Coefs co = load_coefs(fs, dir);
      Cascade q;
      build_cascade(co, q);
      const int hop = fs / 100, L = 5 * hop, C = 100;   // C not a multiple of 32: ragged last group
      const int64_t T = fs == 1000 ? 150000 + 7 : 120000 + 13;
      std::vector<double> mu(C), sigma(C);
      for (int c = 0; c < C; ++c) { mu[c] = 3.0 + 0.01 * c; sigma[c] = 0.4 + 0.001 * c; }
      std::vector<double> ref;
      cpu_reference(q, fs, T, C, L, hop, mu.data(), sigma.data(), ref);
      double *x, *dmu, *dsig;
      CK(cudaMalloc(&x, (size_t)T * C * 8));
      CK(cudaMalloc(&dmu, C * 8));
      CK(cudaMalloc(&dsig, C * 8));
      CK(cudaMemcpy(dmu, mu.data(), C * 8, cudaMemcpyHostToDevice));
      CK(cudaMemcpy(dsig, sigma.data(), C * 8, cudaMemcpyHostToDevice));
      k_fill<<<(unsigned)(((size_t)T * C + 255) / 256), 256>>>(x, T, C, C, fs);
      CK(cudaDeviceSynchronize());
      printf("== parity, fs %d (z-scored features; input has a per-channel DC offset of up to 500)\n", fs);
      auto expect = [&](double err, double tol) { if (!(err <= tol)) { ++fails; printf("   ^^^ FAIL (tolerance %.1e)\n", tol); } };
      if (fs == 1000) {
        for (int64_t Lc : {(int64_t)0, (int64_t)1 << 30, (int64_t)30000, (int64_t)2500, (int64_t)400}) {  // planner, one chunk, long, shorter than the carry, tiny
          expect(check_variant<OneStage<LaneF64, 10, 80, 4, 2, 5>>("f64 TB80 NW4 NS2 U5", co, fs, T, C, Lc, 1e-11, ref, x, dmu, dsig, n_sm), 5e-9);
          expect(check_variant<OneStage<LaneF32, 10, 50, 4, 2, 5>>("f32 TB50 NW4 NS2 U5", co, fs, T, C, Lc, 1e-6, ref, x, dmu, dsig, n_sm), 5e-3);
          expect(check_variant<OneStage<LaneMix, 10, 50, 4, 2, 5>>("mix TB50 NW4 NS2 U5", co, fs, T, C, Lc, 1e-7, ref, x, dmu, dsig, n_sm), 5e-3);
          expect(check_variant<Staged<LaneF32, 10, 50, 2, 2, 2, 3, 5>>("f32 2-stage x2 TB50", co, fs, T, C, Lc, 1e-6, ref, x, dmu, dsig, n_sm), 5e-3);
          // the same with looser carry tolerances: how the error at the seams grows with the dropped tail of the zero-input response
          expect(check_variant<Staged<LaneF32, 10, 50, 2, 2, 2, 3, 5>>("f32 2-stage x2 TB50 tol 1e-5", co, fs, T, C, Lc, 1e-5, ref, x, dmu, dsig, n_sm), 5e-3);
          expect(check_variant<Staged<LaneF32, 10, 50, 2, 2, 2, 3, 5>>("f32 2-stage x2 TB50 tol 1e-4", co, fs, T, C, Lc, 1e-4, ref, x, dmu, dsig, n_sm), 5e-3);
          expect(check_variant<Staged<LaneF32, 10, 40, 2, 4, 2, 2, 5>>("f32 4-stage x2 TB40", co, fs, T, C, Lc, 1e-6, ref, x, dmu, dsig, n_sm), 5e-3);
          expect(check_variant<Staged<LaneF64, 10, 50, 2, 4, 2, 2, 5>>("f64 4-stage x2 TB50", co, fs, T, C, Lc, 1e-11, ref, x, dmu, dsig, n_sm), 5e-9);
        }
        expect(check_variant<OneStage<LaneF64, 10, 40, 8, 2, 5>>("f64 TB40 NW8 NS2 U5", co, fs, T, C, 0, 1e-11, ref, x, dmu, dsig, n_sm), 5e-9);
        expect(check_variant<OneStage<LaneF32, 10, 30, 4, 3, 5>>("f32 TB30 NW4 NS3 U5", co, fs, T, C, 5000, 1e-6, ref, x, dmu, dsig, n_sm), 5e-3);
        expect(check_variant<OneStage<LaneF32, 10, 20, 8, 2, 5>>("f32 TB20 NW8 NS2 U5", co, fs, T, C, 5000, 1e-6, ref, x, dmu, dsig, n_sm), 5e-3);
      } else {
        for (int64_t Lc : {(int64_t)0, (int64_t)40000, (int64_t)3000}) {
          expect(check_variant<OneStage<LaneF64, 20, 80, 4, 2, 5>>("f64 hop20 TB80 NW4 NS2 U5", co, fs, T, C, Lc, 1e-11, ref, x, dmu, dsig, n_sm), 5e-9);
          expect(check_variant<OneStage<LaneF32, 20, 40, 4, 2, 5>>("f32 hop20 TB40 NW4 NS2 U5", co, fs, T, C, Lc, 1e-6, ref, x, dmu, dsig, n_sm), 2e-2);
          expect(check_variant<Staged<LaneF32, 20, 40, 2, 4, 2, 2, 5>>("f32 4-stage x2 hop20 TB40", co, fs, T, C, Lc, 1e-6, ref, x, dmu, dsig, n_sm), 2e-2);
          expect(check_variant<Staged<LaneF64, 20, 40, 2, 4, 2, 2, 10>>("f64 4-stage x2 hop20 TB40 U10", co, fs, T, C, Lc, 1e-11, ref, x, dmu, dsig, n_sm), 5e-9);
          expect(check_variant<OneStage<LaneMix, 20, 40, 4, 2, 10>>("mix hop20 TB40 NW4 NS2 U10", co, fs, T, C, Lc, 1e-7, ref, x, dmu, dsig, n_sm), 2e-2);
        }
      }
      cudaFree(x); cudaFree(dmu); cudaFree(dsig);
    }
  }
  if (mode == "time" || mode == "all") {
    const int fs = 1000, hop = 10, C = 128;
    const Coefs co = load_coefs(fs, dir);
    Cascade q;
    build_cascade(co, q);
    for (int64_t T : {(int64_t)3600000, (int64_t)600000, (int64_t)60000}) {
      double *x, *out;
      CK(cudaMalloc(&x, (size_t)T * C * 8));
      CK(cudaMalloc(&out, (size_t)(T / hop) * C * 8));
      k_fill<<<(unsigned)(((size_t)T * C + 255) / 256), 256>>>(x, T, C, C, fs);
      CK(cudaDeviceSynchronize());
      const int reps = T >= 3000000 ? 10 : 20;
      printf("== timing, T = %lld rows x %d channels\n", (long long)T, C);
#define TIME_X(V, TOL, NAME)                                                                         \
  {                                                                                                  \
    LaneConst k;                                                                                     \
    lane_build_const(q, 5 * hop, hop, 0.01, TOL, k);                                                 \
    LanePlan p;                                                                                      \
    const int64_t s0 = 0, o0 = 0;                                                                    \
    lane_plan(k, hop, C, 32 * V::Pol::CPL, 1, &s0, &T, &o0, 0, (int64_t)n_sm * V::workers_per_sm, true, p, 0); \
    DevPlan dp;                                                                                      \
    upload_plan(k, p, C, hop, sizeof(typename V::Pol::TY), dp, true, TOL);                           \
    run_variant<V>(NAME, k, p, dp, x, C, T, C, nullptr, nullptr, out, n_sm, reps, true);             \
    dp.release();                                                                                    \
  }
#define SINGLE_ARG(...) __VA_ARGS__
#define TIME_V(P, TOL, TB, NW, NS, U, NAME) { using V_ = OneStage<P, 10, TB, NW, NS, U>; TIME_X(V_, TOL, NAME) }
#define TIME_S(P, TOL, TB, NI, NST, NS, NL, U, NAME) { using V_ = Staged<P, 10, TB, NI, NST, NS, NL, U>; TIME_X(V_, TOL, NAME) }
      TIME_S(LaneF32, 1e-6, 40, 2, 4, 2, 2, 5, "f32 4-stage x2 TB40 NS2 NL2")
      TIME_S(LaneF32, 1e-6, 30, 2, 4, 2, 3, 5, "f32 4-stage x2 TB30 NS2 NL3")
      TIME_S(LaneF32, 1e-6, 60, 2, 2, 2, 3, 5, "f32 2-stage x2 TB60")
      TIME_S(LaneF32, 1e-6, 30, 4, 2, 2, 2, 5, "f32 2-stage x4 TB30")
      TIME_S(LaneF64, 1e-11, 50, 2, 4, 2, 2, 5, "f64 4-stage x2 TB50 NS2 NL2")
      TIME_S(LaneF64, 1e-11, 80, 2, 2, 2, 3, 5, "f64 2-stage x2 TB80")
      TIME_V(LaneF32, 1e-6, 50, 4, 2, 5, "f32 TB50 NW4 NS2 U5")
      TIME_V(LaneMix, 1e-7, 50, 4, 2, 5, "mix TB50 NW4 NS2 U5")
      TIME_V(LaneF64, 1e-11, 80, 4, 2, 5, "f64 TB80 NW4 NS2 U5")
      TIME_V(LaneF64, 1e-11, 100, 4, 2, 5, "f64 TB100 NW4 NS2 U5")
      cudaFree(x);
      cudaFree(out);
    }
  }
  if (mode == "fix") {  // the fix-up kernels alone (after one full launch has filled the scratch), per piece size
    const int fs = 1000, hop = 10, C = 128;
    const Coefs co = load_coefs(fs, dir);
    Cascade q;
    build_cascade(co, q);
    int64_t T = 3600000;
    double *x, *out;
    CK(cudaMalloc(&x, (size_t)T * C * 8));
    CK(cudaMalloc(&out, (size_t)(T / hop) * C * 8));
    k_fill<<<(unsigned)(((size_t)T * C + 255) / 256), 256>>>(x, T, C, C, fs);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int reps = 10;
    for (double po : {14000.0, 12000.0, 16000.0, 18000.0, 20000.0, 24000.0}) {
      g_piece_ops = po;
#define FIX_X(V, TOL, NAME)                                                                          \
  {                                                                                                  \
    LaneConst k;                                                                                     \
    lane_build_const(q, 5 * hop, hop, 0.01, TOL, k);                                                 \
    LanePlan p;                                                                                      \
    const int64_t s0 = 0, o0 = 0;                                                                    \
    lane_plan(k, hop, C, 32 * V::Pol::CPL, 1, &s0, &T, &o0, 0, (int64_t)n_sm * V::workers_per_sm, true, p, 0); \
    DevPlan dp;                                                                                      \
    upload_plan(k, p, C, hop, sizeof(typename V::Pol::TY), dp, false, TOL);                          \
    CUtensorMap tm;                                                                                  \
    g_x0 = x;                                                                                        \
    g_ldx = C;                                                                                       \
    lane_tensor_map(encode_fn(), x, C, C, T, 32 * V::Pol::CPL, V::tb, &tm);                          \
    V::template launch<false>(k, p, dp.d, tm, C, nullptr, nullptr, out, n_sm, nullptr);              \
    CK(cudaDeviceSynchronize());                                                                     \
    float sum = 0.f, best = 1e30f;                                                                   \
    for (int i = 0; i < 12; ++i) {                                                                   \
      CK(cudaEventRecord(e0));                                                                       \
      lane_launch_fix<typename V::Pol::TY, 10>(k, p, dp.d, C, nullptr, nullptr, out, C, 0);          \
      CK(cudaEventRecord(e1));                                                                       \
      CK(cudaEventSynchronize(e1));                                                                  \
      float ms;                                                                                      \
      CK(cudaEventElapsedTime(&ms, e0, e1));                                                         \
      if (i >= 2) { sum += ms; best = std::min(best, ms); }                                          \
    }                                                                                                \
    printf("%-14s piece_ops %-6.0f pieces %-3d fix-ups %-4zu blocks %-5zu  fix kernels mean %.4f ms best %.4f ms\n", NAME, po, dp.d.n_sub, p.fix.size(), \
           p.fix.size() * dp.d.n_sub * ((p.groups * p.cw / (int)(8 / sizeof(typename V::Pol::TY)) + 63) / 64), sum / 10, best); \
    dp.release();                                                                                    \
  }
      { using V_ = Staged<LaneF32, 10, 60, 2, 2, 2, 3, 5>; FIX_X(V_, 1e-6, "f32") }
      { using V_ = Staged<LaneF64, 10, 80, 2, 2, 2, 3, 5>; FIX_X(V_, 1e-11, "f64") }
    }
    g_piece_ops = 14000.0;
    { using V_ = Staged<LaneF32, 10, 60, 2, 2, 2, 3, 5>; TIME_X(V_, 1e-6, "f32 2-stage x2 TB60 (whole step)") }
    { using V_ = Staged<LaneF64, 10, 80, 2, 2, 2, 3, 5>; TIME_X(V_, 1e-11, "f64 2-stage x2 TB80 (whole step)") }
    { using V_ = OneStage<LaneMix, 10, 50, 4, 2, 5>; TIME_X(V_, 1e-7, "mix TB50 NW4 (whole step)") }
  }
  if (mode == "prof") {  // one variant, few launches: the target of an ncu run
    const int fs = 1000, hop = 10, C = 128;
    const Coefs co = load_coefs(fs, dir);
    Cascade q;
    build_cascade(co, q);
    int64_t T = 3600000;
    double *x, *out;
    CK(cudaMalloc(&x, (size_t)T * C * 8));
    CK(cudaMalloc(&out, (size_t)(T / hop) * C * 8));
    k_fill<<<(unsigned)(((size_t)T * C + 255) / 256), 256>>>(x, T, C, C, fs);
    CK(cudaDeviceSynchronize());
    const int reps = 2;
    const std::string which = argc > 3 ? argv[3] : "f32";
    if (which == "f32") TIME_S(LaneF32, 1e-6, 40, 2, 4, 2, 2, 5, "f32 4-stage x2 TB40 NS2 NL2")
    else if (which == "mix") TIME_V(LaneMix, 1e-7, 50, 4, 2, 5, "mix TB50 NW4 NS2 U5")
    else TIME_S(LaneF64, 1e-11, 50, 2, 4, 2, 2, 5, "f64 4-stage x2 TB50 NS2 NL2")
  }
  printf(fails ? "RESULT: %d parity failures\n" : "RESULT: parity ok\n", fails);
  return fails ? 1 : 0;
}
