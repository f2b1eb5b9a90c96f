// hga_lane_host.cuh — host side of the lane kernels (hga_kernels_lane.cuh): kernel-basis operands and fresh state, the
// fix-up stage plan, the chunk transition matrix, the time-chunk planner and the launch sequence.  Used by hga_api.cu and
// by tools/microbench/lane_bench.cu.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <functional>
#include <queue>
#include <vector>

#include "hga_kernels_lane.cuh"

namespace hga {

// The lane kernels need the cascade scipy designs for the path: 8 + 8 sections, b0 == b2 in every section, gains only in
// section 0 of each filter, band-pass numerators 4 x [1, 2, 1] and 4 x [1, -2, 1].
inline bool lane_cascade_ok(const Cascade& q) {
  if (!(q.unit && q.ns == 16 && q.ns_a == 8 && (q.has_gain & ~0x101u) == 0)) return false;
  int plus = 0, minus = 0;
  for (int s = 0; s < 8; ++s) {
    if (std::fabs(q.b1u[s] - 2.0) <= 1e-13) ++plus;
    else if (std::fabs(q.b1u[s] + 2.0) <= 1e-13) ++minus;
    else return false;
  }
  return plus == 4 && minus == 4;
}

// tol: relative size below which a section's zero-input response is dropped from the carry (per precision mode).
inline void lane_build_const(const Cascade& q, int L, int hop, double eps, double tol, LaneConst& k) {
  memset(&k, 0, sizeof k);
  double G = 1.0;
  for (int s = 0; s < q.ns; ++s) G *= q.g[s];
  for (int s = 0; s < 8; ++s) {  // A sections: denominators in the caller's order, numerators re-paired to [1, 0, -1]
    k.c1d[s] = q.a1[s];
    k.c2d[s] = q.a2[s];
    k.cbd[s] = 0.0;
  }
  double g = q.g[0];
  for (int s = 8; s < 16; ++s) {  // U sections
    k.c1d[s] = -q.a1[s];
    k.c2d[s] = -q.a2[s];
    k.cbd[s] = q.b1u[s];
    g *= q.g[s];
    k.fresh0[s] = q.zi0[s] / g;  // sosfilt_zi of the notch, divided by the gains in front of the section
    k.fresh1[s] = q.zi1[s] / g;
  }
  // band-pass: sosfilt_zi is the steady state of the history x == 1; in the re-paired form only section 0 holds it
  // (v = w = 1: y = x - v = 0, v' = a1 * 0 + w, w' = a2 * 0 + x), everything behind a zero at DC is at rest.  The
  // reference's zi carries the same history whenever zi_bp is sosfilt_zi of the band-pass (checked by the caller).
  k.fresh0[0] = 1.0;
  k.fresh1[0] = 1.0;
  auto pair = [](float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return ((unsigned long long)u << 32) | u;
  };
  for (int s = 0; s < 16; ++s) {
    k.c1f[s] = (float)k.c1d[s];
    k.c2f[s] = (float)k.c2d[s];
    k.cbf[s] = (float)k.cbd[s];
    k.c1p[s] = pair(k.c1f[s]);
    k.c2p[s] = pair(k.c2f[s]);
    k.cbp[s] = pair(k.cbf[s]);
  }
  k.pow_scale = G * G / (double)L;
  k.eps = eps;
  // fix-up stages: sections [0, s) are dropped once every one of them has decayed below tol (x margin: the band-pass states
  // scale with out-of-band content such as DC offsets, 1e-6; the notch states with the in-band signal and resonate into
  // one another, 1e-2)
  double rows[16], cum = 0.0;
  for (int s = 0; s < 16; ++s) {
    const double a2 = std::fabs(s < 8 ? k.c2d[s] : -k.c2d[s]);
    const double a1 = std::fabs(k.c1d[s]);
    const double disc = a1 * a1 - 4.0 * a2;
    const double r = disc < 0 ? std::sqrt(a2) : (a1 + std::sqrt(disc)) / 2.0;
    const double t = tol * (s < 8 ? 1e-6 : 1e-2);
    rows[s] = r > 0 && r < 1 ? std::ceil(std::log(t) / std::log(r)) : 1.0;
  }
  const int bound[kLaneStages] = {7, 9, 11, 13, 15};
  for (int i = 0, s = 0; i < kLaneStages; ++i) {
    for (; s <= bound[i]; ++s) cum = std::max(cum, rows[s]);
    k.stage_rows[i] = (int)((std::min(cum, 1e8) + hop - 1) / hop) * hop;
  }
}
inline int lane_carry_rows(const LaneConst& k) { return k.stage_rows[kLaneStages - 1]; }

// One zero-input step of the kernel's cascade on state s = (z0[0..15], z1[0..15]), in long double.
inline void lane_step_zero_input(const LaneConst& k, long double* s) {
  long double u = 0.0L;
  for (int i = 0; i < 16; ++i) {
    long double &z0 = s[i], &z1 = s[16 + i];
    if (i < 8) {
      const long double y = u - z0;
      z0 = (long double)k.c1d[i] * y + z1;
      z1 = (long double)k.c2d[i] * y + u;
      u = y;
    } else {
      const long double y = u + z0;
      z0 = (long double)k.c1d[i] * y + ((long double)k.cbd[i] * u + z1);
      z1 = (long double)k.c2d[i] * y + u;
      u = y;
    }
  }
}
// M = A^n (row-major 32 x 32) and its largest absolute entry.
inline double lane_transition(const LaneConst& k, int64_t n, double* M) {
  constexpr int N = kLaneState;
  std::vector<long double> A(N * N), R(N * N, 0.0L), T(N * N);
  for (int c = 0; c < N; ++c) {
    long double s[N] = {};
    s[c] = 1.0L;
    lane_step_zero_input(k, s);
    for (int r = 0; r < N; ++r) A[r * N + c] = s[r];
  }
  for (int i = 0; i < N; ++i) R[i * N + i] = 1.0L;
  auto mul = [&](std::vector<long double>& X, const std::vector<long double>& Y) {  // X = X * Y
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) {
        long double acc = 0.0L;
        for (int l = 0; l < N; ++l) acc += X[i * N + l] * Y[l * N + j];
        T[i * N + j] = acc;
      }
    X = T;
  };
  while (n > 0) {
    if (n & 1) mul(R, A);
    std::vector<long double> A2 = A;
    mul(A2, A);
    A = A2;
    n >>= 1;
  }
  double mx = 0.0;
  for (int i = 0; i < N * N; ++i) {
    M[i] = (double)R[i];
    mx = std::max(mx, std::fabs(M[i]));
  }
  return mx;
}

// End states a chunk's start state has to sum: the smallest w with max |M^w| <= thresh, M = A^Lc (the first dropped term of
// s0(f) = sum_p M^p e(f - p)); 0 when the powers have not decayed after `cap` products (then the sequential chain is used).
constexpr int kLaneCarryWindowMax = 8;   // measured: longer windows lose to the sequential chain (K1: 34 -> 25 G channel-samples/s at 48)
inline int lane_carry_window(const double* M, double thresh, int cap = kLaneCarryWindowMax) {
  constexpr int N = kLaneState;
  std::vector<double> Pw(M, M + N * N), T(N * N);
  for (int w = 1; w <= cap; ++w) {
    double mx = 0.0;
    for (int i = 0; i < N * N; ++i) mx = std::max(mx, std::fabs(Pw[i]));
    if (mx <= thresh) return w;
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) {
        double acc = 0.0;
        for (int l = 0; l < N; ++l) acc += Pw[i * N + l] * M[l * N + j];
        T[i * N + j] = acc;
      }
    Pw.swap(T);
  }
  return 0;
}

// Pieces of the head fix-up (k_lane_fix) and their start-state operators: the staged zero-input evolution of the 32 unit
// states, sampled where a piece starts.  Pieces are cut to roughly equal operation counts so that they finish together;
// every piece but the first starts kLaneNB - 1 hop blocks before the first frame it finishes.
struct LaneFixPlan {
  std::vector<LaneSub> subs;
  std::vector<double> P;
};
inline void lane_build_fix(const LaneConst& k, int hop, LaneFixPlan& fp, double piece_ops = 14000.0) {
  constexpr int N = kLaneState;
  const int first[kLaneStages] = {0, 8, 10, 12, 14};
  const int K = lane_carry_rows(k);
  const int head_blocks = K / hop + (kLaneNB - 1);
  auto stage_of = [&](int row) {
    int st = 0;
    while (st + 1 < kLaneStages && row >= k.stage_rows[st]) ++st;
    return st;
  };
  fp.subs.clear();
  fp.P.clear();
  std::vector<char> start((size_t)K + 1, 0);
  for (int b0 = 0; b0 < head_blocks;) {
    const int row0 = std::max(0, b0 - (kLaneNB - 1)) * hop;
    const int f = first[stage_of(std::min(row0, K - 1))];
    int ops = 1;   // the square and the add of every row
    for (int s = f; s < 16; ++s) ops += s < 8 ? 3 : 4;
    int nb = std::max(2 * kLaneNB, (int)std::ceil(piece_ops / ((double)ops * hop)));
    if (head_blocks - (b0 + nb) < kLaneNB) nb = head_blocks - b0;   // no sliver at the end
    fp.subs.push_back(LaneSub{row0, b0, b0 + nb, f, 0, 0});
    if (row0 < K) start[row0] = 1;
    b0 += nb;
  }
  // evolve the 32 unit states (columns of S) and sample
  std::vector<double> S(N * N, 0.0);
  for (int i = 0; i < N; ++i) S[i * N + i] = 1.0;
  for (int r = 0; r < K; ++r) {
    const int f = first[stage_of(r)];
    if (start[r] && r > 0) {
      for (LaneSub& sub : fp.subs)
        if (sub.row0 == r) {
          sub.p_off = (int32_t)fp.P.size();
          for (int s = f; s < 16; ++s)
            for (int j = 0; j < N; ++j) fp.P.push_back(S[s * N + j]);
          for (int s = f; s < 16; ++s)
            for (int j = 0; j < N; ++j) fp.P.push_back(S[(16 + s) * N + j]);
        }
    }
    for (int j = 0; j < N; ++j) {  // one zero-input step of sections [f, 16) on column j; the dropped sections are at rest
      double u = 0.0;
      for (int s2 = 0; s2 < f; ++s2) S[s2 * N + j] = S[(16 + s2) * N + j] = 0.0;
      for (int s2 = f; s2 < 16; ++s2) {
        double &z0 = S[s2 * N + j], &z1 = S[(16 + s2) * N + j];
        if (s2 < 8) {
          const double y = u - z0;
          z0 = k.c1d[s2] * y + z1;
          z1 = k.c2d[s2] * y + u;
          u = y;
        } else {
          const double y = u + z0;
          z0 = k.c1d[s2] * y + (k.cbd[s2] * u + z1);
          z1 = k.c2d[s2] * y + u;
          u = y;
        }
      }
    }
  }
  if (fp.P.empty()) fp.P.push_back(0.0);
}

struct LanePlan {
  std::vector<LaneWork> work;
  std::vector<LaneFix> fix;
  std::vector<LaneCarrySeg> segs;
  std::vector<int> order;   // item order, longest chunk first
  int64_t y_rows = 0, bs_rows = 0, Lc = 0;
  int groups = 0, cw = 32, max_fix_frames = 0;  // groups of cw channels (one warp each)
};

// Makespan of items of `len` rows (+ a fixed cost) over `workers` greedy workers, longest first.
inline double lane_makespan(std::vector<int64_t> lens, int64_t workers, double fixed) {
  std::sort(lens.begin(), lens.end(), std::greater<int64_t>());
  if ((int64_t)lens.size() <= workers) return lens.empty() ? 0.0 : (double)lens[0] + fixed;
  std::priority_queue<double, std::vector<double>, std::greater<double>> q;
  for (int64_t i = 0; i < workers; ++i) q.push(0.0);
  double end = 0.0;
  for (int64_t v : lens) {
    double t = q.top();
    q.pop();
    t += (double)v + fixed;
    end = std::max(end, t);
    q.push(t);
  }
  return end;
}

// Cut the segments into time-chunks.  workers = warps the device runs at a time; x_row_lo = first addressable x row (the
// TMA tensor starts there).  Cost model: main pass = greedy makespan of (rows + 64) per (chunk, group) item; each chunk
// with a predecessor adds a fix-up of `fix_cost` row-equivalents, spread over the device.
inline void lane_plan(const LaneConst& k, int hop, int C, int cw, int n_seg, const int64_t* seg_start, const int64_t* seg_len, const int64_t* out_row,
                      int64_t x_row_lo, int64_t workers, bool allow_chunking, LanePlan& p, int64_t force_Lc = 0) {
  const int K = lane_carry_rows(k);
  const int groups = (C + cw - 1) / cw;
  int64_t maxT = 0;
  for (int i = 0; i < n_seg; ++i) maxT = std::max(maxT, seg_len[i]);
  const int64_t whole = (maxT + hop - 1) / hop * hop;
  int64_t best_Lc = whole;
  if (force_Lc > 0) best_Lc = (force_Lc + hop - 1) / hop * hop;
  else if (allow_chunking && workers > 0) {
    // fix-up cost of one (chunk, group) in row-equivalents of the main pass: operations per stage / 56, times a latency factor
    const int first[kLaneStages] = {0, 8, 10, 12, 14};
    double fix_ops = 0.0;
    for (int i = 0, r0 = 0; i < kLaneStages; ++i) {
      int ops = 0;
      for (int s = first[i]; s < 16; ++s) ops += s < 8 ? 3 : 4;
      fix_ops += (double)(k.stage_rows[i] - r0) * ops;
      r0 = k.stage_rows[i];
    }
    const double fix_cost = 2.0 * fix_ops / 56.0;   // measured: k_lane_fix costs twice its operation count (profiles/r02_ncu_fix_f32.csv)
    double best = 1e300;
    const int64_t min_Lc = std::max<int64_t>(16 * hop, 320);
    // candidate chunk counts: every n up to 64, then a geometric ladder, plus the counts that fill the workers exactly
    // (n * groups == k * workers) — the ladder alone steps over them (141 -> 149 chunks around the 148 that fill 296 workers)
    std::vector<int64_t> cand;
    for (int64_t n = 1; n <= (1 << 20); n = n < 64 ? n + 1 : n + n / 16) cand.push_back(n);
    for (int64_t kk = 1; kk <= 32; ++kk) {
      const int64_t n = workers * kk / groups;
      if (n >= 1) cand.push_back(n);
    }
    std::sort(cand.begin(), cand.end());
    cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
    for (int64_t n : cand) {
      const int64_t Lc = ((maxT + n - 1) / n + hop - 1) / hop * hop;
      if (n > 1 && Lc < min_Lc) break;
      std::vector<int64_t> lens;
      int64_t n_fix = 0;
      for (int i = 0; i < n_seg; ++i) {
        const int64_t nc = std::max<int64_t>(1, (seg_len[i] + Lc - 1) / Lc);
        for (int64_t j = 0; j < nc; ++j) {
          const int64_t len = std::min(Lc, seg_len[i] - j * Lc);
          for (int g = 0; g < groups; ++g) lens.push_back(len);
        }
        n_fix += nc - 1;
      }
      const double frac = std::min(1.0, (double)Lc / std::max(1, K));  // short chunks: the fix-up covers only Lc rows
      const double cost = lane_makespan(lens, workers, 64.0) + (double)n_fix * groups * fix_cost * frac / (double)workers;
      if (cost < best * 0.999) { best = cost; best_Lc = Lc; }
      if (lens.size() > (size_t)workers * 16) break;
    }
  }
  p = LanePlan{};
  p.Lc = best_Lc;
  p.groups = groups;
  p.cw = cw;
  const int Kd = (K + hop - 1) / hop * hop;
  for (int i = 0; i < n_seg; ++i) {
    const int64_t T = seg_len[i];
    const int64_t nc = std::max<int64_t>(1, (T + best_Lc - 1) / best_Lc);
    LaneCarrySeg cs{(int32_t)p.fix.size(), (int32_t)(nc - 1)};
    for (int64_t j = 0; j < nc; ++j) {
      const int64_t t0 = j * best_Lc, rows = std::min(best_Lc, T - t0);
      const int nblk = (int)(rows / hop);
      LaneWork w{};
      w.out_row0 = out_row[i];
      w.x_row = (int32_t)(seg_start[i] + t0 - x_row_lo);
      w.rows = (int32_t)rows;
      w.blk0 = (int32_t)(t0 / hop);
      w.fresh = j == 0;
      w.seg_row = (int32_t)(seg_start[i] - x_row_lo);
      w.off_rows = T >= 3 ? 3 : 1;
      w.slot = j + 1 < nc ? (int32_t)p.fix.size() + (j > 0 ? 1 : 0) : -1;  // fix-up index of the successor
      w.tail_bs_off = -1;
      w.head_off = 0;
      w.head_rows = 0;
      w.first_clean = 0;
      if (j > 0) {
        LaneFix f{};
        f.out_row0 = out_row[i];
        f.blk0 = w.blk0;
        f.nblk = nblk;
        f.head_rows = std::min(nblk * hop, Kd + (kLaneNB - 1) * hop);
        f.fix_rows = std::min(nblk * hop, Kd);
        const bool all_head = f.head_rows == nblk * hop;
        f.n_frames = all_head ? nblk : Kd / hop + (kLaneNB - 1);
        f.next_fix = (all_head && j + 1 < nc) ? (int32_t)p.fix.size() + 1 : -1;
        f.first = j == 1;
        f.pos = (int32_t)(j - 1);
        f.head_off = p.y_rows;
        f.bs_off = p.bs_rows;
        p.y_rows += f.head_rows;
        p.bs_rows += 2 * (kLaneNB - 1);   // the predecessor's last four block sums and this head's first four
        p.max_fix_frames = std::max(p.max_fix_frames, f.n_frames);
        w.head_off = f.head_off;
        w.head_rows = f.head_rows;
        w.first_clean = all_head ? nblk : Kd / hop;  // all head: every frame of the chunk belongs to the fix-up
        p.fix.push_back(f);
      }
      p.work.push_back(w);
    }
    // the tail block sums of a chunk go in front of its successor's head blocks
    for (int64_t j = 0; j + 1 < nc; ++j) {
      LaneWork& w = p.work[p.work.size() - nc + j];
      w.tail_bs_off = p.fix[cs.f0 + j].bs_off;
    }
    if (cs.n > 0) p.segs.push_back(cs);
  }
  // item order: longest chunks first (the counter hands items out in this order)
  const int n_items = (int)p.work.size() * groups;
  p.order.resize(n_items);
  std::vector<int> widx(p.work.size());
  for (size_t i = 0; i < widx.size(); ++i) widx[i] = (int)i;
  std::stable_sort(widx.begin(), widx.end(), [&](int a, int b) { return p.work[a].rows > p.work[b].rows; });
  int at = 0;
  for (int wi : widx)
    for (int g = 0; g < groups; ++g) p.order[at++] = wi * groups + g;
}

}  // namespace hga

namespace hga {

struct LaneDev {  // device copies of a LanePlan and the scratch it needs (owned by the caller)
  const LaneWork* work = nullptr;
  const LaneFix* fix = nullptr;
  const LaneCarrySeg* segs = nullptr;
  const int* order = nullptr;
  int* counter = nullptr;
  void* ybuf = nullptr;   // plan.y_rows * CP elements of the policy's TU (CP = groups * 32)
  double* bs = nullptr;   // plan.bs_rows * CP
  double* E = nullptr;    // plan.fix.size() * 32 * CP
  double* S0 = nullptr;   // same size (only used when the chunk transition does not underflow)
  const double* M = nullptr;  // 32 x 32 chunk transition, or nullptr when it underflows (chunks longer than the carry)
  const double* MT = nullptr; // M transposed (k_lane_carry_win); used when carry_window > 0
  int carry_window = 0;       // > 0: every start state sums the last carry_window end states, all pairs in parallel
  const LaneSub* subs = nullptr;  // LaneFixPlan (per handle)
  int n_sub = 0;
  const double* P = nullptr;
  // partial launches (host-buffer path: the items whose rows have arrived): `order` then points into an item list of the
  // caller's and n_items counts its entries; the fix-up kernels run once, after the last part
  int n_items = -1;       // -1: all items of the plan
  bool skip_fix = false;
  cudaEvent_t ev_main = nullptr;  // when set: recorded right behind the main kernel (hga_last_kernel_split_ms)
  // statistics partials of the STATS kernels: [items][3][cw], [fix-ups x pieces][3][CP], [fix-ups][3][CP]
  double *st_main = nullptr, *st_fix = nullptr, *st_seam = nullptr;
  // float32 arithmetic: the row of x that LaneWork::x_row == 0 names and its leading dimension (per-channel offset
  // estimate, LaneF32::kOffset); nullptr: no offset is removed
  const double* x0 = nullptr;
  int64_t ldx = 0;
};

typedef CUresult (*LaneEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// TMA tensor over the addressable rows of x: float64 (rows, C) with leading dimension ldx, boxes of 32 channels x TB rows.
inline CUresult lane_tensor_map(LaneEncodeFn enc, const double* x_base, int64_t ldx, int C, int64_t rows, int cw, int TB, CUtensorMap* tm) {
  const cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ldx * 8};
  const cuuint32_t box[2] = {(cuuint32_t)cw, (cuuint32_t)TB};
  const cuuint32_t es[2] = {1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)x_base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

inline void lane_fill_args(const LanePlan& p, const LaneDev& d, int C, const double* mu, const double* sigma, double* out, int64_t ldo, long long* prof, LaneArgs& a) {
  a = LaneArgs{};
  a.work = d.work;
  a.n_work = (int)p.work.size();
  a.groups = p.groups;
  a.n_items = d.n_items >= 0 ? d.n_items : a.n_work * p.groups;
  a.C = C;
  a.CP = p.groups * p.cw;
  a.mu = mu;
  a.sigma = sigma;
  a.out = out;
  a.ldo = ldo;
  a.ybuf = d.ybuf;
  a.bs = d.bs;
  a.E = d.E;
  a.counter = d.counter;
  a.order = d.order;
  a.prof = prof;
  a.st_main = d.st_main;
  a.x0 = d.x0;
  a.ldx = d.ldx;
}
// State carry, head fix-up and the frames touching the heads (after the main kernel, same stream).  Returns the number of
// kernels launched.
template <typename TY, int HOP, bool STATS = false>
inline int lane_launch_fix(const LaneConst& k, const LanePlan& p, const LaneDev& d, int C, const double* mu, const double* sigma, double* out, int64_t ldo,
                           cudaStream_t st) {
  const int n_fix = (int)p.fix.size();
  if (n_fix == 0 || d.skip_fix) return 0;
  const int CP = p.groups * p.cw;
  int launches = 0;
  const double* s0 = d.E;
  if (d.M) {
    if (d.MT && d.carry_window > 0) {
      const int64_t pairs = (int64_t)n_fix * C;
      k_lane_carry_win<<<(unsigned)((pairs + 3) / 4), 128, 0, st>>>(d.E, d.S0, d.MT, d.fix, n_fix, C, CP, d.carry_window);
    } else {
      k_lane_carry<<<(unsigned)(p.segs.size() * ((C + kCarryCh - 1) / kCarryCh)), 32 * kCarryCh, 0, st>>>(d.E, d.S0, d.M, d.segs, (int)p.segs.size(), C, CP);
    }
    s0 = d.S0;
    ++launches;
  }
  constexpr int CPT = LaneFixOf<TY>::type::CPT;
  const int thr = CP / CPT <= 32 ? 32 : 64;   // one warp per CTA when a row of the scratch has no more than 32 thread-columns (64 packed channels)
  const dim3 g1((unsigned)(n_fix * d.n_sub), (unsigned)((CP / CPT + thr - 1) / thr));
  if (STATS) cudaMemsetAsync(d.st_fix, 0, (size_t)n_fix * d.n_sub * 3 * CP * 8, st);   // pieces beyond a short head return without writing
  k_lane_fix<TY, HOP, STATS><<<g1, thr, (size_t)kFixRingRows * thr * 8, st>>>(k, d.fix, d.subs, d.n_sub, d.P, C, CP, d.ybuf, s0, d.bs, mu, sigma, out, ldo,
                                                                          STATS ? d.st_fix : nullptr);
  const dim3 g2((unsigned)n_fix, (unsigned)((C + 127) / 128));
  k_lane_fix_frames<<<g2, 128, 0, st>>>(k, d.fix, n_fix, C, CP, d.bs, mu, sigma, out, ldo, STATS ? d.st_seam : nullptr);
  return launches + 2;
}
template <class KernT>
inline bool lane_set_smem(KernT kern, size_t bytes, bool (&done)[64]) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (!done[dev & 63]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) return false;
    done[dev & 63] = true;
  }
  return true;
}

// The launch sequence of one extraction with the one-warp-per-item kernel: item counter reset, k_hga_lane, then (if any
// chunk has a predecessor) the fix-up kernels.  Returns the number of kernels launched, or -1 on a CUDA error.
template <class P, int HOP, int TB, int NW, int NSLOT, int U, bool PROF = false>
inline int lane_launch_t(const LaneConst& k, const LanePlan& p, const LaneDev& d, const CUtensorMap& tm, int C, const double* mu, const double* sigma,
                         double* out, int64_t ldo, int n_sm, cudaStream_t st, long long* prof = nullptr) {
  using SM = LaneSmem<32 * P::CPL, TB, NW, NSLOT>;
  auto kern = k_hga_lane<P, HOP, TB, NW, NSLOT, U, PROF>;
  static bool attr_set[64] = {};
  if (!lane_set_smem(kern, SM::total, attr_set)) return -1;
  LaneArgs a;
  lane_fill_args(p, d, C, mu, sigma, out, ldo, prof, a);
  if (cudaMemsetAsync(d.counter, 0, sizeof(int), st) != cudaSuccess) return -1;
  const int grid = std::max(1, std::min(n_sm, (a.n_items + NW - 1) / NW));
  kern<<<grid, 32 * NW, SM::total, st>>>(k, a, tm);
  if (d.ev_main) cudaEventRecord(d.ev_main, st);
  const int launches = 1 + lane_launch_fix<typename P::TY, HOP>(k, p, d, C, mu, sigma, out, ldo, st);
  return cudaGetLastError() == cudaSuccess ? launches : -1;
}
// The same with the section-pipelined kernel (NS warps per item, NI items per CTA).
template <class P, int HOP, int TB, int NI, int NS, int NSLOT, int NL, int U, bool PROF = false, bool STATS = false>
inline int lanes_launch_t(const LaneConst& k, const LanePlan& p, const LaneDev& d, const CUtensorMap& tm, int C, const double* mu, const double* sigma,
                          double* out, int64_t ldo, int n_sm, cudaStream_t st, long long* prof = nullptr) {
  using SM = LanesSmem<32 * P::CPL, TB, NI, NS, NSLOT, NL>;
  static_assert(SM::total <= 232448, "more than 227 KB of shared memory");
  auto kern = k_hga_lanes<P, HOP, TB, NI, NS, NSLOT, NL, U, PROF, false, STATS>;
  static bool attr_set[64] = {};
  if (!lane_set_smem(kern, SM::total, attr_set)) return -1;
  LaneArgs a;
  lane_fill_args(p, d, C, mu, sigma, out, ldo, prof, a);
  const int grid = std::max(1, std::min(n_sm, (a.n_items + NI - 1) / NI));
  kern<<<grid, 32 * NI * NS, SM::total, st>>>(k, a, tm);
  if (d.ev_main) cudaEventRecord(d.ev_main, st);
  const int launches = 1 + lane_launch_fix<typename P::TY, HOP, STATS>(k, p, d, C, mu, sigma, out, ldo, st);
  return cudaGetLastError() == cudaSuccess ? launches : -1;
}

// ---- fused channel map: the section-pipelined kernel with one mapper warp per item slot (k_hga_lanes<..., MAP = true>)
struct LaneMapDev {
  const double* raw = nullptr;   // raw row that LaneWork::x_row == 0 names
  int64_t ld_raw = 0, raw_rows = 0;
  int C_in = 0, n_groups = 0, n_idx = 0;
  const int32_t *sel = nullptr, *grp = nullptr, *gptr = nullptr, *gidx = nullptr;
  const double *alpha = nullptr, *beta = nullptr;
};
template <class P, int HOP, int TB, int NI, int NS, int NSLOT, int NL>
inline size_t lanes_map_smem(int c_in, int depth) {
  using SM = LanesSmem<32 * P::CPL, TB, NI, NS, NSLOT, NL>;
  return (SM::total + 15) / 16 * 16 + LaneMapSmem::per_cta + (size_t)NI * (TB / HOP) * LaneMapSmem::per_mapper(HOP, c_in, depth);
}
// Returns the number of kernels launched, -1 on a CUDA error, -2 when the map does not fit the kernel (too wide a raw row,
// too many groups): the caller then runs the channel map as a pass of its own.
template <class P, int HOP, int TB, int NI, int NS, int NSLOT, int NL, int U>
inline int lanes_launch_map_t(const LaneConst& k, const LanePlan& p, const LaneDev& d, const LaneMapDev& m, int C, const double* mu, const double* sigma,
                              double* out, int64_t ldo, int n_sm, cudaStream_t st) {
  constexpr size_t kMaxSmem = 232448;
  int depth = kMapDepth;   // raw hop blocks in flight per mapper warp: 3 when they fit, else 2
  if (lanes_map_smem<P, HOP, TB, NI, NS, NSLOT, NL>(m.C_in, depth) > kMaxSmem) depth = 2;
  const size_t bytes = lanes_map_smem<P, HOP, TB, NI, NS, NSLOT, NL>(m.C_in, depth);
  if (bytes > kMaxSmem || m.C_in > kMapMaxCols || m.n_groups > kMapMaxGroups || m.n_idx > kMapMaxIdx) return -2;
  auto kern = k_hga_lanes<P, HOP, TB, NI, NS, NSLOT, NL, U, false, true>;
  static bool attr_set[64] = {};
  if (!lane_set_smem(kern, kMaxSmem, attr_set)) return -1;
  LaneArgs a;
  lane_fill_args(p, d, C, mu, sigma, out, ldo, nullptr, a);
  a.x0 = nullptr;   // the x rows exist only inside the kernel: the offset estimate forms its three samples from the raw rows
  a.raw = m.raw;
  a.ld_raw = m.ld_raw;
  a.raw_rows = m.raw_rows;
  a.C_in = m.C_in;
  a.n_groups = m.n_groups;
  a.n_idx = m.n_idx;
  a.map_depth = depth;
  a.m_sel = m.sel;
  a.m_grp = m.grp;
  a.m_gptr = m.gptr;
  a.m_gidx = m.gidx;
  a.m_alpha = m.alpha;
  a.m_beta = m.beta;
  CUtensorMap unused;
  memset(&unused, 0, sizeof unused);
  const int grid = std::max(1, std::min(n_sm, (a.n_items + NI - 1) / NI));
  kern<<<grid, 32 * (NI * NS + NI * (TB / HOP)), bytes, st>>>(k, a, unused);
  if (d.ev_main) cudaEventRecord(d.ev_main, st);
  const int launches = 1 + lane_launch_fix<typename P::TY, HOP>(k, p, d, C, mu, sigma, out, ldo, st);
  return cudaGetLastError() == cudaSuccess ? launches : -1;
}

}  // namespace hga
