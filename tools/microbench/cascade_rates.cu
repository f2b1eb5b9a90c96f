// Micro-benchmark of the filter-warp inner loop of k_hga_pipe / k_hga_pipe2 in isolation: one warp = NS biquad sections
// of 32 (f64) or 64 (f32x2) channels, rows read from a shared-memory tile and written to another one, no
// synchronisation.  Prints cycles per row per warp and FP-pipe lane-ops/clk/SM for several warps/SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../delayed_speech_synthesis_b200/csrc -o cascade_rates cascade_rates.cu
#include <cstdio>
#include <cstdlib>
#include "hga_kernels_pipe2.cuh"
using namespace hga;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <typename T, int NS, int TB, int MODE>
__global__ void __launch_bounds__(512) k(const __grid_constant__ Cascade q, const __grid_constant__ PipeConst pc, int ntiles, long long* cyc, double* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  T* tile_in = (T*)smem + (size_t)g * 2 * TB * 32;
  T* tile_out = tile_in + TB * 32;
  for (int i = lane; i < 2 * TB * 32; i += 32) tile_in[i] = PipeTraits<T>::splat(1e-3 * (i % 97));
  T z0[NS], z1[NS], yp[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) { z0[s] = PipeTraits<T>::splat(1e-3 * s); z1[s] = PipeTraits<T>::splat(2e-3 * s); yp[s] = PipeTraits<T>::zero(); }
  const Coef<T, NS, (MODE == 3 ? 0 : -1)> cf(q, pc, g % (kMaxSec / NS));
  __syncthreads();
  const long long t0 = clock64();
  for (int m = 0; m < ntiles; ++m) {
    if (MODE == 0 || MODE == 3) pipe2_tile<T, NS, TB>(cf, tile_in + lane, tile_out + lane, yp, z0, z1);
    else if (MODE == 1) pipe_rows<T, T, 0, NS, TB>(q, pc, tile_in + lane, tile_out + lane, TB, z0, z1);
    else {  // MODE 2: every warp of a sub-partition runs its own copy of the loop (section group = warp index)
      constexpr int GM = kMaxSec / NS;
#define RUNG(G_) if (g % GM == G_) pipe_rows<T, T, (G_ < GM ? G_ : 0), NS, TB>(q, pc, tile_in + lane, tile_out + lane, TB, z0, z1);
      RUNG(0) RUNG(1) RUNG(2) RUNG(3) RUNG(4) RUNG(5) RUNG(6) RUNG(7)
#undef RUNG
    }
  }
  const long long t1 = clock64();
  if (lane == 0) cyc[blockIdx.x * (blockDim.x / 32) + g] = t1 - t0;
  double o[PipeTraits<T>::CPL];
  PipeTraits<T>::split(z0[NS - 1], o);
  if (o[0] == 123.456) sink[0] = o[0];
}

// conversion throughput: MODE 0 cvt.rn.f32.f64, 1 cvt.f64.f32, 2 integer f64->f32 (RNE, normal range), 3 integer f32->f64
__device__ __forceinline__ float d2f_int(double x) {
  const unsigned hi = (unsigned)__double2hiint(x), lo = (unsigned)__double2loint(x);
  const unsigned sign = hi & 0x80000000u;
  const unsigned ah = hi & 0x7fffffffu;
  // mantissa: 20 bits in hi, 32 in lo -> keep 23: (ah << 3 | lo >> 29), exponent re-biased by subtracting 896 << 23
  unsigned m = (ah << 3) | (lo >> 29);
  const unsigned rest = lo << 3;                       // dropped 29 bits, left-aligned
  m += (rest > 0x80000000u) || (rest == 0x80000000u && (m & 1u));
  unsigned r = m - (896u << 23);
  r = ah < (897u << 20) ? 0u : r;                      // |x| < 2^-126: flush to zero
  r = ah >= (1151u << 20) ? 0x7f800000u : r;           // overflow / inf / nan -> inf
  return __uint_as_float(sign | r);
}
__device__ __forceinline__ double f2d_int(float f) {
  const unsigned u = __float_as_uint(f);
  const unsigned sign = u & 0x80000000u, a = u & 0x7fffffffu;
  unsigned hi = (a >> 3) + (896u << 20);
  const unsigned lo = a << 29;
  hi = a < 0x00800000u ? 0u : hi;                      // zero / denormal -> 0
  return __hiloint2double((int)(sign | hi), (int)lo);
}
template <int MODE>
__global__ void __launch_bounds__(512) kcvt(int iters, long long* cyc, double* sink, const double* src) {
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  double a[4]; float f[4];
  for (int i = 0; i < 4; ++i) { a[i] = src[(threadIdx.x * 4 + i) % 64]; f[i] = (float)a[i]; }
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (MODE == 0) { f[i] = (float)a[i] + f[i]; a[i] = __hiloint2double(__double2hiint(a[i]) ^ (it & 1), __double2loint(a[i])); }
      if (MODE == 1) { a[i] = (double)f[i]; f[i] = __uint_as_float(__float_as_uint(f[i]) ^ __double2loint(a[i])); }
      if (MODE == 2) { f[i] = d2f_int(a[i]) + f[i]; a[i] = __hiloint2double(__double2hiint(a[i]) ^ (it & 1), __double2loint(a[i])); }
      if (MODE == 3) { a[i] = f2d_int(f[i]); f[i] = __uint_as_float(__float_as_uint(f[i]) ^ (__double2loint(a[i]) >> 31)); }
    }
  }
  const long long t1 = clock64();
  if (lane == 0) cyc[blockIdx.x * (blockDim.x / 32) + g] = t1 - t0;
  if (a[0] + a[1] + a[2] + a[3] + f[0] + f[1] + f[2] + f[3] == 123.456) sink[0] = a[0];
}
template <int MODE>
void run_cvt(const char* name, int wps) {
  int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  long long* cyc; double *sink, *src;
  CK(cudaMalloc(&cyc, nsm * wps * 8)); CK(cudaMalloc(&sink, 8)); CK(cudaMalloc(&src, 64 * 8));
  double h[64]; for (int i = 0; i < 64; ++i) h[i] = 1e-3 * (i + 1) * ((i & 1) ? -1 : 1);
  CK(cudaMemcpy(src, h, sizeof h, cudaMemcpyHostToDevice));
  const int iters = 2000;
  kcvt<MODE><<<nsm, wps * 32>>>(iters, cyc, sink, src); CK(cudaDeviceSynchronize());
  kcvt<MODE><<<nsm, wps * 32>>>(iters, cyc, sink, src); CK(cudaDeviceSynchronize());
  long long* hc = (long long*)malloc(nsm * wps * 8);
  CK(cudaMemcpy(hc, cyc, nsm * wps * 8, cudaMemcpyDeviceToHost));
  double mean = 0; for (int i = 0; i < nsm * wps; ++i) mean += (double)hc[i]; mean /= nsm * wps;
  printf("%-34s warps/SM=%2d  cycles per conversion per warp=%7.2f  conversions/clk/SM=%6.2f\n", name, wps, mean / (iters * 4.0), wps * 32.0 * iters * 4.0 / mean);
  free(hc); cudaFree(cyc); cudaFree(sink); cudaFree(src);
}

// tile conversion loop of the float32 mode: TB rows x 64 float64 (smem) -> packed float32 pairs (smem)
template <int MODE, int TB>
__global__ void __launch_bounds__(512) ktile(int ntiles, long long* cyc) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  double* in = (double*)smem + (size_t)g * TB * 96;
  f32x2* out = (f32x2*)(in + TB * 64);
  for (int i = lane; i < TB * 64; i += 32) in[i] = 1e-3 * (i % 97) + 1.0;
  __syncthreads();
  const long long t0 = clock64();
  for (int m = 0; m < ntiles; ++m) {
    if (MODE == 2) {
      convert_rows<TB>(in + lane, out + lane);
    } else {
#pragma unroll 10
      for (int r = 0; r < TB; ++r) {
        const double* p = in + r * 64 + lane;
        out[r * 32 + lane] = MODE == 0 ? pack2((float)p[0], (float)p[32]) : pack2(d2f_int(p[0]), d2f_int(p[32]));
      }
    }
    __syncwarp();
  }
  const long long t1 = clock64();
  if (lane == 0) cyc[blockIdx.x * (blockDim.x / 32) + g] = t1 - t0;
}
template <int MODE, int TB>
void run_tile(const char* name, int wps) {
  int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  long long* cyc; CK(cudaMalloc(&cyc, nsm * wps * 8));
  const size_t smem = (size_t)wps * TB * 96 * 8;
  auto kern = ktile<MODE, TB>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int ntiles = 400;
  kern<<<nsm, wps * 32, smem>>>(ntiles, cyc); CK(cudaDeviceSynchronize());
  kern<<<nsm, wps * 32, smem>>>(ntiles, cyc); CK(cudaDeviceSynchronize());
  long long* hc = (long long*)malloc(nsm * wps * 8);
  CK(cudaMemcpy(hc, cyc, nsm * wps * 8, cudaMemcpyDeviceToHost));
  double mean = 0; for (int i = 0; i < nsm * wps; ++i) mean += (double)hc[i]; mean /= nsm * wps;
  printf("%-34s TB=%d warps/SM=%2d  cycles per row (64 ch) per warp=%7.2f\n", name, TB, wps, mean / (ntiles * (double)TB));
  free(hc); cudaFree(cyc);
}

template <typename T, int NS, int TB, int MODE>
void run(const char* name, int wps) {
  Cascade q{}; PipeConst pc{};
  for (int s = 0; s < 16; ++s) {
    q.a1[s] = -1.4 + 0.01 * s; q.a2[s] = 0.95 - 0.01 * s; q.b1u[s] = s < 4 ? 2.0 : -1.45;
    auto pack = [](double v) { float f = (float)v; uint32_t u; memcpy(&u, &f, 4); return ((unsigned long long)u << 32) | u; };
    pc.na1p[s] = pack(-q.a1[s]); pc.na2p[s] = pack(-q.a2[s]); pc.b1p[s] = pack(q.b1u[s]);
    pc.c1d[s] = -q.a1[s]; pc.c2d[s] = -q.a2[s]; pc.cbd[s] = q.b1u[s];   // k_hga_pipe2's own operand arrays (section form U)
  }
  int dev = 0, nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  const int ntiles = 400;
  long long* cyc; double* sink;
  CK(cudaMalloc(&cyc, nsm * wps * 8)); CK(cudaMalloc(&sink, 8));
  const size_t smem = (size_t)wps * 2 * TB * 32 * sizeof(T);
  auto kern = k<T, NS, TB, MODE>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<nsm, wps * 32, smem>>>(q, pc, ntiles, cyc, sink);
  CK(cudaDeviceSynchronize());
  kern<<<nsm, wps * 32, smem>>>(q, pc, ntiles, cyc, sink);
  CK(cudaDeviceSynchronize());
  long long* h = (long long*)malloc(nsm * wps * 8);
  CK(cudaMemcpy(h, cyc, nsm * wps * 8, cudaMemcpyDeviceToHost));
  double mean = 0; for (int i = 0; i < nsm * wps; ++i) mean += (double)h[i]; mean /= nsm * wps;
  const double per_row = mean / ((double)ntiles * TB);
  const double laneops = (double)wps * NS * 4 * 32 * PipeTraits<T>::CPL / per_row;
  printf("%-34s NS=%d TB=%2d warps/SM=%2d  cycles/row/warp=%7.2f  FP lane-ops/clk/SM=%6.1f\n", name, NS, TB, wps, per_row, laneops);
  free(h); cudaFree(cyc); cudaFree(sink);
}

int main() {
  for (int wps : {1, 2, 4}) {
    run_tile<0, 20>("tile convert, F2F", wps);
    run_tile<1, 20>("tile convert, integer RNE", wps);
    run_tile<2, 20>("tile convert, F2F, batched loads", wps);
  }
  for (int wps : {1, 4, 8}) {
    run_cvt<0>("cvt.rn.f32.f64 (F2F)", wps);
    run_cvt<1>("cvt.f64.f32 (F2F)", wps);
    run_cvt<2>("f64->f32 integer RNE", wps);
    run_cvt<3>("f32->f64 integer", wps);
  }
  for (int wps : {8, 16}) {
    run<double, 2, 20, 2>("f64 pipe_rows (v1), own code per warp", wps);
    run<f32x2, 2, 20, 2>("f32x2 pipe_rows (v1), own code/warp", wps);
    run<double, 4, 20, 2>("f64 pipe_rows (v1), own code per warp", wps);
  }
  for (int wps : {4, 8}) {
    run<f32x2, 2, 20, 0>("f32x2 pipe2_tile", wps);
    run<f32x2, 4, 20, 0>("f32x2 pipe2_tile", wps);
    run<f32x2, 4, 40, 0>("f32x2 pipe2_tile", wps);
    run<f32x2, 4, 40, 1>("f32x2 pipe_rows (v1)", wps);
    run<double, 2, 20, 0>("f64 pipe2_tile", wps);
    run<double, 4, 20, 0>("f64 pipe2_tile", wps);
    run<double, 4, 40, 0>("f64 pipe2_tile", wps);
    run<double, 4, 40, 1>("f64 pipe_rows (v1)", wps);
  }
  return 0;
}
