// Micro-benchmark for the section-parallel ("systolic") biquad cascade on sm_100a: how many biquad section-samples per
// clock per SM the FP64 pipe sustains when the cascade's sections are spread over lanes (shuffle hand-off), over warps
// (shared-memory hand-off) or kept in one thread, at several warps/SM.  Also prints DFMA / SHFL / LDS latencies.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o systolic_rates systolic_rates.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int STEPS = 4096;


// MODE 0: NS sections per lane, hand-off by shuffle, no skew (consume the neighbour's output of the previous step)
// MODE 1: same with one extra step of skew (shuffle issued a step ahead)
// MODE 2: NS sections per lane, input from shared memory, output to shared memory (warp-level section groups)
// MODE 3: NS sections per lane, no hand-off at all (sequential kernel's inner loop)
// MODE 4: MODE 1 in float (scalar FFMA)
template <int MODE, int NS, typename T>
__global__ void __launch_bounds__(MODE == 3 ? 256 : NS >= 8 ? 512 : 1024) k(T* out, const T* coef, long long* cycles, int lanes_per_group) {
  extern __shared__ double smem[];
  T z0[NS], z1[NS], a1[NS], a2[NS], b1[NS];
  const int lane = threadIdx.x & 31;
  const int stage = lane % lanes_per_group;
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    z0[s] = (T)(1e-3 * threadIdx.x + s); z1[s] = (T)(2e-3 * threadIdx.x + s);
    a1[s] = coef[(stage * NS + s) % 48]; a2[s] = coef[(stage * NS + s + 16) % 48] * (T)0.25; b1[s] = coef[(stage * NS + s + 32) % 48];
  }
  T* sbuf = reinterpret_cast<T*>(smem) + (threadIdx.x >> 5) * 8 * 32;  // per-warp ring: 8 rows x 32 lanes
  for (int i = lane; i < 8 * 32; i += 32) sbuf[i] = (T)(1e-3 * i);
  __syncthreads();
  T y_out = (T)0, x_next = (T)0;
  long long t0 = clock64();
#pragma unroll 4
  for (int k = 0; k < STEPS; ++k) {
    T x;
    if (MODE == 0) {
      x = __shfl_up_sync(0xffffffffu, y_out, 1);
      if (stage == 0) x = sbuf[(k & 7) * 32 + lane];
    } else if (MODE == 1 || MODE == 4) {
      x = x_next;
      x_next = __shfl_up_sync(0xffffffffu, y_out, 1);
      if (stage == 0) x_next = sbuf[(k & 7) * 32 + lane];
    } else if (MODE == 2) {
      x = sbuf[(k & 7) * 32 + lane];
    } else {
      x = y_out * (T)1e-9 + (T)k;  // cheap synthetic input (1 extra op per NS sections)
    }
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const T y = x + z0[s];
      z0[s] = fma(-a1[s], y, fma(b1[s], x, z1[s]));
      z1[s] = fma(-a2[s], y, x);
      x = y;
    }
    y_out = x;
    if (MODE == 2) sbuf[((k + 4) & 7) * 32 + lane] = y_out;
  }
  long long t1 = clock64();
  T s = y_out + x_next;
#pragma unroll
  for (int i = 0; i < NS; ++i) s += z0[i] + z1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE, int NS, typename T>
void run(const char* name, int warps_per_sm, int lanes_per_group) {
  int nsm; CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  T* out; T* coef; long long* cyc;
  CK(cudaMalloc(&out, sizeof(T) * nsm * 2048)); CK(cudaMalloc(&cyc, sizeof(long long) * nsm * 2)); CK(cudaMalloc(&coef, sizeof(T) * 48));
  T hc[48]; for (int i = 0; i < 48; ++i) hc[i] = (T)(-1.4 + 0.01 * i);
  CK(cudaMemcpy(coef, hc, sizeof hc, cudaMemcpyHostToDevice));
  int blocks_per_sm = warps_per_sm > 32 ? 2 : 1;
  int threads = warps_per_sm / blocks_per_sm * 32;
  size_t smem = (size_t)(threads / 32) * 8 * 32 * sizeof(T);
  CK(cudaFuncSetAttribute(k<MODE, NS, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  k<MODE, NS, T><<<nsm * blocks_per_sm, threads, smem>>>(out, coef, cyc, lanes_per_group); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  k<MODE, NS, T><<<nsm * blocks_per_sm, threads, smem>>>(out, coef, cyc, lanes_per_group);
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  long long h[512]; CK(cudaMemcpy(h, cyc, sizeof(long long) * nsm * blocks_per_sm, cudaMemcpyDeviceToHost));
  double mc = 0; for (int i = 0; i < nsm * blocks_per_sm; ++i) mc += h[i]; mc /= nsm * blocks_per_sm;
  double secsamp = (double)STEPS * NS * warps_per_sm * 32;  // lane section-samples per SM
  // event-timed rate uses the whole launch (all SMs busy the whole time by construction)
  printf("%-34s NS=%2d warps/SM=%2d  cyc/step=%7.1f  section-samples/clk/SM=%6.2f  (FP pipe ops/clk/SM=%6.1f)  ms=%.3f\n", name, NS,
         warps_per_sm, mc / STEPS, secsamp / mc, secsamp / mc * 4, ms);
  CK(cudaFree(out)); CK(cudaFree(cyc)); CK(cudaFree(coef));
}

// latencies: dependent chains, 1 warp per SM
__global__ void k_lat(double* out, long long* cyc, double a, double b) {
  double d = threadIdx.x;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < 4096; ++i) d = fma(d, a, b);
  long long t1 = clock64();
  double e = d;
#pragma unroll 16
  for (int i = 0; i < 4096; ++i) e = e + b;
  long long t2 = clock64();
  double f = e;
#pragma unroll 16
  for (int i = 0; i < 4096; ++i) f = __shfl_up_sync(0xffffffffu, f, 1);
  long long t3 = clock64();
  float g = (float)f;
#pragma unroll 16
  for (int i = 0; i < 4096; ++i) g = fmaf(g, (float)a, (float)b);
  long long t4 = clock64();
  out[threadIdx.x] = d + e + f + g;
  if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; }
}

// shuffle throughput: 8 independent 32-bit shuffles per iteration
__global__ void __launch_bounds__(1024) k_shfl(float* out, long long* cyc) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < 2048; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __shfl_up_sync(0xffffffffu, v[i], 1);
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  printf("device %s  SMs %d  clock %d kHz\n", pr.name, pr.multiProcessorCount, pr.clockRate);
  {
    double* o; long long* c; CK(cudaMalloc(&o, 8 * 32)); CK(cudaMalloc(&c, 8 * 4));
    k_lat<<<1, 32>>>(o, c, 0.999, 0.5); CK(cudaDeviceSynchronize());
    k_lat<<<1, 32>>>(o, c, 0.999, 0.5); CK(cudaDeviceSynchronize());
    long long h[4]; CK(cudaMemcpy(h, c, 32, cudaMemcpyDeviceToHost));
    printf("latency (dependent chain, 1 warp): DFMA %.2f  DADD %.2f  SHFL.64(2xSHFL.32) %.2f  FFMA %.2f cycles\n", h[0] / 4096.0, h[1] / 4096.0,
           h[2] / 4096.0, h[3] / 4096.0);
    int nsm = pr.multiProcessorCount;
    float* of; long long* cc; CK(cudaMalloc(&of, 4 * nsm * 1024)); CK(cudaMalloc(&cc, 8 * nsm));
    for (int w : {4, 8, 16, 32}) {
      k_shfl<<<nsm, w * 32>>>(of, cc); CK(cudaDeviceSynchronize());
      long long hh[256]; CK(cudaMemcpy(hh, cc, 8 * nsm, cudaMemcpyDeviceToHost));
      double mc = 0; for (int i = 0; i < nsm; ++i) mc += hh[i]; mc /= nsm;
      printf("SHFL.32 throughput warps/SM=%2d: %.3f warp-instr/clk/SM\n", w, 2048.0 * 8 * w / mc);
    }
  }
  for (int w : {4, 8, 16, 32, 64}) {
    if (w <= 8) run<3, 16, double>("seq: 16 sections/thread", w, 1);
    run<0, 1, double>("systolic16 shfl, no skew", w, 16);
    run<1, 1, double>("systolic16 shfl, skew 2", w, 16);
    run<0, 2, double>("systolic8 shfl, no skew", w, 8);
    run<1, 2, double>("systolic8 shfl, skew 2", w, 8);
    run<1, 4, double>("systolic4 shfl, skew 2", w, 4);
    if (w <= 32) {
      run<2, 2, double>("warp-groups smem, 2 sec/warp", w, 1);
      run<2, 4, double>("warp-groups smem, 4 sec/warp", w, 1);
      if (w <= 16) run<2, 8, double>("warp-groups smem, 8 sec/warp", w, 1);
    }
    run<4, 1, float>("f32 systolic16 shfl, skew 2", w, 16);
    run<4, 2, float>("f32 systolic8 shfl, skew 2", w, 8);
    if (w <= 8) run<3, 16, float>("f32 seq: 16 sections/thread", w, 1);
  }
  return 0;
}
