// Micro-benchmark: issue rates of the arithmetic the HGA IIR cascade is made of, on sm_100a.
// Prints lane-ops per clock per SM for scalar FFMA (reg / constant-bank coefficient), packed
// fma.rn.f32x2, DFMA (reg / constant), and a DFMA+FFMA2 mix, at 1/2/4 warps per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__constant__ float  c_f[8];
__constant__ double c_d[8];
__constant__ float2 c_f2[8];

constexpr int CH = 8;       // independent dependency chains per thread
constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(512) k(float* out, float a_in, double d_in, long long* cycles) {
  float  f[CH];  double d[CH];  unsigned long long p[CH];
  float a = a_in, b = a_in * 0.5f; double da = d_in, db = d_in * 0.5;
  unsigned long long pa, pb;
  asm("mov.b64 %0, {%1, %2};" : "=l"(pa) : "f"(a), "f"(a));
  asm("mov.b64 %0, {%1, %2};" : "=l"(pb) : "f"(b), "f"(b));
#pragma unroll
  for (int i = 0; i < CH; ++i) { f[i] = threadIdx.x * 1e-3f + i; d[i] = threadIdx.x * 1e-3 + i;
    asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(f[i]), "f"(f[i] + 1.f)); }
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (MODE == 0) f[i] = fmaf(f[i], a, b);                    // FFMA reg,reg,reg
      if (MODE == 1) f[i] = fmaf(f[i], c_f[1], c_f[2]);          // FFMA reg,const,const?
      if (MODE == 2) f[i] = fmaf(f[i], c_f[1], b);               // FFMA reg,const,reg
      if (MODE == 3) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pa), "l"(pb));
      if (MODE == 4) d[i] = fma(d[i], da, db);                   // DFMA reg
      if (MODE == 5) d[i] = fma(d[i], c_d[1], db);               // DFMA const
      if (MODE == 6) { if (i & 1) d[i] = fma(d[i], da, db);      // DFMA + FFMA2 mix (1:1)
                       else asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pa), "l"(pb)); }
      if (MODE == 7) f[i] = f[i] + b;                            // FADD
      if (MODE == 8) { unsigned long long cc; asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c_f[1]));
                       asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cc), "l"(pb)); }
      if (MODE == 9) { if (i & 1) d[i] = fma(d[i], da, db); else f[i] = fmaf(f[i], c_f[1], b); }  // DFMA + FFMA mix
      if (MODE == 10) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pb));
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < CH; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); s += f[i] + (float)d[i] + lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int lanes_per_instr, int warps_per_sm) {
  int nsm; CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  float* out; long long* cyc; CK(cudaMalloc(&out, sizeof(float) * nsm * 1024)); CK(cudaMalloc(&cyc, sizeof(long long) * nsm));
  int threads = warps_per_sm * 32;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  k<MODE><<<nsm, threads>>>(out, 0.999f, 0.999, cyc); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  k<MODE><<<nsm, threads>>>(out, 0.999f, 0.999, cyc);
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  long long h[256]; CK(cudaMemcpy(h, cyc, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
  double mc = 0; for (int i = 0; i < nsm; ++i) mc += h[i]; mc /= nsm;
  double instr = (double)ITERS * CH * warps_per_sm;              // warp-instr per SM
  printf("%-28s warps/SM=%2d  cyc=%9.0f  warp-instr/clk/SM=%6.3f  lane-ops/clk/SM=%7.1f  ms=%.3f  MHz~%.0f\n",
         name, warps_per_sm, mc, instr / mc, instr / mc * 32 * lanes_per_instr, ms, mc / ms / 1e3);
  CK(cudaFree(out)); CK(cudaFree(cyc));
}

int main() {
  float hf[8] = {1, 0.999f, 0.5f, 0, 0, 0, 0, 0}; double hd[8] = {1, 0.999, 0.5, 0, 0, 0, 0, 0};
  CK(cudaMemcpyToSymbol(c_f, hf, sizeof hf)); CK(cudaMemcpyToSymbol(c_d, hd, sizeof hd));
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  printf("device %s  SMs %d  clock %d kHz\n", pr.name, pr.multiProcessorCount, pr.clockRate);
  for (int w : {4, 8, 16}) {
    run<0>("FFMA r,r,r", 1, w);  run<1>("FFMA r,c,c", 1, w);  run<2>("FFMA r,c,r", 1, w);
    run<3>("FFMA2 (f32x2) r,r,r", 2, w);  run<8>("FFMA2 coeff from const", 2, w);
    run<10>("FADD2 (f32x2)", 2, w);
    run<4>("DFMA r,r,r", 1, w);  run<5>("DFMA r,c,r", 1, w);
    run<6>("DFMA+FFMA2 1:1 (instr)", 1, w);  run<9>("DFMA+FFMA 1:1 (instr)", 1, w);
    run<7>("FADD", 1, w);
  }
  return 0;
}
