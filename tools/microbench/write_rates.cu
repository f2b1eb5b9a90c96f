// write_rates.cu — what bounds a write-dominated stream on B200?  Times, on 3.69 GB of float64 output (K2-sized):
//   cudaMemsetAsync; a store-only kernel with 16 B plain / 16 B streaming (st.global.cs) / 32 B (st.global.v4.b64) stores;
//   and the raw-ingest shape (2 B or 4 B read per 8 B written) with each store flavour.
// Build + run on the box: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/write_rates tools/microbench/write_rates.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

enum { PLAIN = 0, CS = 1, V4 = 2 };

template <int MODE>
__device__ __forceinline__ void store4(double* p, double a, double b, double c, double d) {
  if constexpr (MODE == PLAIN) {
    reinterpret_cast<double2*>(p)[0] = make_double2(a, b);
    reinterpret_cast<double2*>(p)[1] = make_double2(c, d);
  } else if constexpr (MODE == CS) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(a, b));
    __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(c, d));
  } else {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
  }
}

template <int MODE>
__global__ void __launch_bounds__(256) k_fill(double* out, int64_t n_quads, double v) {
  const int64_t S = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n_quads; i += S) store4<MODE>(out + 4 * i, v, v + 1, v + 2, v + 3);
}

template <typename Q, int MODE>
__global__ void __launch_bounds__(256) k_widen(const Q* __restrict__ x, double* out, int64_t n_quads, double g) {
  const int64_t S = (int64_t)gridDim.x * 256;
  int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  for (; i + 3 * S < n_quads; i += 4 * S) {
    Q r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = __ldg(x + i + j * S);
#pragma unroll
    for (int j = 0; j < 4; ++j) store4<MODE>(out + 4 * (i + j * S), r[j].x * g, r[j].y * g, r[j].z * g, r[j].w * g);
  }
  for (; i < n_quads; i += S) {
    const Q r = __ldg(x + i);
    store4<MODE>(out + 4 * i, r.x * g, r.y * g, r.z * g, r.w * g);
  }
}

template <typename F>
static float timed(F f, int n = 10) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  f(); f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < n; ++i) f();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms / n;
}

int main() {
  const int64_t T = 3600000, C = 128, N = T * C, nq = N / 4;
  double* out; void* in;
  cudaMalloc(&out, N * 8);
  cudaMalloc(&in, N * 4);
  cudaMemset(in, 1, N * 4);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  auto rep = [&](const char* name, float ms, double bytes) { printf("%-44s %7.3f ms  %7.1f GB/s  (writes %6.1f GB/s)\n", name, ms, bytes / ms / 1e6, N * 8.0 / ms / 1e6); };
  rep("cudaMemsetAsync 3.69 GB", timed([&] { cudaMemsetAsync(out, 0, N * 8); }), N * 8.0);
  for (int per = 4; per <= 16; per *= 2) {
    const unsigned grid = sms * per;
    char nm[96];
    snprintf(nm, sizeof nm, "fill 16 B stores            (%2d CTAs/SM)", per); rep(nm, timed([&] { k_fill<PLAIN><<<grid, 256>>>(out, nq, 1.0); }), N * 8.0);
    snprintf(nm, sizeof nm, "fill 16 B st.cs             (%2d CTAs/SM)", per); rep(nm, timed([&] { k_fill<CS><<<grid, 256>>>(out, nq, 1.0); }), N * 8.0);
    snprintf(nm, sizeof nm, "fill 32 B st.v4.f64         (%2d CTAs/SM)", per); rep(nm, timed([&] { k_fill<V4><<<grid, 256>>>(out, nq, 1.0); }), N * 8.0);
  }
  const unsigned grid = sms * 4;
  rep("int16 -> f64, 16 B stores", timed([&] { k_widen<short4, PLAIN><<<grid, 256>>>((const short4*)in, out, nq, 0.1); }), N * 10.0);
  rep("int16 -> f64, st.cs", timed([&] { k_widen<short4, CS><<<grid, 256>>>((const short4*)in, out, nq, 0.1); }), N * 10.0);
  rep("int16 -> f64, st.v4.f64", timed([&] { k_widen<short4, V4><<<grid, 256>>>((const short4*)in, out, nq, 0.1); }), N * 10.0);
  rep("float32 -> f64, 16 B stores", timed([&] { k_widen<float4, PLAIN><<<grid, 256>>>((const float4*)in, out, nq, 0.1); }), N * 12.0);
  rep("float32 -> f64, st.cs", timed([&] { k_widen<float4, CS><<<grid, 256>>>((const float4*)in, out, nq, 0.1); }), N * 12.0);
  rep("float32 -> f64, st.v4.f64", timed([&] { k_widen<float4, V4><<<grid, 256>>>((const float4*)in, out, nq, 0.1); }), N * 12.0);
  printf("cuda status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
