// hga_async.cuh — sm_100a primitives shared by the HGA kernels: mbarrier, TMA (cp.async.bulk.tensor) and the
// table-driven float64 log used by every framing stage.
#pragma once
#include <cuda.h>
#include <cmath>
#include <cstdint>

namespace hga {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
// TMA: one (box columns x box rows) tile of a 2-D tensor -> shared memory, completion on an mbarrier.  Out-of-range
// columns (C not a multiple of the box width) and rows past the end of the tensor are zero-filled by the hardware.
// per-thread asynchronous 8-byte copies global -> shared (the fix-up kernel's row ring)
__device__ __forceinline__ void cp_async8(void* dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8_s(uint32_t dst_smem, const void* src) {   // shared address already converted
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_s(uint32_t dst_smem, const void* src) {  // both addresses 16-byte aligned
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int row, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(c0), "r"(row), "r"(smem_u32(bar))
      : "memory");
}

// Two float32 lanes in one 64-bit register (fma.rn.f32x2 / add.rn.f32x2, SASS FFMA2 / FADD2): one lane of a warp filters
// TWO channels (c and c + 32) per instruction, which halves the issue slots of the float32 cascade.
struct f32x2 {
  unsigned long long v;
};
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
__device__ __forceinline__ f32x2 operator+(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f32x2 operator-(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
using ::fma;  // keep the scalar overloads visible next to the packed one
__device__ __forceinline__ f32x2 fma(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}

// Table-driven natural log for finite x > 0: x = m * 2^e, m in [1, 2); the top 7 mantissa bits pick c = 1 + (i + 1/2) / 128,
// log m = -log(1/c) + log1p(r) with r = m * (1/c) - 1 (one FMA, |r| <= 2^-8) and a degree-6 series (truncation < 2e-18).
// The table holds the ROUNDED 1/c and -log of that rounded value (computed in long double on the host, hga_api.cu), so
// the split is exact.  ~10 FP64 instructions.  |error| <= 3e-16 over [1e-2, 1e15] (tools/check_fast_log.py).
constexpr int kLogTabN = 128;
__device__ double g_logtab[2 * kLogTabN];
__device__ __forceinline__ double tab_log(double x, const double* __restrict__ tab) {
  const long long bits = __double_as_longlong(x);
  const int hi = (int)(bits >> 32);
  const int e = ((hi >> 20) & 0x7ff) - 1023;
  const int idx = (hi >> 13) & (kLogTabN - 1);
  const double m = __longlong_as_double((bits & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);
  const double2 t = *reinterpret_cast<const double2*>(tab + 2 * idx);
  const double r = fma(m, t.x, -1.0);
  double qv = fma(r, -1.0 / 6.0, 1.0 / 5.0);
  qv = fma(r, qv, -1.0 / 4.0);
  qv = fma(r, qv, 1.0 / 3.0);
  qv = fma(r, qv, -0.5);
  const double ed = (double)e;
  const double lo = fma(r * r, qv, r);
  return fma(ed, 0x1.62e42fee00000p-1, t.y) + fma(ed, 0x1.a39ef35793c76p-33, lo);
}
// Host side of the table: entry i = {rc, -log(rc)} with rc = double(1 / (1 + (i + 1/2) / 128)).
inline void fill_log_table(double (&tab)[2 * kLogTabN]) {
  for (int i = 0; i < kLogTabN; ++i) {
    const double rc = (double)(1.0L / (1.0L + ((long double)i + 0.5L) / (long double)kLogTabN));
    tab[2 * i] = rc;
    tab[2 * i + 1] = (double)(-logl((long double)rc));
  }
}

}  // namespace hga
