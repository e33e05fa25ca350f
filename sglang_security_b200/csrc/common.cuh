// Shared helpers for the sm_100a hot-path kernels: status/error plumbing for the C-ABI,
// launch accounting, small PTX wrappers (cp.async, packed f32x2 FMA, bf16/f16 unpack).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/b200_hotpath.h"

namespace b200 {

// ---- status / error text ------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launch_count;

#define B200_REQUIRE(cond, ...)        \
  do {                                 \
    if (!(cond)) {                     \
      ::b200::set_error(__VA_ARGS__);  \
      return 1;                        \
    }                                  \
  } while (0)

// Called right after a kernel launch: counts it and converts launch errors to a status.
#define B200_CHECK_LAUNCH(name)                                                       \
  do {                                                                                \
    ::b200::g_launch_count.fetch_add(1, std::memory_order_relaxed);                   \
    cudaError_t e__ = cudaPeekAtLastError();                                          \
    if (e__ != cudaSuccess) {                                                         \
      ::b200::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));      \
      (void)cudaGetLastError();                                                       \
      return 2;                                                                       \
    }                                                                                 \
  } while (0)

#define B200_CUDA_OK(expr)                                                            \
  do {                                                                                \
    cudaError_t e__ = (expr);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      ::b200::set_error("%s failed: %s", #expr, cudaGetErrorString(e__));             \
      (void)cudaGetLastError();                                                       \
      return 3;                                                                       \
    }                                                                                 \
  } while (0)

// Optional per-kernel device timing (bench.py's roofline leg): when enabled through
// b200_timing_enable(1), every launch wrapped in a ScopedKernelTimer is bracketed by CUDA events
// on the launching stream; b200_timing_read() accumulates the elapsed times per kernel name.
struct ScopedKernelTimer {
  ScopedKernelTimer(const char* name, cudaStream_t st);
  ~ScopedKernelTimer();
  const char* name_;
  cudaStream_t st_;
  cudaEvent_t start_ = nullptr;
};

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
int device_sm_count();

// ---- device helpers -----------------------------------------------------------------------
__host__ __device__ constexpr int ceil_div(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte asynchronous global->shared copy (LDGSTS), L2-only caching: KV rows are streamed once.
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_addr), "l"(gptr));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ uint4 lds128(uint32_t smem_addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(smem_addr));
  return r;
}

// Packed two-wide fp32 FMA (sm_100 FFMA2): d = a*b + c on both halves.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\t"
      "mov.b64 rb, {%4, %5};\n\t"
      "mov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\t"
      "mov.b64 rb, {%4, %5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// Two packed 16-bit floats -> two fp32 (exact).
template <typename T>
__device__ __forceinline__ float2 unpack2(uint32_t u);
template <>
__device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t u) {
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
template <>
__device__ __forceinline__ float2 unpack2<__half>(uint32_t u) {
  return __half22float2(*reinterpret_cast<const __half2*>(&u));
}

template <typename T>
__device__ __forceinline__ T from_float(float x);
template <>
__device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float x) {
  return __float2bfloat16_rn(x);
}
template <>
__device__ __forceinline__ __half from_float<__half>(float x) {
  return __float2half_rn(x);
}
template <typename T>
__device__ __forceinline__ float to_float(T x);
template <>
__device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 x) {
  return __bfloat162float(x);
}
template <>
__device__ __forceinline__ float to_float<__half>(__half x) {
  return __half2float(x);
}

}  // namespace b200
