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

// Debug / test hook: every dispatcher records WHICH kernel instance it launched last on this host
// thread (tile shape, split factor, multicast, ...) so that parity tests can assert that a shape
// really reached the instance it is meant to cover (b200_last_variant()).
void note_variant(const char* fmt, ...);

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

// ---- launches ------------------------------------------------------------------------------
// Every kernel of the library is launched through launch_kernel(): when programmatic dependent
// launch is on (b200_set_pdl / B200_PDL, default on) the launch carries
// cudaLaunchAttributeProgrammaticStreamSerialization, so inside a stream or a captured CUDA graph
// the NEXT kernel's CTAs may start (and run their prologue: barrier init, TMEM allocation, weight
// prefetch) while this kernel drains.  Device side, a kernel must call griddep_wait() before it
// reads anything an earlier kernel wrote or writes anything an earlier kernel may still read, and
// calls griddep_launch_dependents() once its own dependent-data accesses may begin.
extern std::atomic<int> g_pdl_on;  // bit mask over PdlFamily (all bits = every kernel)
enum PdlFamily { kPdlGemmSmall = 1, kPdlGemmBig = 2, kPdlAttn = 4, kPdlMerge = 8, kPdlRow = 16,
                 kPdlRope = 32, kPdlQuant = 64, kPdlOther = 128 };

// Measured on the Llama-3.1-8B FP8 decode step (profiles/r01_pdl_sweep.md): overlapping launches
// pays while the step is latency-bound (-7..-10 % at batch <= 32) and costs 3-6 % at batch 64+, where
// early-resident dependents only take SM resources from a bandwidth-bound predecessor.  Call sites
// therefore pass pdl_family(tokens, family): the family bit for small token counts, 0 otherwise.
int pdl_max_tokens();
inline int pdl_family(int tokens, int family) { return tokens <= pdl_max_tokens() ? family : 0; }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t st, unsigned cluster_x, int family,
                                 Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (g_pdl_on.load(std::memory_order_relaxed) & family) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
int device_sm_count();

// ---- device helpers -----------------------------------------------------------------------
__host__ __device__ constexpr int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Programmatic dependent launch (no-ops when the kernel was launched without the attribute).
__device__ __forceinline__ void griddep_wait() {
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
}
__device__ __forceinline__ void griddep_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
}

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

// 16-bit float -> e4m3 byte with torch's `.to(torch.float8_e4m3fn)` semantics (round to nearest
// even, NO saturation: |x| beyond the last value that still rounds to 448, i.e. > 464, and NaN/Inf
// become NaN = sign | 0x7f), after the optional in-place `div_(scale)` of set_kv_buffer, whose
// result is rounded back to the 16-bit dtype first (memory_pool.py:390-396).
template <typename T>
__device__ __forceinline__ uint32_t to_e4m3_torch(T x, float inv_is_div_scale, bool has_scale) {
  float f = to_float<T>(x);
  if (has_scale) f = to_float<T>(from_float<T>(f / inv_is_div_scale));
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(0.f), "f"(f));
  uint32_t b = r & 0xffu;
  if (!(fabsf(f) <= 464.f)) b = ((__float_as_uint(f) >> 24) & 0x80u) | 0x7fu;
  return b;
}
// Same for an e5m2 pool (`--kv-cache-dtype fp8_e5m2`, model_runner.py:845-858): torch's
// `.to(torch.float8_e5m2)` is IEEE round-to-nearest-even WITH overflow to infinity (|x| >= 61440,
// the midpoint between the last finite value 57344 and 2^16, becomes sign | 0x7c) and NaN -> sign | 0x7f;
// the hardware conversion saturates to 57344 instead, so the two ranges are patched.
template <typename T>
__device__ __forceinline__ uint32_t to_e5m2_torch(T x, float inv_is_div_scale, bool has_scale) {
  float f = to_float<T>(x);
  if (has_scale) f = to_float<T>(from_float<T>(f / inv_is_div_scale));
  uint16_t r;
  asm("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(0.f), "f"(f));
  uint32_t b = r & 0xffu;
  const uint32_t sign = (__float_as_uint(f) >> 24) & 0x80u;
  if (!(fabsf(f) < 61440.f)) b = sign | (f != f ? 0x7fu : 0x7cu);
  return b;
}
template <typename T>
__device__ __forceinline__ uint32_t to_fp8_torch(T x, float scale, bool has_scale, bool e5m2) {
  return e5m2 ? to_e5m2_torch<T>(x, scale, has_scale) : to_e4m3_torch<T>(x, scale, has_scale);
}
// four fp8 bytes -> two f16x2 registers (exact): lo = bytes 0,1; hi = bytes 2,3
template <bool E5M2>
__device__ __forceinline__ void cvt_fp8x4_f16x4(uint32_t x, uint32_t& lo, uint32_t& hi) {
  if (E5M2) {  // e5m2 is the high byte of an f16
    asm("prmt.b32 %0, %1, 0, 0x1404;\n" : "=r"(lo) : "r"(x));
    asm("prmt.b32 %0, %1, 0, 0x3424;\n" : "=r"(hi) : "r"(x));
  } else {
    asm("{\n\t.reg .b16 a, b;\n\t"
        "mov.b32 {a, b}, %2;\n\t"
        "cvt.rn.f16x2.e4m3x2 %0, a;\n\t"
        "cvt.rn.f16x2.e4m3x2 %1, b;\n\t}"
        : "=r"(lo), "=r"(hi)
        : "r"(x));
  }
}

}  // namespace b200
