// FP8 (e4m3) activation quantisers of the W8A8 linear path.
//   per-token      : scale[m] = absmax(x[m,:]) / 448 ; q = sat(x * (1/scale))       (reference
//                    sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:9-76)
//   per-token-group: scale[m,g] = max(absmax, eps) / 448 ; q = sat(x / scale), group = 128
//                    (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:19-98,
//                     native_per_token_group_quant_fp8, python/sglang/test/test_block_fp8.py:23-48)
//   static         : q = sat(x * (1/scale)) with one given scale             (fp8_kernel.py:336-376)
// The arithmetic (reciprocal-multiply vs. true division, clamp order, RN conversion) follows the
// reference kernels so that the quantised bytes agree.
#include <cuda_fp8.h>

#include "common.cuh"

namespace b200 {

constexpr float kFp8Max = 448.0f;

__device__ __forceinline__ uint8_t to_e4m3(float v) {
  return static_cast<uint8_t>(__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3));
}

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&f)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 v = unpack2<T>(w[e]);
    f[2 * e] = v.x;
    f[2 * e + 1] = v.y;
  }
}

__device__ __forceinline__ float block_max(float v, float* red) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = lane < nw ? red[lane] : 0.f;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

// One CTA per token row.  K % 8 == 0.
// MODE 0: dynamic, q = sat(x * (1/scale))           (per_token_quant_fp8.cu)
// MODE 1: static scale, q = sat(x * (1/scale))      (_static_quant_fp8)
// MODE 2: dynamic, scale = max(absmax, eps)/448, q = sat(x / scale): the group_size == K case of
//         per_token_group_quant_fp8 used for load-time per-channel weight quantisation.
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
fp8_quant_per_token_kernel(const T* __restrict__ x, int64_t ldx, uint8_t* __restrict__ q,
                           int64_t ldq, float* __restrict__ scales,
                           const float* __restrict__ static_scale, int K, float eps) {
  constexpr bool STATIC = MODE == 1;
  __shared__ float red[32];
  griddep_wait();
  // MODE 2 quantises WEIGHTS at load time: its consumers (the GEMMs) prefetch their weight operand
  // before their own dependency wait, so this kernel must not release them early
  if (MODE != 2) griddep_launch_dependents();
  const int m = blockIdx.x;
  const T* row = x + m * ldx;
  uint8_t* out = q + m * ldq;
  const int nvec = K / 8;
  float scale;
  if (STATIC) {
    scale = __ldg(static_scale);
  } else {
    float amax = 0.f;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      float f[8];
      load8(row + i * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(f[e]));
    }
    amax = block_max(amax, red);
    if (MODE == 2) amax = fmaxf(amax, eps);
    scale = amax / kFp8Max;
    if (threadIdx.x == 0) scales[m] = scale;
  }
  // an all-zero row keeps scale 0 and quantises to zeros (the reference would emit NaN bytes)
  const float inv = scale > 0.f ? 1.0f / scale : 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    float f[8];
    load8(row + i * 8, f);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a = MODE == 2 ? f[e] / scale : f[e] * inv;
      const float b = MODE == 2 ? f[4 + e] / scale : f[4 + e] * inv;
      lo |= static_cast<uint32_t>(to_e4m3(fmaxf(fminf(a, kFp8Max), -kFp8Max))) << (8 * e);
      hi |= static_cast<uint32_t>(to_e4m3(fmaxf(fminf(b, kFp8Max), -kFp8Max))) << (8 * e);
    }
    *reinterpret_cast<uint2*>(out + i * 8) = make_uint2(lo, hi);
  }
}

// 16 lanes per 128-element group (8 elements per lane), 2 groups per warp, 16 groups per CTA.
// scales[m * s_row + g * s_col]: row-major (s_row = K/128, s_col = 1) or column-major / TMA
// aligned (s_row = 1, s_col = padded M).
template <typename T>
__global__ void __launch_bounds__(256)
fp8_quant_group128_kernel(const T* __restrict__ x, int64_t ldx, uint8_t* __restrict__ q,
                          int64_t ldq, float* __restrict__ scales, int64_t s_row, int64_t s_col,
                          int M, int groups_per_row, float eps) {
  griddep_wait();
  griddep_launch_dependents();
  const int64_t gid = static_cast<int64_t>(blockIdx.x) * 16 + (threadIdx.x >> 4);
  const int lane = threadIdx.x & 15;
  const int64_t total = static_cast<int64_t>(M) * groups_per_row;
  const bool active = gid < total;
  const int m = active ? static_cast<int>(gid / groups_per_row) : 0;
  const int g = active ? static_cast<int>(gid - static_cast<int64_t>(m) * groups_per_row) : 0;
  float f[8];
  if (active) {
    load8(x + m * ldx + g * 128 + lane * 8, f);
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = 0.f;
  }
  float amax = eps;
#pragma unroll
  for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(f[e]));
#pragma unroll
  for (int off = 8; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if (!active) return;
  const float s = amax / kFp8Max;
  if (lane == 0) scales[m * s_row + g * s_col] = s;
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    lo |= static_cast<uint32_t>(to_e4m3(fminf(fmaxf(f[e] / s, -kFp8Max), kFp8Max))) << (8 * e);
    hi |= static_cast<uint32_t>(to_e4m3(fminf(fmaxf(f[4 + e] / s, -kFp8Max), kFp8Max))) << (8 * e);
  }
  *reinterpret_cast<uint2*>(q + m * ldq + g * 128 + lane * 8) = make_uint2(lo, hi);
}

}  // namespace b200

extern "C" {

int b200_fp8_quant_per_token(const void* x, int64_t ldx, void* q, int64_t ldq, float* scales,
                             const float* static_scale, int M, int K, int use_division, float eps,
                             int dtype, void* stream) {
  using namespace b200;
  if (M == 0) return 0;
  B200_REQUIRE(x && q && (scales || static_scale), "b200_fp8_quant_per_token: null pointer");
  B200_REQUIRE(K > 0 && K % 8 == 0 && ldx % 8 == 0 && ldq % 8 == 0 && aligned16(x) &&
                   (reinterpret_cast<uintptr_t>(q) & 7u) == 0,
               "b200_fp8_quant_per_token: K/ld must be multiples of 8 and pointers aligned (K=%d)",
               K);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_fp8_quant_per_token: dtype %d",
               dtype);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ScopedKernelTimer timer("fp8_quant", st);
  const int threads = K >= 2048 ? 256 : 128;
#define LAUNCH(T, MODE)                                                                          \
  B200_CUDA_OK(launch_kernel(fp8_quant_per_token_kernel<T, MODE>, dim3(M), dim3(threads), 0, st, 1, pdl_family(M, kPdlQuant), \
                             static_cast<const T*>(x), ldx, static_cast<uint8_t*>(q), ldq, scales,  \
                             static_scale, K, eps))
  if (dtype == B200_BF16) {
    if (static_scale) LAUNCH(__nv_bfloat16, 1);
    else if (use_division) LAUNCH(__nv_bfloat16, 2);
    else LAUNCH(__nv_bfloat16, 0);
  } else {
    if (static_scale) LAUNCH(__half, 1);
    else if (use_division) LAUNCH(__half, 2);
    else LAUNCH(__half, 0);
  }
#undef LAUNCH
  B200_CHECK_LAUNCH("fp8_quant_per_token_kernel");
  return 0;
}

int b200_fp8_quant_group128(const void* x, int64_t ldx, void* q, int64_t ldq, float* scales,
                            int64_t scale_row_stride, int64_t scale_col_stride, int M, int K,
                            float eps, int dtype, void* stream) {
  using namespace b200;
  if (M == 0) return 0;
  B200_REQUIRE(x && q && scales, "b200_fp8_quant_group128: null pointer");
  B200_REQUIRE(K > 0 && K % 128 == 0 && ldx % 8 == 0 && ldq % 8 == 0 && aligned16(x) &&
                   (reinterpret_cast<uintptr_t>(q) & 7u) == 0,
               "b200_fp8_quant_group128: K must be a multiple of 128 (K=%d)", K);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_fp8_quant_group128: dtype %d", dtype);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ScopedKernelTimer timer("fp8_quant", st);
  const int gpr = K / 128;
  const int64_t total = static_cast<int64_t>(M) * gpr;
  const unsigned blocks = static_cast<unsigned>((total + 15) / 16);
  if (dtype == B200_BF16)
    B200_CUDA_OK(launch_kernel(fp8_quant_group128_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0,
                               st, 1, pdl_family(M, kPdlQuant), static_cast<const __nv_bfloat16*>(x), ldx,
                               static_cast<uint8_t*>(q), ldq, scales, scale_row_stride,
                               scale_col_stride, M, gpr, eps));
  else
    B200_CUDA_OK(launch_kernel(fp8_quant_group128_kernel<__half>, dim3(blocks), dim3(256), 0, st, 1, pdl_family(M, kPdlQuant),
                               static_cast<const __half*>(x), ldx, static_cast<uint8_t*>(q), ldq,
                               scales, scale_row_stride, scale_col_stride, M, gpr, eps));
  B200_CHECK_LAUNCH("fp8_quant_group128_kernel");
  return 0;
}

}  // extern "C"
