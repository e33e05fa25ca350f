// KV-pool side of the hot path: the cache-location write (set_kv_buffer) and the
// req_to_token -> flat kv_indices gather.  Both are pure byte/index movement and must be
// bit-exact with the reference (memory_pool.py:378-412, layers/attention/utils.py:5-41).
#include "common.cuh"

namespace b200 {

// One thread moves one VEC-byte chunk of one token row of K or V.
//   blockIdx.y == 0 -> K, 1 -> V.   Rows are [Hkv*D] elements = row_bytes contiguous bytes.
template <typename Vec>
__global__ void kv_write_kernel(const char* __restrict__ k, const char* __restrict__ v,
                                int64_t k_stride, int64_t v_stride, char* __restrict__ k_cache,
                                char* __restrict__ v_cache, const int64_t* __restrict__ loc,
                                int num_tokens, int k_row_bytes, int v_row_bytes) {
  griddep_wait();
  griddep_launch_dependents();
  const bool is_v = blockIdx.y != 0;
  const char* src = is_v ? v : k;
  char* dst = is_v ? v_cache : k_cache;
  const int64_t stride = is_v ? v_stride : k_stride;
  const int row_bytes = is_v ? v_row_bytes : k_row_bytes;
  const int chunks = row_bytes / static_cast<int>(sizeof(Vec));
  const int64_t total = static_cast<int64_t>(num_tokens) * chunks;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / chunks);
    const int c = static_cast<int>(i - static_cast<int64_t>(t) * chunks);
    const int64_t slot = __ldg(loc + t);
    const Vec val = *reinterpret_cast<const Vec*>(src + t * stride + c * sizeof(Vec));
    *reinterpret_cast<Vec*>(dst + slot * row_bytes + c * sizeof(Vec)) = val;
  }
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

// One thread quantises 8 consecutive elements of one token row of K or V (blockIdx.y = 0 / 1).
template <typename T>
__global__ void kv_write_fp8_kernel(const T* __restrict__ k, const T* __restrict__ v,
                                    int64_t k_stride, int64_t v_stride,
                                    uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                                    const int64_t* __restrict__ loc, int num_tokens, int row_elems,
                                    float k_scale, float v_scale, int has_scale) {
  griddep_wait();
  griddep_launch_dependents();
  const bool is_v = blockIdx.y != 0;
  const T* src = is_v ? v : k;
  uint8_t* dst = is_v ? v_cache : k_cache;
  const int64_t stride = is_v ? v_stride : k_stride;
  const float scale = is_v ? v_scale : k_scale;
  const int chunks = row_elems / 8;
  const int64_t total = static_cast<int64_t>(num_tokens) * chunks;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / chunks);
    const int c = static_cast<int>(i - static_cast<int64_t>(t) * chunks);
    const int64_t slot = __ldg(loc + t);
    const uint4 raw = *reinterpret_cast<const uint4*>(src + t * stride + c * 8);
    const T* e = reinterpret_cast<const T*>(&raw);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      lo |= to_e4m3_torch<T>(e[j], scale, has_scale != 0) << (8 * j);
      hi |= to_e4m3_torch<T>(e[4 + j], scale, has_scale != 0) << (8 * j);
    }
    *reinterpret_cast<uint2*>(dst + slot * row_elems + c * 8) = make_uint2(lo, hi);
  }
}

// One CTA per request: copy lens[b] int32 entries of its req_to_token row.
__global__ void kv_indices_kernel(const int32_t* __restrict__ req_to_token, int64_t stride,
                                  const int64_t* __restrict__ req_pool_indices,
                                  const int64_t* __restrict__ lens,
                                  const int32_t* __restrict__ kv_indptr,
                                  const int32_t* __restrict__ kv_start,
                                  int32_t* __restrict__ kv_indices) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x;
  const int64_t row = req_pool_indices[b];
  const int start = kv_start ? kv_start[b] : 0;
  const int n = static_cast<int>(lens[b]);
  const int32_t* src = req_to_token + row * stride + start;
  int32_t* dst = kv_indices + kv_indptr[b];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < n; i += gridDim.y * blockDim.x)
    dst[i] = __ldg(src + i);
}

}  // namespace b200

extern "C" {

int b200_kv_write(const void* k, const void* v, int64_t k_token_stride_bytes,
                  int64_t v_token_stride_bytes, void* k_cache, void* v_cache, const int64_t* loc,
                  int num_tokens, int k_row_bytes, int v_row_bytes, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return 0;
  B200_REQUIRE(k && v && k_cache && v_cache && loc, "b200_kv_write: null pointer");
  B200_REQUIRE(num_tokens > 0 && k_row_bytes > 0 && v_row_bytes > 0,
               "b200_kv_write: bad sizes (tokens=%d, row_bytes=%d/%d)", num_tokens, k_row_bytes,
               v_row_bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Pick the widest vector every pointer/stride/row length allows.
  uintptr_t bits = reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
                   reinterpret_cast<uintptr_t>(k_cache) | reinterpret_cast<uintptr_t>(v_cache) |
                   static_cast<uintptr_t>(k_token_stride_bytes) |
                   static_cast<uintptr_t>(v_token_stride_bytes) |
                   static_cast<uintptr_t>(k_row_bytes) | static_cast<uintptr_t>(v_row_bytes);
  const int max_row = k_row_bytes > v_row_bytes ? k_row_bytes : v_row_bytes;
  ScopedKernelTimer timer("kv_write", st);
  cudaError_t lerr = cudaSuccess;
  auto launch = [&](auto vec_tag) {
    using Vec = decltype(vec_tag);
    const int64_t total = static_cast<int64_t>(num_tokens) * (max_row / sizeof(Vec));
    const int threads = 256;
    int64_t blocks = (total + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    dim3 grid(static_cast<unsigned>(blocks), 2);
    lerr = launch_kernel(kv_write_kernel<Vec>, grid, dim3(threads), 0, st, 1, pdl_family(num_tokens, kPdlOther),
                         static_cast<const char*>(k), static_cast<const char*>(v),
                         k_token_stride_bytes, v_token_stride_bytes, static_cast<char*>(k_cache),
                         static_cast<char*>(v_cache), loc, num_tokens, k_row_bytes, v_row_bytes);
  };
  if ((bits & 15u) == 0) launch(uint4{});
  else if ((bits & 7u) == 0) launch(uint2{});
  else if ((bits & 3u) == 0) launch(uint32_t{});
  else if ((bits & 1u) == 0) launch(uint16_t{});
  else launch(uint8_t{});
  B200_CUDA_OK(lerr);
  B200_CHECK_LAUNCH("b200_kv_write");
  return 0;
}

int b200_kv_write_fp8(const void* k, const void* v, int64_t k_token_stride,
                      int64_t v_token_stride, void* k_cache, void* v_cache, const int64_t* loc,
                      int num_tokens, int row_elems, float k_scale, float v_scale, int has_scale,
                      int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return 0;
  B200_REQUIRE(k && v && k_cache && v_cache && loc, "b200_kv_write_fp8: null pointer");
  B200_REQUIRE(num_tokens > 0 && row_elems > 0 && row_elems % 8 == 0,
               "b200_kv_write_fp8: bad sizes (tokens=%d, row_elems=%d)", num_tokens, row_elems);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_kv_write_fp8: source dtype %d", dtype);
  B200_REQUIRE(aligned16(k) && aligned16(v) && k_token_stride % 8 == 0 && v_token_stride % 8 == 0 &&
                   (reinterpret_cast<uintptr_t>(k_cache) & 7u) == 0 &&
                   (reinterpret_cast<uintptr_t>(v_cache) & 7u) == 0,
               "b200_kv_write_fp8: k/v must be 16-byte aligned with strides %% 8 == 0, caches 8-byte "
               "aligned");
  B200_REQUIRE(!has_scale || (k_scale > 0.f && v_scale > 0.f),
               "b200_kv_write_fp8: scales must be positive");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t total = static_cast<int64_t>(num_tokens) * (row_elems / 8);
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  dim3 grid(static_cast<unsigned>(blocks), 2);
  ScopedKernelTimer timer("kv_write_fp8", st);
  cudaError_t lerr;
  if (dtype == B200_BF16)
    lerr = launch_kernel(kv_write_fp8_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, 1,
                         pdl_family(num_tokens, kPdlOther), static_cast<const __nv_bfloat16*>(k),
                         static_cast<const __nv_bfloat16*>(v), k_token_stride, v_token_stride,
                         static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), loc,
                         num_tokens, row_elems, k_scale, v_scale, has_scale);
  else
    lerr = launch_kernel(kv_write_fp8_kernel<__half>, grid, dim3(256), 0, st, 1,
                         pdl_family(num_tokens, kPdlOther), static_cast<const __half*>(k),
                         static_cast<const __half*>(v), k_token_stride, v_token_stride,
                         static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), loc,
                         num_tokens, row_elems, k_scale, v_scale, has_scale);
  B200_CUDA_OK(lerr);
  B200_CHECK_LAUNCH("b200_kv_write_fp8");
  return 0;
}

int b200_kv_indices(const int32_t* req_to_token, int64_t req_to_token_stride,
                    const int64_t* req_pool_indices, const int64_t* lens,
                    const int32_t* kv_indptr, const int32_t* kv_start, int32_t* kv_indices, int bs,
                    void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(req_to_token && req_pool_indices && lens && kv_indptr && kv_indices,
               "b200_kv_indices: null pointer");
  B200_REQUIRE(bs > 0, "b200_kv_indices: bad batch size %d", bs);
  // Few requests: spread each row over several CTAs so a lone long request still fills SMs.
  int per_req = bs >= 512 ? 1 : (bs >= 64 ? 4 : 16);
  dim3 grid(static_cast<unsigned>(bs), static_cast<unsigned>(per_req));
  B200_CUDA_OK(launch_kernel(kv_indices_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream),
                             1, pdl_family(bs, kPdlOther), req_to_token, req_to_token_stride, req_pool_indices, lens, kv_indptr,
                             kv_start, kv_indices));
  B200_CHECK_LAUNCH("b200_kv_indices");
  return 0;
}

}  // extern "C"
