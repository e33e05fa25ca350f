// KV-pool side of the hot path: the cache-location write (set_kv_buffer) and the
// req_to_token -> flat kv_indices gather.  Both are pure byte/index movement and must be
// bit-exact with the reference (memory_pool.py:378-412, layers/attention/utils.py:5-41).
#include "common.cuh"

#include <type_traits>

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

// to_e4m3_torch (common.cuh): torch's 16-bit -> float8_e4m3fn conversion, shared with the fused
// RoPE + KV write kernel.
// One thread quantises 8 consecutive elements of one token row of K or V (blockIdx.y = 0 / 1).
template <typename T>
__global__ void kv_write_fp8_kernel(const T* __restrict__ k, const T* __restrict__ v,
                                    int64_t k_stride, int64_t v_stride,
                                    uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                                    const int64_t* __restrict__ loc, int num_tokens, int row_elems,
                                    float k_scale, float v_scale, int has_scale, int e5m2) {
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
      lo |= to_fp8_torch<T>(e[j], scale, has_scale != 0, e5m2 != 0) << (8 * j);
      hi |= to_fp8_torch<T>(e[4 + j], scale, has_scale != 0, e5m2 != 0) << (8 * j);
    }
    *reinterpret_cast<uint2*>(dst + slot * row_elems + c * 8) = make_uint2(lo, hi);
  }
}

// e4m3 pool -> dense 16-bit rows of one request (exact: every e4m3 value is a bf16 and an f16
// value), plus the identity page table the extend kernel will index the dense copy with.
// grid (chunks of tokens, bs); one thread converts 8 bytes of one row.
template <typename T>
__global__ void kv_dequant_gather_kernel(const uint8_t* __restrict__ k8, const uint8_t* __restrict__ v8,
                                         int row_elems, const int32_t* __restrict__ req_to_token,
                                         int64_t rtt_stride,
                                         const int64_t* __restrict__ req_pool_indices,
                                         const int64_t* __restrict__ seq_lens,
                                         const int32_t* __restrict__ kv_indptr, T* __restrict__ k_out,
                                         T* __restrict__ v_out, int32_t* __restrict__ dense_rtt,
                                         int64_t dense_rtt_stride, int e5m2) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.y;
  const int len = static_cast<int>(seq_lens[b]);
  const int base = kv_indptr[b];
  const int32_t* rtt = req_to_token + req_pool_indices[b] * rtt_stride;
  const int chunks = row_elems / 8;
  const int64_t total = static_cast<int64_t>(len) * chunks;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / chunks);
    const int c = static_cast<int>(i - static_cast<int64_t>(t) * chunks);
    const int64_t slot = __ldg(rtt + t);
    if (c == 0) dense_rtt[b * dense_rtt_stride + t] = base + t;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      const uint8_t* src = (which ? v8 : k8) + slot * row_elems + c * 8;
      T* dst = (which ? v_out : k_out) + static_cast<int64_t>(base + t) * row_elems + c * 8;
      const uint2 raw = *reinterpret_cast<const uint2*>(src);
      const uint32_t w[2] = {raw.x, raw.y};
      uint32_t o[4];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t lo, hi;
        if (e5m2) cvt_fp8x4_f16x4<true>(w[h], lo, hi);
        else cvt_fp8x4_f16x4<false>(w[h], lo, hi);
        if (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) {
          const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&lo));
          const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&hi));
          const __nv_bfloat162 b0 = __floats2bfloat162_rn(f0.x, f0.y);
          const __nv_bfloat162 b1 = __floats2bfloat162_rn(f1.x, f1.y);
          lo = *reinterpret_cast<const uint32_t*>(&b0);
          hi = *reinterpret_cast<const uint32_t*>(&b1);
        }
        o[2 * h] = lo;
        o[2 * h + 1] = hi;
      }
      *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
    }
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
  const int e5m2 = (dtype & B200_KV_FP8_E5M2) ? 1 : 0;
  dtype &= ~B200_KV_FP8_E5M2;
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
                         num_tokens, row_elems, k_scale, v_scale, has_scale, e5m2);
  else
    lerr = launch_kernel(kv_write_fp8_kernel<__half>, grid, dim3(256), 0, st, 1,
                         pdl_family(num_tokens, kPdlOther), static_cast<const __half*>(k),
                         static_cast<const __half*>(v), k_token_stride, v_token_stride,
                         static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), loc,
                         num_tokens, row_elems, k_scale, v_scale, has_scale, e5m2);
  B200_CUDA_OK(lerr);
  B200_CHECK_LAUNCH("b200_kv_write_fp8");
  return 0;
}

int b200_kv_dequant_gather(const void* k_cache, const void* v_cache, int row_elems,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens,
                           const int32_t* kv_indptr, void* k_out, void* v_out, int32_t* dense_rtt,
                           int64_t dense_rtt_stride, int bs, int64_t max_len, int dtype,
                           void* stream) {
  using namespace b200;
  if (bs == 0 || max_len == 0) return 0;
  B200_REQUIRE(k_cache && v_cache && req_to_token && req_pool_indices && seq_lens && kv_indptr &&
                   k_out && v_out && dense_rtt,
               "b200_kv_dequant_gather: null pointer");
  B200_REQUIRE(row_elems > 0 && row_elems % 8 == 0 && bs > 0 && bs <= 65535,
               "b200_kv_dequant_gather: bad sizes (row_elems=%d bs=%d)", row_elems, bs);
  const int e5m2 = (dtype & B200_KV_FP8_E5M2) ? 1 : 0;
  dtype &= ~B200_KV_FP8_E5M2;
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_kv_dequant_gather: dtype %d", dtype);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(k_cache) & 7u) == 0 &&
                   (reinterpret_cast<uintptr_t>(v_cache) & 7u) == 0 && aligned16(k_out) &&
                   aligned16(v_out),
               "b200_kv_dequant_gather: misaligned pointers");
  const int64_t per_req = max_len * (row_elems / 8);
  int64_t bx = (per_req + 255) / 256;
  const int64_t cap = (148 * 16 + bs - 1) / bs;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  dim3 grid(static_cast<unsigned>(bx), static_cast<unsigned>(bs));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ScopedKernelTimer timer("kv_dequant_gather", st);
  cudaError_t lerr;
  if (dtype == B200_BF16)
    lerr = launch_kernel(kv_dequant_gather_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, 1,
                         kPdlOther & 0, static_cast<const uint8_t*>(k_cache),
                         static_cast<const uint8_t*>(v_cache), row_elems, req_to_token,
                         req_to_token_stride, req_pool_indices, seq_lens, kv_indptr,
                         static_cast<__nv_bfloat16*>(k_out), static_cast<__nv_bfloat16*>(v_out),
                         dense_rtt, dense_rtt_stride, e5m2);
  else
    lerr = launch_kernel(kv_dequant_gather_kernel<__half>, grid, dim3(256), 0, st, 1, kPdlOther & 0,
                         static_cast<const uint8_t*>(k_cache), static_cast<const uint8_t*>(v_cache),
                         row_elems, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                         kv_indptr, static_cast<__half*>(k_out), static_cast<__half*>(v_out),
                         dense_rtt, dense_rtt_stride, e5m2);
  B200_CUDA_OK(lerr);
  B200_CHECK_LAUNCH("b200_kv_dequant_gather");
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
