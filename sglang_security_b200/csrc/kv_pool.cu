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
