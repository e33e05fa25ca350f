// Scheduler-side index kernels around the attention hot path: pure int64 /
// int32 index arithmetic, bit-exact with the reference's Triton kernels.
//   b200_write_req_to_token  write_req_to_token_pool_triton   managers/schedule_batch.py:1789-1823
//   b200_get_last_loc        get_last_loc_kernel / _torch     managers/schedule_batch.py:1838-1893
//   b200_alloc_extend        alloc_extend_kernel              mem_cache/paged_allocator.py:29-115
//   b200_alloc_decode        alloc_decode_kernel              mem_cache/paged_allocator.py:118-155
//   b200_compute_position    compute_position_kernel          model_executor/forward_batch_info.py:639-692
// The reference computes every request's exclusive prefix sums with an O(bs) scalar loop per
// program ("can be slow for large bs"); here one CTA per request reduces them cooperatively.
#include "common.cuh"

namespace b200 {

constexpr int kIdxThreads = 256;

// Sum of f(i) for i < n over the CTA (all threads get the result).
template <typename F>
__device__ __forceinline__ int64_t cta_prefix_sum(int n, F f) {
  __shared__ int64_t warp_part[kIdxThreads / 32];
  __shared__ int64_t total;
  int64_t v = 0;
  for (int i = threadIdx.x; i < n; i += kIdxThreads) v += f(i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();  // protects warp_part / total against the previous call
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    int64_t t = 0;
#pragma unroll
    for (int w = 0; w < kIdxThreads / 32; ++w) t += warp_part[w];
    total = t;
  }
  __syncthreads();
  return total;
}

__global__ void __launch_bounds__(kIdxThreads)
write_req_to_token_kernel(int32_t* __restrict__ req_to_token, int64_t stride,
                          const int64_t* __restrict__ req_pool_indices,
                          const int64_t* __restrict__ pre_lens, const int64_t* __restrict__ seq_lens,
                          const int64_t* __restrict__ extend_lens,
                          const int64_t* __restrict__ out_cache_loc) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x;
  const int64_t start = cta_prefix_sum(b, [&](int i) { return extend_lens[i]; });
  const int64_t pre = pre_lens[b];
  const int64_t n = seq_lens[b] - pre;
  int32_t* row = req_to_token + req_pool_indices[b] * stride + pre;
  for (int64_t i = threadIdx.x; i < n; i += kIdxThreads)
    row[i] = static_cast<int32_t>(out_cache_loc[start + i]);
}

__global__ void get_last_loc_kernel(const int32_t* __restrict__ req_to_token, int64_t stride,
                                    const int64_t* __restrict__ req_pool_indices,
                                    const int64_t* __restrict__ prefix_lens,
                                    int64_t* __restrict__ result, int n) {
  griddep_wait();
  griddep_launch_dependents();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t pl = prefix_lens[i];
  result[i] = pl > 0 ? static_cast<int64_t>(req_to_token[req_pool_indices[i] * stride + pl - 1])
                     : static_cast<int64_t>(-1);
}

__device__ __forceinline__ int64_t pages_of(int64_t len, int64_t page) {
  return (len + page - 1) / page;
}

__global__ void __launch_bounds__(kIdxThreads)
alloc_extend_kernel(const int64_t* __restrict__ pre_lens, const int64_t* __restrict__ seq_lens,
                    const int64_t* __restrict__ last_loc, const int64_t* __restrict__ free_pages,
                    int64_t* __restrict__ out_indices, int64_t* __restrict__ ret_values,
                    int64_t page) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x;
  const int64_t out_start = cta_prefix_sum(b, [&](int i) { return seq_lens[i] - pre_lens[i]; });
  const int64_t page_start = cta_prefix_sum(
      b, [&](int i) { return pages_of(seq_lens[i], page) - pages_of(pre_lens[i], page); });
  const int64_t seq = seq_lens[b], pre = pre_lens[b];
  const int64_t my_pages = pages_of(seq, page) - pages_of(pre, page);
  if (b == static_cast<int>(gridDim.x) - 1 && threadIdx.x == 0)
    *ret_values = ((page_start + my_pages) << 32) | (out_start + (seq - pre));
  // part 1: finish the partially filled last page of the prefix
  const int64_t pre_up = pages_of(pre, page) * page;
  const int64_t n1 = (seq < pre_up ? seq : pre_up) - pre;
  const int64_t ll = last_loc[b];
  int64_t* out = out_indices + out_start;
  for (int64_t i = threadIdx.x; i < n1; i += kIdxThreads) out[i] = ll + 1 + i;
  if (pre + n1 == seq) return;
  // part 2: whole new pages
  const int64_t n2 = seq / page * page - pre_up;
  for (int64_t i = threadIdx.x; i < n2; i += kIdxThreads)
    out[n1 + i] = free_pages[page_start + i / page] * page + i % page;
  if (pre + n1 + n2 == seq) return;
  // part 3: the new partial page
  const int64_t n3 = seq - seq / page * page;
  const int64_t last_page = free_pages[page_start + my_pages - 1];
  for (int64_t i = threadIdx.x; i < n3; i += kIdxThreads) out[n1 + n2 + i] = last_page * page + i;
}

__global__ void __launch_bounds__(kIdxThreads)
alloc_decode_kernel(const int64_t* __restrict__ seq_lens, const int64_t* __restrict__ last_loc,
                    const int64_t* __restrict__ free_pages, int64_t* __restrict__ out_indices,
                    int64_t* __restrict__ ret_values, int64_t page) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x;
  const int64_t page_start = cta_prefix_sum(
      b, [&](int i) { return pages_of(seq_lens[i], page) - pages_of(seq_lens[i] - 1, page); });
  const int64_t seq = seq_lens[b];
  const int64_t mine = pages_of(seq, page) - pages_of(seq - 1, page);
  if (threadIdx.x != 0) return;
  if (b == static_cast<int>(gridDim.x) - 1) *ret_values = page_start + mine;
  out_indices[b] = mine == 0 ? last_loc[b] + 1 : free_pages[page_start] * page;
}

__global__ void __launch_bounds__(kIdxThreads)
compute_position_kernel(int64_t* __restrict__ positions, int32_t* __restrict__ extend_start_loc,
                        const int32_t* __restrict__ extend_prefix_lens,
                        const int32_t* __restrict__ extend_seq_lens) {
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x;
  const int64_t start =
      cta_prefix_sum(b, [&](int i) { return static_cast<int64_t>(extend_seq_lens[i]); });
  const int64_t prefix = extend_prefix_lens ? extend_prefix_lens[b] : 0;
  const int64_t n = extend_seq_lens[b];
  for (int64_t i = threadIdx.x; i < n; i += kIdxThreads) positions[start + i] = prefix + i;
  if (threadIdx.x == 0) extend_start_loc[b] = static_cast<int32_t>(start);
}

}  // namespace b200

extern "C" {

int b200_write_req_to_token(int32_t* req_to_token, int64_t req_to_token_stride,
                            const int64_t* req_pool_indices, const int64_t* pre_lens,
                            const int64_t* seq_lens, const int64_t* extend_lens,
                            const int64_t* out_cache_loc, int bs, void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(req_to_token && req_pool_indices && pre_lens && seq_lens && extend_lens &&
                   out_cache_loc && bs > 0,
               "b200_write_req_to_token: bad arguments");
  B200_CUDA_OK(launch_kernel(write_req_to_token_kernel, dim3(bs), dim3(kIdxThreads), 0,
                             static_cast<cudaStream_t>(stream), 1, pdl_family(bs, kPdlOther),
                             req_to_token, req_to_token_stride, req_pool_indices, pre_lens, seq_lens,
                             extend_lens, out_cache_loc));
  B200_CHECK_LAUNCH("write_req_to_token_kernel");
  return 0;
}

int b200_get_last_loc(const int32_t* req_to_token, int64_t req_to_token_stride,
                      const int64_t* req_pool_indices, const int64_t* prefix_lens, int64_t* result,
                      int n, void* stream) {
  using namespace b200;
  if (n == 0) return 0;
  B200_REQUIRE(req_to_token && req_pool_indices && prefix_lens && result && n > 0,
               "b200_get_last_loc: bad arguments");
  B200_CUDA_OK(launch_kernel(get_last_loc_kernel, dim3((n + 255) / 256), dim3(256), 0,
                             static_cast<cudaStream_t>(stream), 1, pdl_family(n, kPdlOther),
                             req_to_token, req_to_token_stride, req_pool_indices, prefix_lens, result,
                             n));
  B200_CHECK_LAUNCH("get_last_loc_kernel");
  return 0;
}

int b200_alloc_extend(const int64_t* pre_lens, const int64_t* seq_lens, const int64_t* last_loc,
                      const int64_t* free_pages, int64_t* out_indices, int64_t* ret_values, int bs,
                      int page_size, void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(pre_lens && seq_lens && last_loc && free_pages && out_indices && ret_values &&
                   bs > 0 && page_size > 0,
               "b200_alloc_extend: bad arguments");
  B200_CUDA_OK(launch_kernel(alloc_extend_kernel, dim3(bs), dim3(kIdxThreads), 0,
                             static_cast<cudaStream_t>(stream), 1, 0, pre_lens, seq_lens, last_loc,
                             free_pages, out_indices, ret_values, static_cast<int64_t>(page_size)));
  B200_CHECK_LAUNCH("alloc_extend_kernel");
  return 0;
}

int b200_alloc_decode(const int64_t* seq_lens, const int64_t* last_loc, const int64_t* free_pages,
                      int64_t* out_indices, int64_t* ret_values, int bs, int page_size,
                      void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(seq_lens && last_loc && free_pages && out_indices && ret_values && bs > 0 &&
                   page_size > 0,
               "b200_alloc_decode: bad arguments");
  B200_CUDA_OK(launch_kernel(alloc_decode_kernel, dim3(bs), dim3(kIdxThreads), 0,
                             static_cast<cudaStream_t>(stream), 1, 0, seq_lens, last_loc, free_pages,
                             out_indices, ret_values, static_cast<int64_t>(page_size)));
  B200_CHECK_LAUNCH("alloc_decode_kernel");
  return 0;
}

int b200_compute_position(int64_t* positions, int32_t* extend_start_loc,
                          const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens, int bs,
                          void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(positions && extend_start_loc && extend_seq_lens && bs > 0,
               "b200_compute_position: bad arguments");
  B200_CUDA_OK(launch_kernel(compute_position_kernel, dim3(bs), dim3(kIdxThreads), 0,
                             static_cast<cudaStream_t>(stream), 1, pdl_family(bs, kPdlOther),
                             positions, extend_start_loc, extend_prefix_lens, extend_seq_lens));
  B200_CHECK_LAUNCH("compute_position_kernel");
  return 0;
}

}  // extern "C"
