// Paged split-KV decode attention whose split-KV scratch lives in the library itself — for callers
// that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
// Existing entry points and kernels are untouched: this file only owns the scratch
// array, resolves its device address and hands it to b200_attn_decode as the workspace.
//
// The scratch is ONE file-scope array per GPU that loads the library (held until process exit),
// B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES long.  It holds the fp32 partial outputs and log-sum-exps
// of the splits, [bs, Hq, num_splits, Dv] + [bs, Hq, num_splits].  Nothing in it has to start at
// zero: every non-empty split writes its partials before the merge kernel (same stream) reads them,
// and the merge never reads an empty split, so the previous call's contents are harmless.  All
// calls on a device share the array: two calls that overlap on the GPU (different streams, or
// concurrent graph launches) race on it — serialise them on one stream.
#include <mutex>

#include "common.cuh"

namespace b200 {

// In host code this name is only a host-side stand-in for the symbol, never a device address: the
// address a kernel may use comes from cudaGetSymbolAddress (scratch_for_current_device below).
static __device__ __align__(256) unsigned char g_decode_scratch[B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static std::mutex g_scratch_mu;
static void* g_scratch_ptr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device, resolved once per device while that
// device is current (cudaGetSymbolAddress answers for the current device's copy of the symbol).
static int scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "b200_attn_decode_noalloc: device ordinal %d out of range",
               dev);
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void* ptr = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&ptr, g_decode_scratch));
    B200_REQUIRE(ptr != nullptr, "b200_attn_decode_noalloc: cudaGetSymbolAddress returned null");
    g_scratch_ptr[dev] = ptr;
  }
  *out = g_scratch_ptr[dev];
  return 0;
}

}  // namespace b200

extern "C" {

int b200_attn_decode_noalloc(const void* q, int64_t q_stride, const void* k_cache,
                             const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                             const int32_t* req_to_token, int64_t req_to_token_stride,
                             const int64_t* req_pool_indices, const int64_t* seq_lens, int bs,
                             int Hq, int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                             int num_splits, void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(bs > 0 && Hq > 0 && Dv > 0 && num_splits >= 1 && num_splits <= 1024,
               "b200_attn_decode_noalloc: bad sizes bs=%d Hq=%d Dv=%d num_splits=%d", bs, Hq, Dv,
               num_splits);
  // in 64-bit, before anything is narrowed: the request must fit the built-in scratch
  const unsigned long long rows =
      static_cast<unsigned long long>(bs) * static_cast<unsigned long long>(Hq) * num_splits;
  const unsigned long long need = num_splits > 1 ? rows * (Dv + 1ull) * sizeof(float) : 0ull;
  B200_REQUIRE(need <= B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES,
               "b200_attn_decode_noalloc: bs=%d Hq=%d Dv=%d num_splits=%d needs %llu bytes of split-KV "
               "scratch, the built-in scratch holds %llu (use fewer splits, a smaller batch, or "
               "b200_attn_decode with your own workspace)",
               bs, Hq, Dv, num_splits, need,
               static_cast<unsigned long long>(B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES));
  void* scratch = nullptr;
  if (int rc = scratch_for_current_device(&scratch)) return rc;
  return b200_attn_decode(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride, req_to_token,
                          req_to_token_stride, req_pool_indices, seq_lens, bs, Hq, Hkv, D, Dv, scale,
                          logit_cap, dtype, scratch, B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES, num_splits,
                          stream);
}

}  // extern "C"
