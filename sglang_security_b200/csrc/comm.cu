// One-shot all-reduce over NVLink peer memory + the IPC plumbing of the communicator (comm.cuh).
// Replaces tensor_model_parallel_all_reduce's custom path for decode-sized messages
// (python/sglang/srt/distributed/parallel_state.py:398-461 -> custom_all_reduce.py:452-498 ->
// sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-225); larger messages stay on NCCL.
#include "comm.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {

template <typename T>
__device__ __forceinline__ void accum8(float (&acc)[8], const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = unpack2<T>(w[i]);
    acc[2 * i] += f.x;
    acc[2 * i + 1] += f.y;
  }
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&acc)[8]);
template <>
__device__ __forceinline__ uint4 pack8<__nv_bfloat16>(const float (&acc)[8]) {
  uint4 o;
  uint32_t* w = &o.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t*>(&h);
  }
  return o;
}
template <>
__device__ __forceinline__ uint4 pack8<__half>(const float (&acc)[8]) {
  uint4 o;
  uint32_t* w = &o.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half2 h = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t*>(&h);
  }
  return o;
}

// T = bf16 / f16: 8 elements per 16-byte vector; float: 4 per vector (kF32 = true).
template <typename T, bool kF32>
__global__ void __launch_bounds__(256)
all_reduce_oneshot_kernel(const CommDev c, const uint4* __restrict__ in, uint4* __restrict__ out,
                          int64_t num_vec, int vec_per_cta) {
  griddep_wait();
  griddep_launch_dependents();
  const uint32_t e = comm_begin(c);
  const int cta = blockIdx.x;
  const int64_t v0 = static_cast<int64_t>(cta) * vec_per_cta;
  const int64_t v1 = min(num_vec, v0 + vec_per_cta);
  // push: my part of the input into slot [my rank] of every peer (start at rank+1 to spread links)
  for (int64_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
    const uint4 x = in[v];
    for (int i = 1; i < c.world; ++i) {
      int p = c.rank + i;
      if (p >= c.world) p -= c.world;
      reinterpret_cast<uint4*>(comm_slot(c, p, c.rank, e))[v] = x;
    }
  }
  __syncthreads();
  if (threadIdx.x < c.world) {
    comm_signal(c, e, cta, threadIdx.x);
    comm_wait(c, e, cta, threadIdx.x);
  }
  __syncthreads();
  // reduce in rank order (bit-identical on every rank)
  for (int64_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
    if constexpr (kF32) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < c.world; ++r) {
        const uint4 x = (r == c.rank)
                            ? in[v]
                            : __ldcg(reinterpret_cast<const uint4*>(comm_slot(c, c.rank, r, e)) + v);
        acc.x += __uint_as_float(x.x);
        acc.y += __uint_as_float(x.y);
        acc.z += __uint_as_float(x.z);
        acc.w += __uint_as_float(x.w);
      }
      out[v] = make_uint4(__float_as_uint(acc.x), __float_as_uint(acc.y), __float_as_uint(acc.z),
                          __float_as_uint(acc.w));
    } else {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int r = 0; r < c.world; ++r) {
        const uint4 x = (r == c.rank)
                            ? in[v]
                            : __ldcg(reinterpret_cast<const uint4*>(comm_slot(c, c.rank, r, e)) + v);
        accum8<T>(acc, x);
      }
      out[v] = pack8<T>(acc);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) comm_end(c, e, gridDim.x);
}

// LL variant: every 16-byte input vector travels as four 8-byte {payload, epoch} words; the
// receiver polls the words themselves (comm.cuh).  No barrier, no fence: one NVLink trip.
template <typename T, bool kF32>
__global__ void __launch_bounds__(128)
all_reduce_ll_kernel(const CommDev c, const uint4* __restrict__ in, uint4* __restrict__ out,
                     int64_t num_vec) {
  griddep_wait();
  griddep_launch_dependents();
  const uint32_t e = comm_begin(c);
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t first = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (int64_t v = first; v < num_vec; v += stride) {
    const uint4 x = in[v];
    for (int i = 1; i < c.world; ++i) {
      int p = c.rank + i;
      if (p >= c.world) p -= c.world;
      uint2* dst = comm_ll_slot(c, p, c.rank, e) + 4 * v;
      ll_store2(dst, x.x, x.y, e);
      ll_store2(dst + 2, x.z, x.w, e);
    }
  }
  for (int64_t v = first; v < num_vec; v += stride) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < c.world; ++r) {
      uint4 x;
      if (r == c.rank) {
        x = in[v];
      } else {
        const uint2* src = comm_ll_slot(c, c.rank, r, e) + 4 * v;
        const uint2 a = ll_load2(src, e), b = ll_load2(src + 2, e);
        x = make_uint4(a.x, a.y, b.x, b.y);
      }
      if constexpr (kF32) {
        acc[0] += __uint_as_float(x.x);
        acc[1] += __uint_as_float(x.y);
        acc[2] += __uint_as_float(x.z);
        acc[3] += __uint_as_float(x.w);
      } else {
        accum8<T>(acc, x);
      }
    }
    if constexpr (kF32)
      out[v] = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]),
                          __float_as_uint(acc[3]));
    else
      out[v] = pack8<T>(acc);
  }
  __syncthreads();
  if (threadIdx.x == 0) comm_end(c, e, gridDim.x);
}

// LL is used while its doubled wire size is cheaper than the flag protocol's fence + flag round
// trip (env B200_COMM_LL_MAX overrides the payload threshold in bytes).
// Default 4 MiB / world: a rank receives (world - 1) x 2 x payload over its NVLink ports, so the
// break-even payload shrinks with the group size (measured at TP = 2: LL 6.1 us at 512 KiB, flag
// protocol 11.1 us at 1 MiB).
static int64_t ll_max_bytes(int world) {
  static int64_t v = [] {
    const char* e = std::getenv("B200_COMM_LL_MAX");
    return (e && *e) ? std::atoll(e) : static_cast<int64_t>(-1);
  }();
  return v >= 0 ? v : static_cast<int64_t>(4 << 20) / world;
}

}  // namespace b200

extern "C" {

int b200_comm_handle_bytes(void) { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }

size_t b200_comm_buffer_bytes(int world, size_t max_bytes) {
  const size_t slot = (max_bytes + 1023) / 1024 * 1024;
  return b200::kCommDataOffset + 4 * static_cast<size_t>(world) * slot;  // 2 LL + 2 raw parities
}

int b200_comm_alloc(size_t bytes, void** ptr, void* handle_out) {
  using namespace b200;
  B200_REQUIRE(ptr && handle_out && bytes > 0, "b200_comm_alloc: bad arguments");
  void* p = nullptr;
  B200_CUDA_OK(cudaMalloc(&p, bytes));
  B200_CUDA_OK(cudaMemset(p, 0, bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  std::memcpy(handle_out, &h, sizeof(h));
  *ptr = p;
  return 0;
}

int b200_comm_open(const void* handle, void** ptr) {
  using namespace b200;
  B200_REQUIRE(handle && ptr, "b200_comm_open: null pointer");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  B200_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return 0;
}

int b200_comm_close(void* ptr) {
  using namespace b200;
  if (ptr) B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int b200_comm_free(void* ptr) {
  using namespace b200;
  if (ptr) B200_CUDA_OK(cudaFree(ptr));
  return 0;
}

int b200_comm_create(int rank, int world, void* const* peer_bufs, size_t buf_bytes,
                     void** comm_out) {
  using namespace b200;
  B200_REQUIRE(comm_out && peer_bufs, "b200_comm_create: null pointer");
  B200_REQUIRE(world >= 2 && world <= kCommMaxRanks && rank >= 0 && rank < world,
               "b200_comm_create: rank %d / world %d (2..%d ranks supported)", rank, world,
               kCommMaxRanks);
  B200_REQUIRE(buf_bytes > kCommDataOffset + 4 * static_cast<size_t>(world) * 1024,
               "b200_comm_create: buffer of %zu bytes is too small", buf_bytes);
  Comm* c = new Comm();
  for (int r = 0; r < world; ++r) {
    if (!peer_bufs[r]) {
      delete c;
      set_error("b200_comm_create: null buffer for rank %d", r);
      return 1;
    }
    c->dev.peer[r] = static_cast<uint8_t*>(peer_bufs[r]);
  }
  for (int r = world; r < kCommMaxRanks; ++r) c->dev.peer[r] = nullptr;
  c->dev.rank = rank;
  c->dev.world = world;
  c->dev.slot_bytes = ((buf_bytes - kCommDataOffset) / (4 * static_cast<size_t>(world))) / 1024 * 1024;
  c->buf_bytes = buf_bytes;
  *comm_out = c;
  return 0;
}

int b200_comm_destroy(void* comm) {
  delete static_cast<b200::Comm*>(comm);
  return 0;
}

size_t b200_comm_max_bytes(const void* comm) {
  return comm ? static_cast<const b200::Comm*>(comm)->dev.slot_bytes : 0;
}

int b200_comm_all_reduce(void* comm, const void* in, void* out, int64_t numel, int dtype,
                         void* stream) {
  using namespace b200;
  B200_REQUIRE(comm && in && out, "b200_comm_all_reduce: null pointer");
  if (numel == 0) return 0;
  const Comm* c = static_cast<const Comm*>(comm);
  const int esize = dtype == B200_F32 ? 4 : 2;
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16 || dtype == B200_F32,
               "b200_comm_all_reduce: dtype %d", dtype);
  const int64_t bytes = numel * esize;
  B200_REQUIRE(bytes % 16 == 0 && aligned16(in) && aligned16(out),
               "b200_comm_all_reduce: byte size must be a multiple of 16 and pointers 16-byte "
               "aligned (same rule as custom_all_reduce.py:418-420)");
  B200_REQUIRE(static_cast<uint64_t>(bytes) <= c->dev.slot_bytes,
               "b200_comm_all_reduce: %lld bytes exceed the communicator's max_size %llu",
               static_cast<long long>(bytes), static_cast<unsigned long long>(c->dev.slot_bytes));
  const int64_t num_vec = bytes / 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    const int rc = comm_all_reduce_nvls(c, in, out, bytes, dtype, st);  // -1: not attached / too small
    if (rc >= 0) return rc;
  }
  const uint4* i4 = static_cast<const uint4*>(in);
  uint4* o4 = static_cast<uint4*>(out);
  if (bytes <= ll_max_bytes(c->dev.world) && static_cast<uint64_t>(2 * bytes) <= c->dev.slot_bytes) {
    int64_t ctas64 = (num_vec + 127) / 128;
    const int64_t cap = 4 * static_cast<int64_t>(device_sm_count());
    const int ctas = static_cast<int>(ctas64 < cap ? ctas64 : cap);
    {
      ScopedKernelTimer timer("all_reduce_ll", st);
      cudaError_t err;
      if (dtype == B200_BF16)
        err = launch_kernel(all_reduce_ll_kernel<__nv_bfloat16, false>, dim3(ctas), dim3(128), 0, st,
                            1, kPdlOther, c->dev, i4, o4, num_vec);
      else if (dtype == B200_F16)
        err = launch_kernel(all_reduce_ll_kernel<__half, false>, dim3(ctas), dim3(128), 0, st, 1,
                            kPdlOther, c->dev, i4, o4, num_vec);
      else
        err = launch_kernel(all_reduce_ll_kernel<float, true>, dim3(ctas), dim3(128), 0, st, 1,
                            kPdlOther, c->dev, i4, o4, num_vec);
      B200_CUDA_OK(err);
    }
    B200_CHECK_LAUNCH("all_reduce_ll_kernel");
    return 0;
  }
  // 8 KiB per CTA (2 vectors per thread) up to 2 CTAs per SM, then grow the per-CTA share.
  int64_t vec_per_cta = 512;
  const int64_t sm2 = 2 * static_cast<int64_t>(device_sm_count());
  const int64_t max_ctas = sm2 < kCommMaxBlocks ? sm2 : kCommMaxBlocks;
  if ((num_vec + vec_per_cta - 1) / vec_per_cta > max_ctas)
    vec_per_cta = (num_vec + max_ctas - 1) / max_ctas;
  const int ctas = static_cast<int>((num_vec + vec_per_cta - 1) / vec_per_cta);
  {
    ScopedKernelTimer timer("all_reduce_oneshot", st);
    cudaError_t err;
    if (dtype == B200_BF16)
      err = launch_kernel(all_reduce_oneshot_kernel<__nv_bfloat16, false>, dim3(ctas), dim3(256), 0,
                          st, 1, kPdlOther, c->dev, i4, o4, num_vec, static_cast<int>(vec_per_cta));
    else if (dtype == B200_F16)
      err = launch_kernel(all_reduce_oneshot_kernel<__half, false>, dim3(ctas), dim3(256), 0, st, 1,
                          kPdlOther, c->dev, i4, o4, num_vec, static_cast<int>(vec_per_cta));
    else
      err = launch_kernel(all_reduce_oneshot_kernel<float, true>, dim3(ctas), dim3(256), 0, st, 1,
                          kPdlOther, c->dev, i4, o4, num_vec, static_cast<int>(vec_per_cta));
    B200_CUDA_OK(err);
  }
  B200_CHECK_LAUNCH("all_reduce_oneshot_kernel");
  return 0;
}

}  // extern "C"
