// Greedy sampling step of the decode loop: row-wise argmax of the logits (first index of the maximum,
// torch.argmax semantics).  The reference does this with torch.argmax in Sampler.forward
// (python/sglang/srt/layers/sampler.py:60-66); torch's reduce kernel puts ONE CTA on a 128k-entry row
// (46 us per decode step in the whole-model launch list, profiles/r01_ncu_decode_model_bs1_launches.json).
// Here a row is spread over a cluster of 8 CTAs (16-byte loads), the per-CTA candidates are combined
// through distributed shared memory by rank 0.
#include "common.cuh"

namespace b200 {

constexpr int kAmThreads = 256;
constexpr int kAmCluster = 8;

struct Cand {
  float v;
  int i;
};
__device__ __forceinline__ Cand better(Cand a, Cand b) {  // larger value, then smaller index
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}

template <typename T>
__global__ void __launch_bounds__(kAmThreads)
argmax_rows_kernel(const T* __restrict__ x, int64_t ldx, int64_t* __restrict__ out, int V, int per) {
  __shared__ float sv[kAmThreads / 32];
  __shared__ int si[kAmThreads / 32];
  __shared__ float cv[kAmCluster];
  __shared__ int ci[kAmCluster];
  uint32_t S, rank;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(S));
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(rank));
  if (S > 1)  // every CTA of the cluster runs before any remote store is issued
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
  griddep_wait();
  griddep_launch_dependents();
  const int m = blockIdx.x / S;
  const T* row = x + m * ldx;
  const int lo = static_cast<int>(rank) * per, hi = min(V, lo + per);  // per % 8 == 0
  Cand best{-INFINITY, 0x7fffffff};
  const bool vec_ok = (ldx % 8 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15u) == 0);
  if (vec_ok) {
    const int nfull = lo + ((hi - lo) / 8) * 8;
    for (int i = lo + threadIdx.x * 8; i < nfull; i += kAmThreads * 8) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(row + i));
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack2<T>(w[e]);
        if (f.x > best.v) best = Cand{f.x, i + 2 * e};       // ascending i: strict > keeps the first
        if (f.y > best.v) best = Cand{f.y, i + 2 * e + 1};
      }
    }
    for (int i = nfull + threadIdx.x; i < hi; i += kAmThreads) {
      const float f = to_float<T>(row[i]);
      if (f > best.v) best = Cand{f, i};
    }
  } else {
    for (int i = lo + threadIdx.x; i < hi; i += kAmThreads) {
      const float f = to_float<T>(row[i]);
      if (f > best.v) best = Cand{f, i};
    }
  }
  // a thread's candidates are in ascending index order, but threads interleave: full tie-break from here
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    Cand o{__shfl_xor_sync(0xffffffffu, best.v, off), __shfl_xor_sync(0xffffffffu, best.i, off)};
    best = better(best, o);
  }
  if ((threadIdx.x & 31) == 0) {
    sv[threadIdx.x >> 5] = best.v;
    si[threadIdx.x >> 5] = best.i;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kAmThreads / 32; ++w) best = better(best, Cand{sv[w], si[w]});
    if (S > 1) {
      uint32_t rv, ri;
      const uint32_t lv = smem_u32(cv + rank), li = smem_u32(ci + rank);
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(rv) : "r"(lv), "r"(0));
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(ri) : "r"(li), "r"(0));
      asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(rv), "f"(best.v) : "memory");
      asm volatile("st.shared::cluster.u32 [%0], %1;\n" ::"r"(ri), "r"(best.i) : "memory");
    }
  }
  if (S > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    if (rank == 0 && threadIdx.x == 0) {
      Cand b{cv[0], ci[0]};
      for (uint32_t r = 1; r < S; ++r) b = better(b, Cand{cv[r], ci[r]});
      out[m] = b.i == 0x7fffffff ? 0 : b.i;
    }
  } else if (threadIdx.x == 0) {
    out[m] = best.i == 0x7fffffff ? 0 : best.i;
  }
}

}  // namespace b200

extern "C" int b200_argmax_rows(const void* x, int64_t ldx, int64_t* out, int M, int V, int dtype,
                                void* stream) {
  using namespace b200;
  if (M == 0) return 0;
  B200_REQUIRE(x && out && V > 0 && M > 0, "b200_argmax_rows: bad arguments");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_argmax_rows: dtype %d", dtype);
  // long rows on few tokens: a cluster per row; many tokens fill the machine with one CTA per row
  int S = kAmCluster;
  while (S > 1 && (M * S > 4 * device_sm_count() || V < S * kAmThreads * 8)) S >>= 1;
  const int per = ceil_div(ceil_div(V, S), 8) * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    ScopedKernelTimer timer("argmax_rows", st);
    if (dtype == B200_BF16)
      B200_CUDA_OK(launch_kernel(argmax_rows_kernel<__nv_bfloat16>, dim3(M * S), dim3(kAmThreads), 0, st, S,
                                 pdl_family(M, kPdlOther), static_cast<const __nv_bfloat16*>(x), ldx, out,
                                 V, per));
    else
      B200_CUDA_OK(launch_kernel(argmax_rows_kernel<__half>, dim3(M * S), dim3(kAmThreads), 0, st, S,
                                 pdl_family(M, kPdlOther), static_cast<const __half*>(x), ldx, out, V, per));
  }
  B200_CHECK_LAUNCH("argmax_rows_kernel");
  return 0;
}
