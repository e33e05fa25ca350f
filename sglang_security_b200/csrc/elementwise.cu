// The steps either side of every GEMM / attention call of a decoder layer,
// fused so that a decode step launches one kernel where the reference launches two or three:
//
//   b200_rmsnorm        (residual +) RMSNorm [-> FP8 per-token / per-128-group quant]
//                       reference: RMSNorm.forward_native, python/sglang/srt/layers/layernorm.py:88-108
//                       (+ sgl_per_token_quant_fp8 / per_token_group_quant_fp8 that follow it in
//                       Fp8LinearMethod.apply, fp8_utils.py:454-648)
//   b200_silu_and_mul   SiLU(gate) * up [-> FP8 quant]
//                       reference: SiluAndMul.forward_native, layers/activation.py:42-45
//   b200_rope_kv_write  rotary embedding of q and k in place [+ set_kv_buffer of k, v]
//                       reference: RotaryEmbedding.forward_native + _apply_rotary_emb,
//                       layers/rotary_embedding.py:30-57,119-147; memory_pool.py:378-412
//
// All three are row-wise and bandwidth/latency-bound: one CTA per token row, the row lives in
// registers between the reduction and the output pass, 16-byte loads and stores.  Rounding points
// follow the torch-native reference: every tensor the reference materialises in 16-bit is rounded
// to 16-bit here too, so a fused call produces exactly what the unfused chain of our own kernels
// produces (norm -> 16-bit -> quant) and differs from the CPU oracle only by reduction order.
#include <cuda_fp8.h>

#include <cstdlib>

#include "common.cuh"

namespace b200 {

constexpr float kFp8MaxE = 448.0f;
constexpr int kRowThreads = 256;
constexpr int kMaxIters = 8;  // 256 threads x 8 elements x 8 iterations = 16384 columns

__device__ __forceinline__ uint8_t f32_to_e4m3(float v) {
  return static_cast<uint8_t>(__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3));
}

template <typename T>
__device__ __forceinline__ void ld8(const T* p, float (&f)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 v = unpack2<T>(w[e]);
    f[2 * e] = v.x;
    f[2 * e + 1] = v.y;
  }
}
template <typename T>
__device__ __forceinline__ uint32_t pack16x2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack16x2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack16x2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack16x2<T>(f[0], f[1]), pack16x2<T>(f[2], f[3]),
                                            pack16x2<T>(f[4], f[5]), pack16x2<T>(f[6], f[7]));
}
// round a float to T and back (the value a 16-bit tensor would hold)
template <typename T>
__device__ __forceinline__ float rnd(float x) {
  return to_float<T>(from_float<T>(x));
}

template <bool MAX>
__device__ __forceinline__ float block_reduce(float v, float* red) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float o = __shfl_xor_sync(0xffffffffu, v, off);
    v = MAX ? fmaxf(v, o) : v + o;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // red may still be read from a previous reduction
  if (lane == 0) red[warp] = v;
  __syncthreads();
  v = lane < (kRowThreads / 32) ? red[lane] : 0.f;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float o = __shfl_xor_sync(0xffffffffu, v, off);
    v = MAX ? fmaxf(v, o) : v + o;
  }
  return v;
}

struct RowOut {
  void* out;       // 16-bit output or null
  int64_t ldo;
  uint8_t* q;      // e4m3 output or null
  int64_t ldq;
  float* scales;   // per-token [M] (s_row = 1) or per-group [M, K/128] at m*s_row + g*s_col
  int64_t s_row, s_col;
  int quant_mode;  // 0 none, 1 per token (reciprocal multiply), 2 per 128-group (division, eps)
  float eps;       // group quant epsilon
};

// Writes one row held in registers as y[it][8] (values already rounded to T) to the 16-bit output
// and/or quantises it.  The arithmetic of the quantisers is that of fp8_quant.cu, so fused and
// unfused paths emit identical bytes.
// A row may be split over the CTAs of a cluster (small token counts): this CTA then holds the
// vectors i0 + threadIdx.x + it * kRowThreads < nvec (nvec = end of its window, i0 a multiple of 16 so
// that 128-groups stay inside one CTA) and the per-token amax is combined across the cluster through
// distributed shared memory — a maximum, so the result does not depend on the split.
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_size() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ float cluster_max(float v, float* slots /* [8] shared */) {
  const uint32_t S = cluster_size();
  if (S == 1) return v;
  const uint32_t rank = cluster_rank();
  if (threadIdx.x < S) {  // thread j publishes this CTA's value into CTA j's slot [rank]
    uint32_t local = static_cast<uint32_t>(__cvta_generic_to_shared(slots + rank)), remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"(threadIdx.x));
    asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(remote), "f"(v) : "memory");
  }
  cluster_barrier();
  float r = slots[0];
  for (uint32_t j = 1; j < S; ++j) r = fmaxf(r, slots[j]);
  return r;
}

// sum over the cluster in rank order (the same value on every CTA, deterministic for a given S)
__device__ __forceinline__ float cluster_sum(float v, float* slots /* [8] shared */) {
  const uint32_t S = cluster_size();
  if (S == 1) return v;
  const uint32_t rank = cluster_rank();
  if (threadIdx.x < S) {
    uint32_t local = static_cast<uint32_t>(__cvta_generic_to_shared(slots + rank)), remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"(threadIdx.x));
    asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(remote), "f"(v) : "memory");
  }
  cluster_barrier();
  float r = slots[0];
  for (uint32_t j = 1; j < S; ++j) r += slots[j];
  return r;
}

template <typename T>
__device__ __forceinline__ void emit_row(const float (&y)[kMaxIters][8], int nvec, int m,
                                         const RowOut& o, float* red, int i0 = 0,
                                         float* cl_slots = nullptr) {
  if (o.out) {
    T* orow = static_cast<T*>(o.out) + m * o.ldo;
#pragma unroll
    for (int it = 0; it < kMaxIters; ++it) {
      const int i = i0 + threadIdx.x + it * kRowThreads;
      if (i < nvec) st8(orow + i * 8, y[it]);
    }
  }
  if (o.quant_mode == 1) {
    float amax = 0.f;
#pragma unroll
    for (int it = 0; it < kMaxIters; ++it) {
      const int i = i0 + threadIdx.x + it * kRowThreads;
      if (i < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(y[it][e]));
      }
    }
    amax = block_reduce<true>(amax, red);
    if (cl_slots) amax = cluster_max(amax, cl_slots);
    const float scale = amax / kFp8MaxE;
    if (threadIdx.x == 0 && (!cl_slots || cluster_rank() == 0)) o.scales[m * o.s_row] = scale;
    const float inv = scale > 0.f ? 1.0f / scale : 0.f;
    uint8_t* qrow = o.q + m * o.ldq;
#pragma unroll
    for (int it = 0; it < kMaxIters; ++it) {
      const int i = i0 + threadIdx.x + it * kRowThreads;
      if (i < nvec) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          lo |= static_cast<uint32_t>(
                    f32_to_e4m3(fmaxf(fminf(y[it][e] * inv, kFp8MaxE), -kFp8MaxE))) << (8 * e);
          hi |= static_cast<uint32_t>(
                    f32_to_e4m3(fmaxf(fminf(y[it][4 + e] * inv, kFp8MaxE), -kFp8MaxE))) << (8 * e);
        }
        *reinterpret_cast<uint2*>(qrow + i * 8) = make_uint2(lo, hi);
      }
    }
  } else if (o.quant_mode == 2) {
    uint8_t* qrow = o.q + m * o.ldq;
#pragma unroll
    for (int it = 0; it < kMaxIters; ++it) {
      const int i = i0 + threadIdx.x + it * kRowThreads;  // 16 consecutive threads = one 128-group
      float amax = o.eps;
      if (i < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(y[it][e]));
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1)
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
      if (i < nvec) {
        const float s = amax / kFp8MaxE;
        if ((threadIdx.x & 15) == 0) o.scales[m * o.s_row + (i >> 4) * o.s_col] = s;
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          lo |= static_cast<uint32_t>(
                    f32_to_e4m3(fminf(fmaxf(y[it][e] / s, -kFp8MaxE), kFp8MaxE))) << (8 * e);
          hi |= static_cast<uint32_t>(
                    f32_to_e4m3(fminf(fmaxf(y[it][4 + e] / s, -kFp8MaxE), kFp8MaxE))) << (8 * e);
        }
        *reinterpret_cast<uint2*>(qrow + i * 8) = make_uint2(lo, hi);
      }
    }
  }
}

// ---- (residual +) RMSNorm -----------------------------------------------------------------------
// layernorm.py:95-104: x32 = x (+ residual); residual' = T(x32); var = mean(x32^2);
//                      y = T( (x32 * rsqrt(var + eps)) * weight )
// Small token counts: a row is split over a cluster of gridDim.x / M CTAs (`per` vectors each); the sum
// of squares is combined over distributed shared memory in rank order, the per-token amax as a maximum.
template <typename T>
__global__ void __launch_bounds__(kRowThreads)
rmsnorm_kernel(const T* __restrict__ x, int64_t ldx, T* __restrict__ residual, int64_t ldr,
               const T* __restrict__ weight, int K, float eps, int per, RowOut o) {
  __shared__ float red[32];
  __shared__ float cl_sum[8];
  __shared__ float cl_max[8];
  const uint32_t S = cluster_size();
  if (S > 1) cluster_barrier();  // every CTA of the cluster runs before any remote store is issued
  const int m = blockIdx.x / S;
  const int i0 = static_cast<int>(cluster_rank()) * per;
  const int nvec = min(K / 8, i0 + per);
  // the norm weight does not depend on the previous kernel: fetch it before the dependency wait so
  // its latency overlaps the predecessor's tail instead of following the reduction
  uint4 wraw[kMaxIters];
#pragma unroll
  for (int it = 0; it < kMaxIters; ++it) {
    const int i = i0 + threadIdx.x + it * kRowThreads;
    wraw[it] = i < nvec ? __ldg(reinterpret_cast<const uint4*>(weight + i * 8)) : make_uint4(0, 0, 0, 0);
  }
  griddep_wait();
  griddep_launch_dependents();
  const T* xrow = x + m * ldx;
  T* rrow = residual ? residual + m * ldr : nullptr;
  float y[kMaxIters][8];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kMaxIters; ++it) {
    const int i = i0 + threadIdx.x + it * kRowThreads;
    if (i < nvec) {
      ld8(xrow + i * 8, y[it]);
      if (rrow) {
        float r[8];
        ld8(rrow + i * 8, r);
#pragma unroll
        for (int e = 0; e < 8; ++e) y[it][e] = __fadd_rn(y[it][e], r[e]);
        st8(rrow + i * 8, y[it]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(y[it][e], y[it][e], ss);
    }
  }
  ss = block_reduce<false>(ss, red);
  ss = cluster_sum(ss, cl_sum);
  const float inv = 1.0f / sqrtf(__fadd_rn(ss / static_cast<float>(K), eps));
#pragma unroll
  for (int it = 0; it < kMaxIters; ++it) {
    const int i = i0 + threadIdx.x + it * kRowThreads;
    if (i < nvec) {
      const uint32_t wr[4] = {wraw[it].x, wraw[it].y, wraw[it].z, wraw[it].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 w = unpack2<T>(wr[e]);
        y[it][2 * e] = rnd<T>(__fmul_rn(__fmul_rn(y[it][2 * e], inv), w.x));
        y[it][2 * e + 1] = rnd<T>(__fmul_rn(__fmul_rn(y[it][2 * e + 1], inv), w.y));
      }
    }
  }
  emit_row<T>(y, nvec, m, o, red, i0, S > 1 ? cl_max : nullptr);
}

// ---- SiLU(gate) * up ------------------------------------------------------------------------------
// activation.py:43-45: F.silu(x[..., :d]) * x[..., d:] on 16-bit tensors = T( T(silu(g)) * u )
// Small token counts: the row is split over a cluster of gridDim.x / M CTAs (one CTA per row leaves
// a 14336-wide row to 8 warps of one SM: 11 us at M = 1).
template <typename T>
__global__ void __launch_bounds__(kRowThreads)
silu_mul_kernel(const T* __restrict__ x, int64_t ldx, int d, int per, RowOut o) {
  __shared__ float red[32];
  __shared__ float cl_slots[8];
  const uint32_t S = cluster_size();
  if (S > 1) cluster_barrier();  // every CTA of the cluster runs before any remote store is issued
  griddep_wait();
  griddep_launch_dependents();
  const int m = blockIdx.x / S;
  const int i0 = static_cast<int>(cluster_rank()) * per;
  const int nvec = min(d / 8, i0 + per);
  const T* grow = x + m * ldx;
  const T* urow = grow + d;
  float y[kMaxIters][8];
#pragma unroll
  for (int it = 0; it < kMaxIters; ++it) {
    const int i = i0 + threadIdx.x + it * kRowThreads;
    if (i < nvec) {
      float u[8];
      ld8(grow + i * 8, y[it]);
      ld8(urow + i * 8, u);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float g = y[it][e];
        const float s = rnd<T>(g / (1.0f + expf(-g)));
        y[it][e] = rnd<T>(__fmul_rn(s, u[e]));
      }
    }
  }
  emit_row<T>(y, nvec, m, o, red, i0, S > 1 ? cl_slots : nullptr);
}

// ---- rotary embedding (+ KV-cache write) ------------------------------------------------------------
struct RopeParams {
  const int64_t* positions;
  void* q;
  int64_t q_stride;
  void* k;
  int64_t k_stride;
  const void* v;
  int64_t v_stride;
  const void* cache;   // cos_sin_cache [max_pos, rot]: cos (rot/2) then sin (rot/2)
  int cache_f32;
  int rot, is_neox, native_rounding;
  void* k_cache;
  void* v_cache;
  int64_t cache_row_stride;
  const int64_t* loc;
  int Hq, Hkv, D;
  // FP8 pool (b200_rope_kv_write_fp8): rows are bytes, values quantised as b200_kv_write_fp8 does;
  // 0 = 16-bit pool, 1 = e4m3, 2 = e5m2
  int cache_fp8;
  float k_scale, v_scale;
  int has_scale;
};

// 8 values (already rounded to T) into a cache row at element offset `off`
template <typename T>
__device__ __forceinline__ void st_cache8(const RopeParams& p, uint8_t* row, int off,
                                          const float (&f)[8], float scale) {
  if (p.cache_fp8) {
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      lo |= to_fp8_torch<T>(from_float<T>(f[j]), scale, p.has_scale != 0, p.cache_fp8 == 2) << (8 * j);
      hi |= to_fp8_torch<T>(from_float<T>(f[4 + j]), scale, p.has_scale != 0, p.cache_fp8 == 2) << (8 * j);
    }
    *reinterpret_cast<uint2*>(row + off) = make_uint2(lo, hi);
  } else {
    st8(reinterpret_cast<T*>(row) + off, f);
  }
}
// raw 16-bit chunk -> cache row (pass-through part of k, all of v)
template <typename T>
__device__ __forceinline__ void copy_cache8(const RopeParams& p, uint8_t* row, int off, const T* src,
                                            float scale) {
  if (p.cache_fp8) {
    float f[8];
    ld8(src, f);
    st_cache8<T>(p, row, off, f, scale);
  } else {
    *reinterpret_cast<uint4*>(reinterpret_cast<T*>(row) + off) = *reinterpret_cast<const uint4*>(src);
  }
}

template <typename T>
__device__ __forceinline__ void load_cs(const RopeParams& p, int64_t pos, int j0, float (&c)[8],
                                        float (&s)[8]) {
  const int half = p.rot / 2;
  if (p.cache_f32) {
    const float* row = static_cast<const float*>(p.cache) + pos * p.rot;
#pragma unroll
    for (int e = 0; e < 8; e += 4) {
      const float4 cv = *reinterpret_cast<const float4*>(row + j0 + e);
      const float4 sv = *reinterpret_cast<const float4*>(row + half + j0 + e);
      c[e] = cv.x; c[e + 1] = cv.y; c[e + 2] = cv.z; c[e + 3] = cv.w;
      s[e] = sv.x; s[e + 1] = sv.y; s[e + 2] = sv.z; s[e + 3] = sv.w;
    }
  } else {
    const T* row = static_cast<const T*>(p.cache) + pos * p.rot;
    ld8(row + j0, c);
    ld8(row + half + j0, s);
  }
}

// o1 = x1*cos - x2*sin ; o2 = x2*cos + x1*sin  (rotary_embedding.py:50-51).  native: every product
// and the sum are rounded to T with cos/sin cast to T first (what torch does on 16-bit tensors);
// otherwise fp32 throughout with one final rounding (what the reference's CUDA kernel does).
template <typename T>
__device__ __forceinline__ void rotate(float x1, float x2, float c, float s, bool native, float& o1,
                                       float& o2) {
  if (native) {
    c = rnd<T>(c);
    s = rnd<T>(s);
    o1 = rnd<T>(rnd<T>(__fmul_rn(x1, c)) - rnd<T>(__fmul_rn(x2, s)));
    o2 = rnd<T>(rnd<T>(__fmul_rn(x2, c)) + rnd<T>(__fmul_rn(x1, s)));
  } else {
    o1 = fmaf(x1, c, -__fmul_rn(x2, s));
    o2 = fmaf(x2, c, __fmul_rn(x1, s));
  }
}

// One CTA per token.  Work items: for every q and k head the chunks (8 elements) of the first
// rotation half (neox: paired with the chunk rot/2 further; gpt-j: pairs are adjacent elements),
// for k additionally the pass-through chunks (cache write only), for v all chunks (cache write).
template <typename T>
__global__ void __launch_bounds__(256)
rope_kv_write_kernel(const RopeParams p) {
  griddep_wait();
  griddep_launch_dependents();
  const int t = blockIdx.x;
  const int64_t pos = p.positions[t];
  const int cpd = p.D / 8;                                 // chunks per head row
  const int rot_items = p.is_neox ? p.rot / 16 : p.rot / 8;  // rotating work items per head
  const int k_items = rot_items + (p.D - p.rot) / 8;       // + pass-through chunks (k only)
  const bool write = p.k_cache != nullptr;
  const int64_t slot = write ? p.loc[t] : 0;
  T* qrow = static_cast<T*>(p.q) + t * p.q_stride;
  T* krow = static_cast<T*>(p.k) + t * p.k_stride;
  const T* vrow = write ? static_cast<const T*>(p.v) + t * p.v_stride : nullptr;
  const int esz = p.cache_fp8 ? 1 : static_cast<int>(sizeof(T));
  uint8_t* kc = write ? static_cast<uint8_t*>(p.k_cache) + slot * p.cache_row_stride * esz : nullptr;
  uint8_t* vc = write ? static_cast<uint8_t*>(p.v_cache) + slot * p.cache_row_stride * esz : nullptr;
  const int n_q = p.Hq * rot_items;
  const int n_k = p.Hkv * (write ? k_items : rot_items);
  const int n_v = write ? p.Hkv * cpd : 0;
  const bool native = p.native_rounding != 0;
  for (int w = threadIdx.x; w < n_q + n_k + n_v; w += blockDim.x) {
    if (w >= n_q + n_k) {  // v: plain copy into the cache
      const int c = w - n_q - n_k;
      copy_cache8<T>(p, vc, c * 8, vrow + c * 8, p.v_scale);
      continue;
    }
    const bool is_k = w >= n_q;
    const int per_head = is_k && write ? k_items : rot_items;
    const int idx = is_k ? w - n_q : w;
    const int h = idx / per_head, it = idx - h * per_head;
    T* row = (is_k ? krow : qrow) + h * p.D;
    uint8_t* crow = is_k && write ? kc + h * p.D * esz : nullptr;
    if (it >= rot_items) {  // pass-through chunk of k
      const int c = p.rot / 8 + (it - rot_items);
      copy_cache8<T>(p, crow, c * 8, row + c * 8, p.k_scale);
      continue;
    }
    if (p.is_neox) {
      const int j0 = it * 8;
      float x1[8], x2[8], c[8], s[8], o1[8], o2[8];
      ld8(row + j0, x1);
      ld8(row + p.rot / 2 + j0, x2);
      load_cs<T>(p, pos, j0, c, s);
#pragma unroll
      for (int e = 0; e < 8; ++e) rotate<T>(x1[e], x2[e], c[e], s[e], native, o1[e], o2[e]);
      st8(row + j0, o1);
      st8(row + p.rot / 2 + j0, o2);
      if (crow) {
        st_cache8<T>(p, crow, j0, o1, p.k_scale);
        st_cache8<T>(p, crow, p.rot / 2 + j0, o2, p.k_scale);
      }
    } else {
      // gpt-j: elements (2j, 2j+1) of this chunk use cos/sin index j = it*4 + e
      float x[8], o[8];
      ld8(row + it * 8, x);
      const int half = p.rot / 2;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int j = it * 4 + e;
        float c, s;
        if (p.cache_f32) {
          const float* cr = static_cast<const float*>(p.cache) + pos * p.rot;
          c = cr[j];
          s = cr[half + j];
        } else {
          const T* cr = static_cast<const T*>(p.cache) + pos * p.rot;
          c = to_float<T>(cr[j]);
          s = to_float<T>(cr[half + j]);
        }
        rotate<T>(x[2 * e], x[2 * e + 1], c, s, native, o[2 * e], o[2 * e + 1]);
      }
      st8(row + it * 8, o);
      if (crow) st_cache8<T>(p, crow, it * 8, o, p.k_scale);
    }
  }
}

}  // namespace b200

extern "C" {

int b200_rmsnorm(const void* x, int64_t ldx, void* residual, int64_t ldr, const void* weight,
                 void* out, int64_t ldo, void* q, int64_t ldq, float* scales,
                 int64_t scale_row_stride, int64_t scale_col_stride, int M, int K, float eps,
                 int quant_mode, float quant_eps, int dtype, void* stream) {
  using namespace b200;
  if (M == 0) return 0;
  B200_REQUIRE(x && weight && (out || q), "b200_rmsnorm: null pointer");
  B200_REQUIRE(K > 0 && K % 8 == 0 && K <= kRowThreads * 8 * kMaxIters,
               "b200_rmsnorm: K must be a multiple of 8 and <= %d (K=%d)",
               kRowThreads * 8 * kMaxIters, K);
  B200_REQUIRE(quant_mode >= 0 && quant_mode <= 2 && (quant_mode == 0 || (q && scales)),
               "b200_rmsnorm: bad quant_mode %d / missing q or scales", quant_mode);
  B200_REQUIRE(quant_mode != 2 || K % 128 == 0, "b200_rmsnorm: group quant needs K %% 128 == 0");
  B200_REQUIRE(aligned16(x) && aligned16(weight) && ldx % 8 == 0 &&
                   (!residual || (aligned16(residual) && ldr % 8 == 0)) &&
                   (!out || (aligned16(out) && ldo % 8 == 0)) &&
                   (!q || ((reinterpret_cast<uintptr_t>(q) & 7u) == 0 && ldq % 8 == 0)),
               "b200_rmsnorm: pointers must be 16-byte aligned with strides %% 8 == 0");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_rmsnorm: dtype %d", dtype);
  RowOut o{out, ldo, static_cast<uint8_t*>(q), ldq, scales, quant_mode == 1 ? 1 : scale_row_stride,
           scale_col_stride, q ? quant_mode : 0, quant_eps};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    ScopedKernelTimer timer("rmsnorm", st);
    // Optional cluster split of a row (B200_RMSNORM_TARGET_CTAS = CTA count to aim for).  Off by default:
    // unlike SiLU*mul (56 exp + divide per thread on a 14336-wide row) a 4096-wide norm row is a short
    // chain of dependent latencies that the extra cluster barriers only lengthen — measured 10.65 vs
    // 9.74 us per launch, whole step 8.238 vs 8.232 ms at 64 tokens (profiles/r02_bench_rmsnorm_split.json).
    static const int target = [] {
      const char* e = std::getenv("B200_RMSNORM_TARGET_CTAS");
      return (e && *e) ? std::atoi(e) : 0;
    }();
    const int nvec = K / 8;
    int S = 8;
    while (S > 1 && (M * S > target || nvec < S * 128)) S >>= 1;
    const int per = ceil_div(ceil_div(nvec, S), 16) * 16;  // 128-groups stay inside one CTA
    if (dtype == B200_BF16)
      B200_CUDA_OK(launch_kernel(rmsnorm_kernel<__nv_bfloat16>, dim3(M * S), dim3(kRowThreads), 0, st, S,
                                 pdl_family(M, kPdlRow), static_cast<const __nv_bfloat16*>(x), ldx,
                                 static_cast<__nv_bfloat16*>(residual), ldr,
                                 static_cast<const __nv_bfloat16*>(weight), K, eps, per, o));
    else
      B200_CUDA_OK(launch_kernel(rmsnorm_kernel<__half>, dim3(M * S), dim3(kRowThreads), 0, st, S,
                                 pdl_family(M, kPdlRow), static_cast<const __half*>(x), ldx,
                                 static_cast<__half*>(residual), ldr, static_cast<const __half*>(weight), K,
                                 eps, per, o));
  }
  B200_CHECK_LAUNCH("rmsnorm_kernel");
  return 0;
}

int b200_silu_and_mul(const void* x, int64_t ldx, void* out, int64_t ldo, void* q, int64_t ldq,
                      float* scales, int64_t scale_row_stride, int64_t scale_col_stride, int M,
                      int d, int quant_mode, float quant_eps, int dtype, void* stream) {
  using namespace b200;
  if (M == 0) return 0;
  B200_REQUIRE(x && (out || q), "b200_silu_and_mul: null pointer");
  // a CTA holds at most kRowThreads * kMaxIters vectors of a row in registers; longer rows (Qwen2.5-32B
  // without TP: d = 27648) are split over a cluster of up to 8 CTAs
  B200_REQUIRE(d > 0 && d % 8 == 0 && d <= 8 * kRowThreads * 8 * kMaxIters,
               "b200_silu_and_mul: d must be a multiple of 8 and <= %d (d=%d)",
               8 * kRowThreads * 8 * kMaxIters, d);
  B200_REQUIRE(quant_mode >= 0 && quant_mode <= 2 && (quant_mode == 0 || (q && scales)),
               "b200_silu_and_mul: bad quant_mode %d / missing q or scales", quant_mode);
  B200_REQUIRE(quant_mode != 2 || d % 128 == 0, "b200_silu_and_mul: group quant needs d %% 128 == 0");
  B200_REQUIRE(aligned16(x) && ldx % 8 == 0 && (!out || (aligned16(out) && ldo % 8 == 0)) &&
                   (!q || ((reinterpret_cast<uintptr_t>(q) & 7u) == 0 && ldq % 8 == 0)),
               "b200_silu_and_mul: pointers must be 16-byte aligned with strides %% 8 == 0");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_silu_and_mul: dtype %d", dtype);
  RowOut o{out, ldo, static_cast<uint8_t*>(q), ldq, scales, quant_mode == 1 ? 1 : scale_row_stride,
           scale_col_stride, q ? quant_mode : 0, quant_eps};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    ScopedKernelTimer timer("silu_and_mul", st);
    // cluster size: spread few rows over more SMs while every CTA keeps >= 256 vectors
    const int nvec = d / 8;
    // (one CTA per row leaves a 14336-wide row — 56 exp + divide per thread — to 8 warps of one SM:
    // 11.7 us at M = 64 in the whole-model launch list, profiles/r02_ncu_launches.json), so rows are
    // split until about two CTAs per SM exist; B200_SILU_TARGET_CTAS overrides the target
    static const int target = [] {
      const char* e = std::getenv("B200_SILU_TARGET_CTAS");
      return (e && *e) ? std::atoi(e) : 2 * device_sm_count();
    }();
    int S = 8;
    while (S > 1 && (M * S > target || nvec < S * kRowThreads)) S >>= 1;
    while (ceil_div(ceil_div(nvec, S), 16) * 16 > kRowThreads * kMaxIters) S <<= 1;  // register window
    const int per = ceil_div(ceil_div(nvec, S), 16) * 16;  // 128-groups stay inside one CTA
    if (dtype == B200_BF16)
      B200_CUDA_OK(launch_kernel(silu_mul_kernel<__nv_bfloat16>, dim3(M * S), dim3(kRowThreads), 0, st,
                                 S, pdl_family(M, kPdlRow), static_cast<const __nv_bfloat16*>(x), ldx, d, per, o));
    else
      B200_CUDA_OK(launch_kernel(silu_mul_kernel<__half>, dim3(M * S), dim3(kRowThreads), 0, st, S, pdl_family(M, kPdlRow),
                                 static_cast<const __half*>(x), ldx, d, per, o));
  }
  B200_CHECK_LAUNCH("silu_mul_kernel");
  return 0;
}

static int rope_kv_write_impl(const int64_t* positions, void* q, int64_t q_stride, void* k,
                              int64_t k_stride, const void* v, int64_t v_stride,
                              const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                              int native_rounding, void* k_cache, void* v_cache,
                              int64_t cache_row_stride, const int64_t* loc, int cache_fp8,
                              float k_scale, float v_scale, int has_scale, int T, int Hq, int Hkv,
                              int D, int dtype, void* stream) {
  using namespace b200;
  if (T == 0) return 0;
  B200_REQUIRE(positions && q && k && cos_sin_cache, "b200_rope_kv_write: null pointer");
  B200_REQUIRE((k_cache == nullptr) == (v_cache == nullptr) && (!k_cache || (v && loc)),
               "b200_rope_kv_write: k_cache, v_cache, v and loc must be given together");
  B200_REQUIRE(D % 8 == 0 && rotary_dim > 0 && rotary_dim <= D && rotary_dim % 16 == 0,
               "b200_rope_kv_write: head_dim %% 8 and rotary_dim %% 16 must be 0 (D=%d rot=%d)", D,
               rotary_dim);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_rope_kv_write: dtype %d", dtype);
  B200_REQUIRE(cache_dtype == B200_F32 || cache_dtype == dtype,
               "b200_rope_kv_write: cos_sin_cache must be fp32 or the activation dtype");
  B200_REQUIRE(aligned16(q) && aligned16(k) && aligned16(cos_sin_cache) && q_stride % 8 == 0 &&
                   k_stride % 8 == 0 &&
                   (!k_cache || (aligned16(v) && aligned16(k_cache) && aligned16(v_cache) &&
                                 v_stride % 8 == 0 && cache_row_stride % 8 == 0)),
               "b200_rope_kv_write: pointers must be 16-byte aligned with strides %% 8 == 0");
  RopeParams p{positions, q, q_stride, k, k_stride, v, v_stride, cos_sin_cache,
               cache_dtype == B200_F32 ? 1 : 0, rotary_dim, is_neox, native_rounding, k_cache,
               v_cache, cache_row_stride, loc, Hq, Hkv, D, cache_fp8, k_scale, v_scale, has_scale};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    ScopedKernelTimer timer("rope_kv_write", st);
    if (dtype == B200_BF16)
      B200_CUDA_OK(launch_kernel(rope_kv_write_kernel<__nv_bfloat16>, dim3(T), dim3(256), 0, st, 1, pdl_family(T, kPdlRope), p));
    else
      B200_CUDA_OK(launch_kernel(rope_kv_write_kernel<__half>, dim3(T), dim3(256), 0, st, 1, pdl_family(T, kPdlRope), p));
  }
  B200_CHECK_LAUNCH("rope_kv_write_kernel");
  return 0;
}

int b200_rope_kv_write(const int64_t* positions, void* q, int64_t q_stride, void* k,
                       int64_t k_stride, const void* v, int64_t v_stride,
                       const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                       int native_rounding, void* k_cache, void* v_cache, int64_t cache_row_stride,
                       const int64_t* loc, int T, int Hq, int Hkv, int D, int dtype, void* stream) {
  return rope_kv_write_impl(positions, q, q_stride, k, k_stride, v, v_stride, cos_sin_cache,
                            cache_dtype, rotary_dim, is_neox, native_rounding, k_cache, v_cache,
                            cache_row_stride, loc, 0, 1.f, 1.f, 0, T, Hq, Hkv, D, dtype, stream);
}

int b200_rope_kv_write_fp8(const int64_t* positions, void* q, int64_t q_stride, void* k,
                           int64_t k_stride, const void* v, int64_t v_stride,
                           const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                           int native_rounding, void* k_cache, void* v_cache,
                           int64_t cache_row_stride, const int64_t* loc, float k_scale,
                           float v_scale, int has_scale, int T, int Hq, int Hkv, int D, int dtype,
                           void* stream) {
  B200_REQUIRE(k_cache && v_cache, "b200_rope_kv_write_fp8: the FP8 pool rows are required");
  const int fmt = (dtype & B200_KV_FP8_E5M2) ? 2 : 1;
  dtype &= ~B200_KV_FP8_E5M2;
  B200_REQUIRE(!has_scale || (k_scale > 0.f && v_scale > 0.f),
               "b200_rope_kv_write_fp8: scales must be positive");
  return rope_kv_write_impl(positions, q, q_stride, k, k_stride, v, v_stride, cos_sin_cache,
                            cache_dtype, rotary_dim, is_neox, native_rounding, k_cache, v_cache,
                            cache_row_stride, loc, fmt, k_scale, v_scale, has_scale, T, Hq, Hkv, D,
                            dtype, stream);
}

}  // extern "C"
