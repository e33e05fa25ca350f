// INT4 weight-only (W4A16) linear kernels: AWQ and GPTQ packings.
//
// Reference path: AWQLinearMethod.apply = awq_dequantize(qweight, scales, qzeros) -> [K, N] 16-bit,
// then torch.matmul (python/sglang/srt/layers/quantization/awq.py:197-215; dequant kernel
// sgl-kernel/csrc/gemm/awq_kernel.cu:126-219, torch twin sgl-kernel/tests/test_awq_dequant.py:9-57).
// GPTQ arithmetic lives in vllm 0.8.4 (absent): restated from the public AutoGPTQ v1 format,
//   w[k,n] = (nib(qweight[k/8, n], k%8) - (nib(qzeros[g, n/8], n%8) + 1)) * scales[g, n].
//
// Internal ("B200") weight format, produced once in process_weights_after_loading:
//   qweight [K, N/8] int32 in AWQ nibble order: output column 8c + j sits at bit 4*ORDER[j],
//            ORDER = [0,4,1,5,2,6,3,7]  (pairs (2i, 2i+1) = ((x >> 4i) & 0x000f000f): one LOP3 each)
//   zeros   [K/g, N] 16-bit float holding MAGIC + z  (MAGIC = 1024 for f16, 128 for bf16) so that
//            (MAGIC + w) - (MAGIC + z) = w - z is exact in one packed subtract
//   scales  [K/g, N] 16-bit float, unchanged.
// Dequantised value = round16((w - z) * s): same two-step arithmetic as the reference kernel.
//
// Kernels
//   awq_dequant_kernel      : bit-exact standalone dequant (parity entry / reference semantics)
//   w4a16_unpack_zeros      : AWQ / GPTQ packed zeros -> biased 16-bit zeros
//   gptq_repack_kernel      : GPTQ [K/8, N] (K-packed) -> B200 [K, N/8] (N-packed), optional row perm
//   w4a16_gemm_kernel       : fused tcgen05 GEMM — x tiles by TMA (K-major A operand), weight tiles
//                             dequantised in registers by 4 warps straight into 128-byte-swizzled
//                             shared memory as an MN-major B operand; TMEM accumulators; same
//                             mbarrier rings as gemm_tc.cuh.
#include <type_traits>

#include "gemm_tc.cuh"

namespace b200 {

int w4a16_gemv_max_m();                                        // w4a16_gemv.cu
bool w4a16_gemv_supported(int M, int N, int K, int group);

template <typename T>
struct W4Magic;
template <>
struct W4Magic<__half> {
  static constexpr uint32_t kBits = 0x64006400u;  // 1024.0 | 1024.0
  static constexpr float kValue = 1024.f;
};
template <>
struct W4Magic<__nv_bfloat16> {
  static constexpr uint32_t kBits = 0x43004300u;  // 128.0 | 128.0
  static constexpr float kValue = 128.f;
};

template <typename T>
__device__ __forceinline__ uint32_t hsub2_u32(uint32_t a, uint32_t b);
template <typename T>
__device__ __forceinline__ uint32_t hmul2_u32(uint32_t a, uint32_t b);
template <>
__device__ __forceinline__ uint32_t hsub2_u32<__half>(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("sub.f16x2 %0, %1, %2;\n" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
template <>
__device__ __forceinline__ uint32_t hmul2_u32<__half>(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.f16x2 %0, %1, %2;\n" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
template <>
__device__ __forceinline__ uint32_t hsub2_u32<__nv_bfloat16>(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("sub.bf16x2 %0, %1, %2;\n" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
template <>
__device__ __forceinline__ uint32_t hmul2_u32<__nv_bfloat16>(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;\n" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

// 8 nibbles of `q` (AWQ order) -> 8 dequantised 16-bit values (w - z) * s, as 4 packed pairs in
// output-column order.  zb = biased zeros (MAGIC + z), s = scales, both 4 packed pairs.
template <typename T>
__device__ __forceinline__ uint4 dequant8(uint32_t q, const uint4& zb, const uint4& s) {
  constexpr uint32_t kMagic = W4Magic<T>::kBits;
  const uint32_t w0 = (q & 0x000f000fu) | kMagic;
  const uint32_t w1 = ((q >> 4) & 0x000f000fu) | kMagic;
  const uint32_t w2 = ((q >> 8) & 0x000f000fu) | kMagic;
  const uint32_t w3 = ((q >> 12) & 0x000f000fu) | kMagic;
  uint4 r;
  r.x = hmul2_u32<T>(hsub2_u32<T>(w0, zb.x), s.x);
  r.y = hmul2_u32<T>(hsub2_u32<T>(w1, zb.y), s.y);
  r.z = hmul2_u32<T>(hsub2_u32<T>(w2, zb.z), s.z);
  r.w = hmul2_u32<T>(hsub2_u32<T>(w3, zb.w), s.w);
  return r;
}

// ---- standalone AWQ dequant: reference layouts in, [K, N] out ------------------------------------
template <typename T>
__global__ void awq_dequant_kernel(const uint32_t* __restrict__ qweight,
                                   const T* __restrict__ scales,
                                   const uint32_t* __restrict__ qzeros, T* __restrict__ out, int K,
                                   int NC /* = N/8 */, int group) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y * blockDim.y + threadIdx.y;
  if (c >= NC || k >= K) return;
  const int g = k / group;
  const uint32_t zq = qzeros[static_cast<int64_t>(g) * NC + c];
  constexpr uint32_t kMagic = W4Magic<T>::kBits;
  uint4 zb;
  zb.x = (zq & 0x000f000fu) | kMagic;
  zb.y = ((zq >> 4) & 0x000f000fu) | kMagic;
  zb.z = ((zq >> 8) & 0x000f000fu) | kMagic;
  zb.w = ((zq >> 12) & 0x000f000fu) | kMagic;
  const uint4 s = *reinterpret_cast<const uint4*>(scales + (static_cast<int64_t>(g) * NC + c) * 8);
  const uint4 r = dequant8<T>(qweight[static_cast<int64_t>(k) * NC + c], zb, s);
  *reinterpret_cast<uint4*>(out + (static_cast<int64_t>(k) * NC + c) * 8) = r;
}

// ---- packed zeros -> biased 16-bit zeros [G, N] -------------------------------------------------
// awq_order = 1: AWQ interleaved nibbles; 0: GPTQ sequential nibbles.  add = 0 (AWQ) or 1 (GPTQ v1).
template <typename T>
__global__ void w4a16_unpack_zeros_kernel(const uint32_t* __restrict__ qzeros, T* __restrict__ out,
                                          int64_t total /* G * N/8 */, int awq_order, int add) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const uint32_t zq = qzeros[i];
  const int order[8] = {0, 4, 1, 5, 2, 6, 3, 7};
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int shift = 4 * (awq_order ? order[j] : j);
    const int z = static_cast<int>((zq >> shift) & 0xfu) + add;
    out[i * 8 + j] = from_float<T>(W4Magic<T>::kValue + static_cast<float>(z));
  }
}

// ---- GPTQ [K/8, N] -> B200 [K, N/8] -------------------------------------------------------------
// out[k', c] packs columns 8c..8c+7 of source row perm[k'] (perm == null: identity).
__global__ void gptq_repack_kernel(const uint32_t* __restrict__ qweight,
                                   const int32_t* __restrict__ perm, uint32_t* __restrict__ out,
                                   int K, int N) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int kd = blockIdx.y * blockDim.y + threadIdx.y;
  const int NC = N / 8;
  if (c >= NC || kd >= K) return;
  const int ks = perm ? perm[kd] : kd;
  const uint32_t* src = qweight + static_cast<int64_t>(ks >> 3) * N + c * 8;
  const int shift = (ks & 7) * 4;
  const int order[8] = {0, 4, 1, 5, 2, 6, 3, 7};
  uint32_t packed = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) packed |= ((src[j] >> shift) & 0xfu) << (4 * order[j]);
  out[static_cast<int64_t>(kd) * NC + c] = packed;
}

// ---- fused tcgen05 W4A16 GEMM -------------------------------------------------------------------
struct W4GemmParams {
  const uint32_t* qweight;  // [K, N/8] B200 order
  const void* zeros;        // [G, N] biased
  const void* scales;       // [G, N]
  const void* bias;         // [N] or null
  void* out;
  int64_t ldc;
  int M, N, K, group;
};

constexpr int kW4Stages = 5;
constexpr int kW4Threads = 320;  // TMA, MMA, 4 epilogue, 4 dequant warps

template <typename T>
__global__ void __launch_bounds__(kW4Threads, 1)
w4a16_gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const W4GemmParams p) {
  constexpr int BM = 128, BN = 128, BK = 64, STAGES = kW4Stages;
  constexpr int A_BYTES = BM * 128, B_BYTES = BK * BN * 2, STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES * STAGE_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int m_blocks = ceil_div(p.M, BM), n_blocks = ceil_div(p.N, BN);
  const int num_tiles = m_blocks * n_blocks;
  const int num_kb = ceil_div(p.K, BK);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmap_x);
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(full_bar(s), 1 + 128);  // TMA expect_tx arrive + 128 dequant threads
      tc::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(tfull_bar(a), 1);
      tc::mbar_init(tempty_bar(a), 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_slot, 2 * BN);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer: activation tiles =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile % m_blocks;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(empty_bar(stage), phase ^ 1u);
          tc::mbar_arrive_expect_tx(full_bar(stage), A_BYTES);
          tc::tma_load_2d(smem_base + stage * STAGE_BYTES, &tmap_x, full_bar(stage), kb * BK,
                          m_blk * BM);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      constexpr uint32_t idesc = tc::make_idesc(fmt, fmt, BM, BN, 0, 1);  // B is MN-major
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int acc = local & 1;
        tc::mbar_wait(tempty_bar(acc), ((local >> 1) & 1u) ^ 1u);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(full_bar(stage), phase);
          tc::tc_fence_after();
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t adesc = tc::make_smem_desc_sw128(sa + kk * 32, 1024, 16);
            // B rows kk*16.. of the [64 k x 128 n] tile; panels of 64 n are 64*128 bytes apart
            const uint64_t bdesc = tc::make_smem_desc_sw128(sb + kk * 16 * 128, 1024, BK * 128);
            tc::mma_f16(tmem_d, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          tc::mma_commit(empty_bar(stage));
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
        tc::mma_commit(tfull_bar(acc));
      }
    }
  } else if (warp < 6) {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;
    T* out = static_cast<T*>(p.out);
    const T* bias = static_cast<const T*>(p.bias);
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
      const int acc = local & 1;
      tc::mbar_wait(tfull_bar(acc), (local >> 1) & 1u);
      tc::tc_fence_after();
      const int row = m_blk * BM + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      T* orow = out + static_cast<int64_t>(row) * p.ldc;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tc::tmem_ld32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN + c * 32,
                      r);
        tc::tmem_ld_wait();
        const int col0 = n_blk * BN + c * 32;
        if (row_ok && col0 < p.N) {
          if (col0 + 32 <= p.N && (p.ldc & 7) == 0) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              uint32_t packed[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int j = v * 8 + e * 2;
                float x0 = __uint_as_float(r[j]), x1 = __uint_as_float(r[j + 1]);
                if (bias) {
                  x0 += to_float<T>(bias[col0 + j]);
                  x1 += to_float<T>(bias[col0 + j + 1]);
                }
                packed[e] = pack_out2<T>(x0, x1);
              }
              *reinterpret_cast<uint4*>(orow + col0 + v * 8) =
                  make_uint4(packed[0], packed[1], packed[2], packed[3]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (col0 + j >= p.N) break;
              const float b = bias ? to_float<T>(bias[col0 + j]) : 0.f;
              orow[col0 + j] = from_float<T>(__uint_as_float(r[j]) + b);
            }
          }
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tempty_bar(acc));
    }
  } else {
    // ===================== dequant warps (6..9): int4 -> 16-bit -> swizzled smem B tile ============
    const int t = threadIdx.x - 6 * 32;  // 0..127
    const int col = t & 15;              // int32 column inside the 128-wide n tile
    const int row0 = t >> 4;             // 0..7, rows row0 + 8i
    const int NC = p.N / 8;
    const T* scales = static_cast<const T*>(p.scales);
    const T* zeros = static_cast<const T*>(p.zeros);
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int n_blk = tile / m_blocks;
      const int c = n_blk * (BN / 8) + col;  // global int32 column
      const bool c_ok = c < NC;
      int g_cached = -1;
      uint4 zb = make_uint4(0, 0, 0, 0), sc = make_uint4(0, 0, 0, 0);
      uint32_t q[8];
      auto load_q = [&](int kb) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int k = kb * BK + row0 + 8 * i;
          q[i] = (c_ok && k < p.K) ? __ldg(p.qweight + static_cast<int64_t>(k) * NC + c) : 0u;
        }
      };
      load_q(0);
      for (int kb = 0; kb < num_kb; ++kb) {
        uint32_t qc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) qc[i] = q[i];
        if (kb + 1 < num_kb) load_q(kb + 1);  // prefetch the next K block's packed weights
        tc::mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sb = smem_base + stage * STAGE_BYTES + A_BYTES;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = row0 + 8 * i;
          const int k = kb * BK + r;
          uint4 w = make_uint4(0, 0, 0, 0);
          if (c_ok && k < p.K) {
            const int g = k / p.group;
            if (g != g_cached) {
              g_cached = g;
              zb = *reinterpret_cast<const uint4*>(zeros + (static_cast<int64_t>(g) * NC + c) * 8);
              sc = *reinterpret_cast<const uint4*>(scales + (static_cast<int64_t>(g) * NC + c) * 8);
            }
            w = dequant8<T>(qc[i], zb, sc);
          }
          const uint32_t addr =
              sb + (col >> 3) * (BK * 128) + r * 128 + (((col & 7) ^ (r & 7)) << 4);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(addr), "r"(w.x), "r"(w.y),
                       "r"(w.z), "r"(w.w)
                       : "memory");
        }
        tc::fence_proxy_async_smem();
        tc::mbar_arrive(full_bar(stage));
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 2 * BN);
}

template <typename T>
static int launch_w4a16_gemm(const void* x, int64_t ldx, const W4GemmParams& p, cudaStream_t st) {
  constexpr int SMEM = kW4Stages * (128 * 128 + 64 * 128 * 2) + (2 * kW4Stages + 4) * 8 + 16 + 1024;
  CUtensorMap tx;
  int rc = make_tmap_2d(&tx, x, p.M, p.K, ldx * 2, 128, 64, 2, false);
  if (rc) return rc;
  auto kern = w4a16_gemm_kernel<T>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_done = true;
  }
  const int tiles = ceil_div(p.M, 128) * ceil_div(p.N, 128);
  const int sms = device_sm_count();
  {
    ScopedKernelTimer timer("w4a16_gemm", st);
    kern<<<tiles < sms ? tiles : sms, kW4Threads, SMEM, st>>>(tx, p);
  }
  B200_CHECK_LAUNCH("w4a16_gemm_kernel");
  return 0;
}

}  // namespace b200

extern "C" {

int b200_awq_dequant(const void* qweight, const void* scales, const void* qzeros, void* out, int K,
                     int N, int group_size, int dtype, void* stream) {
  using namespace b200;
  if (K == 0 || N == 0) return 0;
  B200_REQUIRE(qweight && scales && qzeros && out, "b200_awq_dequant: null pointer");
  B200_REQUIRE(K > 0 && N > 0 && N % 8 == 0 && group_size > 0 && K % group_size == 0,
               "b200_awq_dequant: bad shape K=%d N=%d group=%d", K, N, group_size);
  B200_REQUIRE(aligned16(scales) && aligned16(out), "b200_awq_dequant: scales/out must be 16-byte aligned");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_awq_dequant: dtype %d", dtype);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int NC = N / 8;
  dim3 block(32, 8), grid(ceil_div(NC, 32), ceil_div(K, 8));
  ScopedKernelTimer timer("awq_dequant", st);
  if (dtype == B200_F16)
    awq_dequant_kernel<__half><<<grid, block, 0, st>>>(
        static_cast<const uint32_t*>(qweight), static_cast<const __half*>(scales),
        static_cast<const uint32_t*>(qzeros), static_cast<__half*>(out), K, NC, group_size);
  else
    awq_dequant_kernel<__nv_bfloat16><<<grid, block, 0, st>>>(
        static_cast<const uint32_t*>(qweight), static_cast<const __nv_bfloat16*>(scales),
        static_cast<const uint32_t*>(qzeros), static_cast<__nv_bfloat16*>(out), K, NC, group_size);
  B200_CHECK_LAUNCH("awq_dequant_kernel");
  return 0;
}

int b200_w4a16_unpack_zeros(const void* qzeros, void* zeros_out, int num_groups, int N,
                            int awq_order, int add_one, int dtype, void* stream) {
  using namespace b200;
  if (num_groups == 0 || N == 0) return 0;
  B200_REQUIRE(qzeros && zeros_out && N % 8 == 0, "b200_w4a16_unpack_zeros: bad arguments");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_w4a16_unpack_zeros: dtype %d", dtype);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t total = static_cast<int64_t>(num_groups) * (N / 8);
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  if (dtype == B200_F16)
    w4a16_unpack_zeros_kernel<__half><<<blocks, 256, 0, st>>>(
        static_cast<const uint32_t*>(qzeros), static_cast<__half*>(zeros_out), total, awq_order,
        add_one);
  else
    w4a16_unpack_zeros_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(
        static_cast<const uint32_t*>(qzeros), static_cast<__nv_bfloat16*>(zeros_out), total,
        awq_order, add_one);
  B200_CHECK_LAUNCH("w4a16_unpack_zeros_kernel");
  return 0;
}

int b200_w4a16_repack_gptq(const void* qweight, const int32_t* perm, void* out, int K, int N,
                           void* stream) {
  using namespace b200;
  if (K == 0 || N == 0) return 0;
  B200_REQUIRE(qweight && out && K % 8 == 0 && N % 8 == 0, "b200_w4a16_repack_gptq: bad arguments");
  dim3 block(32, 8), grid(ceil_div(N / 8, 32), ceil_div(K, 8));
  gptq_repack_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint32_t*>(qweight), perm, static_cast<uint32_t*>(out), K, N);
  B200_CHECK_LAUNCH("gptq_repack_kernel");
  return 0;
}

int b200_w4a16_gemm(const void* x, int64_t ldx, const void* qweight, const void* zeros,
                    const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                    int K, int group_size, int dtype, void* stream) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(x && qweight && zeros && scales && out, "b200_w4a16_gemm: null pointer");
  B200_REQUIRE(M > 0 && N > 0 && K > 0 && N % 8 == 0 && group_size > 0 && K % group_size == 0,
               "b200_w4a16_gemm: bad shape M=%d N=%d K=%d group=%d", M, N, K, group_size);
  B200_REQUIRE(aligned16(x) && ldx % 8 == 0 && aligned16(zeros) && aligned16(scales),
               "b200_w4a16_gemm: x/zeros/scales must be 16-byte aligned, ldx %% 8 == 0");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_w4a16_gemm: dtype %d", dtype);
  // decode regime: HBM-bound GEMV on the packed weights (w4a16_gemv.cu)
  if (M <= w4a16_gemv_max_m() && w4a16_gemv_supported(M, N, K, group_size) && aligned16(qweight))
    return b200_w4a16_gemv(x, ldx, qweight, zeros, scales, bias, out, ldc, M, N, K, group_size,
                           dtype, stream);
  W4GemmParams p;
  p.qweight = static_cast<const uint32_t*>(qweight);
  p.zeros = zeros;
  p.scales = scales;
  p.bias = bias;
  p.out = out;
  p.ldc = ldc;
  p.M = M;
  p.N = N;
  p.K = K;
  p.group = group_size;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return dtype == B200_F16 ? launch_w4a16_gemm<__half>(x, ldx, p, st)
                           : launch_w4a16_gemm<__nv_bfloat16>(x, ldx, p, st);
}

}  // extern "C"
