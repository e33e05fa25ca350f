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
#include <cstdlib>
#include <type_traits>

#include "gemm_tc.cuh"

namespace b200 {

int w4a16_gemv_max_m(int N);                                   // w4a16_gemv.cu
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
  // (x & 0x000f000f) | MAGIC as ONE LOP3 (mask and magic live in registers, LUT 0xEA = (a & b) | c)
  uint32_t mask = 0x000f000fu, magic = W4Magic<T>::kBits;
  asm("" : "+r"(mask), "+r"(magic));  // keep them out of the immediate field
  auto nib = [&](uint32_t x) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;\n" : "=r"(d) : "r"(x), "r"(mask), "r"(magic));
    return d;
  };
  const uint32_t w0 = nib(q), w1 = nib(q >> 4), w2 = nib(q >> 8), w3 = nib(q >> 12);
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
// Warp roles: 0 = TMA producer of activation tiles, 1 = MMA issuer, 2..5 = epilogue, 6..6+DQ-1 =
// dequant (int4 -> 16-bit -> swizzled MN-major operand), last warp = producer of the PACKED ring.
//   operand ring : stage = x tile (TMA) + dequantised weight tile (written by the dequant warps);
//                  `full` counts the TMA transaction plus one arrive per dequant warp of the owning
//                  group, `empty` is one tcgen05.commit.
//   packed ring  : the raw int4 words of a stage plus that stage's (zero, scale) rows, all by TMA and
//                  deeper than the operand ring, so the dequant warps only ever read shared memory
//                  (a global load in flight would stall the MEMBAR that fence.proxy.async implies).
// A stage spans KS sub-blocks of 64 K rows: every barrier hand-off (a fixed ~0.3 us round trip per
// role, measured) is amortised over KS * 64 rows.  The dequant warps form DG groups that own alternate
// stages.  BM = 256 (two 128-row accumulators sharing one dequantised stage) halves the dequant work
// per flop for large M; SWAP exchanges the operand roles for M <= 64 (see W4Cfg).
struct W4GemmParams {
  const uint32_t* qweight;  // [K, N/8] B200 order
  const void* zeros;        // [G, N] biased
  const void* scales;       // [G, N]
  const void* bias;         // [N] or null
  void* out;
  int64_t ldc;
  int M, N, K, group;
  int raw_tma;              // 1: packed tiles by TMA (N % 32 == 0), 0: by the producer warp's lanes
  int kb_per_split;         // SWAP kernels launched as clusters: stages per cluster rank (split-K)
};

constexpr int kW4PackedBytes = 64 * 64;  // one sub-block: 64 k rows x 16 int32 (= 128 output columns)
constexpr int kW4ZsBytes = 512;          // one sub-block's zero row + scale row, 256 B each

// BM = rows of x per tile (128 or 256); in SWAP mode BM = MT (32 or 64) tokens and the roles of the
// operands are exchanged: the dequantised weights are the 128-row MN-major A operand, x is the
// K-major B operand with UMMA N = MT, so the MMA time scales with the token count.
template <int BM, bool SWAP>
struct W4Cfg {
  static constexpr int BN = 128, BK = 64;
  static constexpr int KS = (!SWAP && BM == 256) ? 1 : 2;   // 64-row sub-blocks per stage
  static constexpr int XSUB = BM * 128, BSUB = BK * BN * 2;
  static constexpr int X_BYTES = KS * XSUB, B_BYTES = KS * BSUB, STAGE_BYTES = X_BYTES + B_BYTES;
  static constexpr int RAW_BYTES = KS * (kW4PackedBytes + kW4ZsBytes);
  static constexpr int ST = BM == 256 ? (SWAP ? 2 : 4) : 3;  // operand stages
  static constexpr int RAWS = SWAP ? (BM == 32 ? 10 : BM == 64 ? 8 : 3) : (BM == 256 ? 5 : 3);  // packed stages
  static constexpr int ACC_COLS = BM;  // fp32 columns per accumulator set
  static constexpr int TMEM_COLS = 2 * ACC_COLS < 32 ? 32 : 2 * ACC_COLS;
  static constexpr int NBARS = 2 * ST + 2 * RAWS + 4;
  static constexpr int SMEM = ST * STAGE_BYTES + RAWS * RAW_BYTES + NBARS * 8 + 16 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget");
};

template <typename T, int BM, bool SWAP, int DQ, int DG>
__global__ void __launch_bounds__((7 + DQ) * 32, 1)
w4a16_gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_q,
                  const __grid_constant__ CUtensorMap tmap_z, const __grid_constant__ CUtensorMap tmap_s,
                  const W4GemmParams p) {
  using C = W4Cfg<BM, SWAP>;
  constexpr int BN = C::BN, BK = C::BK, KS = C::KS, ST = C::ST, RAWS = C::RAWS;
  constexpr int MH = SWAP ? 1 : BM / 128;          // 128-row x boxes (and accumulators) per sub-block
  constexpr int XBOX = SWAP ? BM : 128;            // rows of one x TMA box
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t raw_base = smem_base + ST * C::STAGE_BYTES;
  const uint32_t bars = raw_base + RAWS * C::RAW_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (ST + s); };
  auto rfull_bar = [&](int s) { return bars + 8u * (2 * ST + s); };
  auto rempty_bar = [&](int s) { return bars + 8u * (2 * ST + RAWS + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * ST + 2 * RAWS + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * ST + 2 * RAWS + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * C::NBARS;
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int m_blocks = ceil_div(p.M, BM), n_blocks = ceil_div(p.N, BN);
  const int num_tiles = m_blocks * n_blocks;
  const int num_kb = ceil_div(p.K, BK * KS);       // stages per tile
  const int NC = p.N / 8;
  const bool zs_rows = (p.group % BK) == 0;        // one (zero, scale) row per sub-block
  // split-K (SWAP kernels only): a cluster of S CTAs shares one tile, rank r owns stages
  // [r * kb_per_split, ...); the partial sums meet in the epilogue over distributed shared memory.
  const uint32_t S = SWAP ? cluster_nctarank() : 1u;
  const uint32_t crank = SWAP ? cluster_ctarank() : 0u;
  const int tile0 = S > 1 ? static_cast<int>(blockIdx.x / S) : static_cast<int>(blockIdx.x);
  const int tile_step = S > 1 ? num_tiles : static_cast<int>(gridDim.x);
  const int kb_begin = S > 1 ? static_cast<int>(crank) * p.kb_per_split : 0;
  const int kb_end = S > 1 ? min(num_kb, kb_begin + p.kb_per_split) : num_kb;
  const int sub_per_group = zs_rows ? p.group / BK : 1;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmap_x);
    if (p.raw_tma) {
      tc::prefetch_tmap(&tmap_q);
      tc::prefetch_tmap(&tmap_z);
      tc::prefetch_tmap(&tmap_s);
    }
    for (int s = 0; s < ST; ++s) {
      tc::mbar_init(full_bar(s), 1 + DQ / DG);  // TMA expect_tx arrive + the owning group's warps
      tc::mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < RAWS; ++s) {
      tc::mbar_init(rfull_bar(s), 1);
      tc::mbar_init(rempty_bar(s), DQ / DG);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(tfull_bar(a), 1);
      tc::mbar_init(tempty_bar(a), 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
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
      for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        const int m_blk = tile % m_blocks;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          tc::mbar_wait(empty_bar(stage), phase ^ 1u);
          tc::mbar_arrive_expect_tx(full_bar(stage), C::X_BYTES);
#pragma unroll
          for (int j = 0; j < KS; ++j)
#pragma unroll
            for (int h = 0; h < MH; ++h)
              tc::tma_load_2d(smem_base + stage * C::STAGE_BYTES + (j * MH + h) * (XBOX * 128),
                              &tmap_x, full_bar(stage), (kb * KS + j) * BK, m_blk * BM + h * XBOX);
          if (++stage == ST) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 6 + DQ) {
    // ===================== producer of the packed ring =====================
    int rs = 0;
    uint32_t rphase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
      const int n_blk = tile / m_blocks;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        const uint32_t dst = raw_base + rs * C::RAW_BYTES;
        const uint32_t zdst = dst + KS * kW4PackedBytes;
        if (p.raw_tma) {
          if (lane == 0) {
            tc::mbar_wait(rempty_bar(rs), rphase ^ 1u);
            tc::mbar_arrive_expect_tx(rfull_bar(rs),
                                      KS * (kW4PackedBytes + (zs_rows ? kW4ZsBytes : 0)));
            tc::tma_load_2d_hint(dst, &tmap_q, rfull_bar(rs), n_blk * (BN / 8), kb * KS * BK,
                                 m_blocks > 1 ? tc::kEvictLast : tc::kEvictFirst);
            if (zs_rows) {
#pragma unroll
              for (int j = 0; j < KS; ++j) {
                const int g = (kb * KS + j) / sub_per_group;  // past the last group: TMA zero fill
                tc::tma_load_2d(zdst + j * kW4ZsBytes, &tmap_z, rfull_bar(rs), n_blk * BN, g);
                tc::tma_load_2d(zdst + j * kW4ZsBytes + 256, &tmap_s, rfull_bar(rs), n_blk * BN, g);
              }
            }
          }
        } else {
          // rows whose pitch is not a multiple of 16 bytes: the 32 lanes copy the tile word by word
          tc::mbar_wait(rempty_bar(rs), rphase ^ 1u);
#pragma unroll 1
          for (int j = 0; j < KS; ++j) {
            uint32_t v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int w = i * 32 + lane, r = w >> 4, cc = n_blk * (BN / 8) + (w & 15);
              const int k = (kb * KS + j) * BK + r;
              v[i] = (k < p.K && cc < NC) ? __ldg(p.qweight + static_cast<int64_t>(k) * NC + cc) : 0u;
            }
#pragma unroll
            for (int i = 0; i < 32; ++i)
              asm volatile("st.shared.u32 [%0], %1;\n" ::"r"(dst + j * kW4PackedBytes + (i * 32 + lane) * 4),
                           "r"(v[i])
                           : "memory");
            if (zs_rows) {  // lanes 0..15: zeros, 16..31: scales, 8 columns (16 bytes) each
              const int cc = n_blk * (BN / 8) + (lane & 15);
              const int g = (kb * KS + j) / sub_per_group;
              const uint4* src = static_cast<const uint4*>(lane < 16 ? p.zeros : p.scales);
              const uint4 zs = (cc < NC && static_cast<int64_t>(g) * p.group < p.K)
                                   ? __ldg(src + static_cast<int64_t>(g) * NC + cc)
                                   : make_uint4(0, 0, 0, 0);
              asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(zdst + j * kW4ZsBytes + lane * 16),
                           "r"(zs.x), "r"(zs.y), "r"(zs.z), "r"(zs.w)
                           : "memory");
            }
          }
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(rfull_bar(rs));
        }
        if (++rs == RAWS) {
          rs = 0;
          rphase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      // normal: A = x (K-major, 128 rows), B = weights (MN-major, N = 128)
      // swap  : A = weights (MN-major, M = 128 output columns), B = x (K-major, N = MT tokens)
      constexpr uint32_t idesc = SWAP ? tc::make_idesc(fmt, fmt, 128, BM, 1, 0)
                                      : tc::make_idesc(fmt, fmt, 128, BN, 0, 1);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_step, ++local) {
        const int acc = local & 1;
        tc::mbar_wait(tempty_bar(acc), ((local >> 1) & 1u) ^ 1u);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * C::ACC_COLS;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          tc::mbar_wait(full_bar(stage), phase);
          tc::tc_fence_after();
          const uint32_t sx = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sw = sx + C::X_BYTES;
#pragma unroll
          for (int j = 0; j < KS; ++j) {
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk) {
              // weight rows kk*16.. of a [64 k x 128 n] sub-tile; panels of 64 n are 64*128 B apart
              const uint64_t wdesc =
                  tc::make_smem_desc_sw128(sw + j * C::BSUB + kk * 16 * 128, 1024, BK * 128);
#pragma unroll
              for (int h = 0; h < MH; ++h) {
                const uint64_t xdesc = tc::make_smem_desc_sw128(
                    sx + (j * MH + h) * (XBOX * 128) + kk * 32, 1024, 16);
                const uint32_t accum = ((kb - kb_begin) | j | kk) != 0 ? 1u : 0u;
                if (SWAP)
                  tc::mma_f16(tmem_d, wdesc, xdesc, idesc, accum);
                else
                  tc::mma_f16(tmem_d + h * BN, xdesc, wdesc, idesc, accum);
              }
            }
          }
          tc::mma_commit(empty_bar(stage));
          if (++stage == ST) {
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
    for (int tile = tile0; tile < num_tiles; tile += tile_step, ++local) {
      const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
      const int acc = local & 1;
      tc::mbar_wait(tfull_bar(acc), (local >> 1) & 1u);
      tc::tc_fence_after();
      if (SWAP) {
        // accumulator lanes = output columns, TMEM columns = tokens
        const int ch = quarter * 32 + lane;
        const int n = n_blk * BN + ch;
        const float bv = (bias && n < p.N) ? to_float<T>(bias[n]) : 0.f;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * C::ACC_COLS;
        if (S == 1) {
#pragma unroll 1
          for (int c = 0; c < BM / 32; ++c) {
            if (c * 32 >= p.M) break;
            uint32_t r[32];
            tc::tmem_ld32(lane_addr + c * 32, r);
            tc::tmem_ld_wait();
            if (n < p.N) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const int m = c * 32 + j;
                if (m < p.M) out[static_cast<int64_t>(m) * p.ldc + n] = from_float<T>(__uint_as_float(r[j]) + bv);
              }
            }
          }
        } else {
          // fixed-order cluster reduction (same scheme as fp8_gemm_smallm.cuh): after barrier 1 every
          // peer's operand ring is idle; token m of my partial is pushed to rank m % S, slot
          // [my rank][m / S][channel] of ITS ring; after barrier 2 each rank sums the S slots of its
          // own tokens from local shared memory in rank order (deterministic) and stores.
          const int tpr = ceil_div(p.M, static_cast<int>(S));
          const uint32_t push_base = smem_base + (crank * tpr * 128 + ch) * 4;
          __syncwarp();
          cluster_sync_all();  // barrier 1
          uint32_t dst_rank = 0, dst_slot = 0;
#pragma unroll 1
          for (int c = 0; c < BM / 32; ++c) {
            if (c * 32 >= p.M) break;
            uint32_t r[32];
            tc::tmem_ld32(lane_addr + c * 32, r);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (c * 32 + j < p.M) {
                const uint32_t a = map_to_rank(push_base + dst_slot * 512, dst_rank);
                asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(a), "f"(__uint_as_float(r[j])) : "memory");
                if (++dst_rank == S) {
                  dst_rank = 0;
                  ++dst_slot;
                }
              }
            }
          }
          __syncwarp();
          cluster_sync_all();  // barrier 2: every peer's pushes into my ring are visible
          const float* part = reinterpret_cast<const float*>(smem_raw + (smem_base - smem_u32(smem_raw))) + ch;
          int m = static_cast<int>(crank);
          for (int j = 0; j < tpr; ++j, m += S) {
            if (m >= p.M) break;
            float sum = 0.f;
            for (uint32_t q = 0; q < S; ++q) sum += part[(q * tpr + j) * 128];
            if (n < p.N) out[static_cast<int64_t>(m) * p.ldc + n] = from_float<T>(sum + bv);
          }
        }
      } else
#pragma unroll 1
      for (int h = 0; h < MH; ++h) {
        const int row = m_blk * BM + h * 128 + quarter * 32 + lane;
        const bool row_ok = row < p.M;
        T* orow = out + static_cast<int64_t>(row) * p.ldc;
        if (m_blk * BM + h * 128 >= p.M) break;  // whole half is padding (warp-uniform)
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tc::tmem_ld32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                            acc * C::ACC_COLS + h * BN + c * 32,
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
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tempty_bar(acc));
    }
  } else {
    // ===================== dequant warps (6..6+DQ-1): packed ring -> 16-bit swizzled weight tile ====
    // The warps form DG groups; group j owns every DG-th stage of this CTA's stage sequence (the
    // sequence runs on across tiles), so the latency chain of one group's stage (barrier polls,
    // ld.shared, fence.proxy.async) overlaps with the other groups' stages.
    constexpr int NT = (DQ / DG) * 32;    // threads per group
    constexpr int RSTEP = NT / 16;        // row stride of one thread's rows inside a sub-block
    constexpr int RPT = BK / RSTEP;       // rows per thread per sub-block
    static_assert(DQ % DG == 0 && RSTEP % 8 == 0, "the swizzle term must not depend on the row index i");
    static_assert(DG <= ST && DG <= RAWS, "a group advances its ring position by DG with one wrap");
    const int grp = (warp - 6) / (DQ / DG);
    const int t = threadIdx.x - 6 * 32 - grp * NT;
    const int col = t & 15;               // int32 column inside the 128-wide n tile
    const int row0 = t >> 4;              // rows row0 + RSTEP * i
    const uint4* zeros = static_cast<const uint4*>(p.zeros);    // rows of NC uint4 (8 columns each)
    const uint4* scales = static_cast<const uint4*>(p.scales);
    // byte offset of this thread's 16-byte chunk of row row0 inside a weight sub-tile; rows
    // row0 + RSTEP*i keep the same swizzle term because RSTEP is a multiple of 8
    const uint32_t b_off =
        C::X_BYTES + (col >> 3) * (BK * 128) + row0 * 128 + (((col & 7) ^ (row0 & 7)) << 4);
    const uint32_t r_off = row0 * 64 + col * 4;
    int stage = grp, rs = grp;            // ring positions of this group's next stage
    uint32_t phase = 0, rphase = 0;
    int tile = tile0, kb = kb_begin + grp;
    while (true) {
      while (kb >= kb_end) {
        kb -= kb_end - kb_begin;
        tile += tile_step;
      }
      if (tile >= num_tiles) break;
      tc::mbar_wait(rfull_bar(rs), rphase);
      const uint32_t rsrc = raw_base + rs * C::RAW_BYTES;
      uint32_t q[KS][RPT];
      uint4 zb[KS], sc[KS];
#pragma unroll
      for (int j = 0; j < KS; ++j) {
#pragma unroll
        for (int i = 0; i < RPT; ++i)
          asm volatile("ld.shared.u32 %0, [%1];\n"
                       : "=r"(q[j][i])
                       : "r"(rsrc + j * kW4PackedBytes + r_off + RSTEP * i * 64));
        zb[j] = sc[j] = make_uint4(0, 0, 0, 0);
        if (zs_rows) {  // the sub-block's (zero, scale) rows travel with the packed tile
          const uint32_t zsrc = rsrc + KS * kW4PackedBytes + j * kW4ZsBytes + col * 16;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                       : "=r"(zb[j].x), "=r"(zb[j].y), "=r"(zb[j].z), "=r"(zb[j].w)
                       : "r"(zsrc));
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                       : "=r"(sc[j].x), "=r"(sc[j].y), "=r"(sc[j].z), "=r"(sc[j].w)
                       : "r"(zsrc + 256));
        }
      }
      // The slot is refilled by TMA (async proxy) as soon as it is released, and the mbarrier release
      // only orders this thread's generic-proxy accesses: without a proxy fence the refill may
      // overtake the ld.shared above (write-after-read across proxies).  It showed whenever the
      // MMA pipe, not the dequant warps, was the bottleneck (256-row tiles with K >= 3072, second
      // tiles of persistent CTAs): a few stale weights per stage, different on every run.
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(rempty_bar(rs));
      rs += DG;
      if (rs >= RAWS) {
        rs -= RAWS;
        rphase ^= 1u;
      }
      const uint32_t dst = smem_base + stage * C::STAGE_BYTES + b_off;
      if (zs_rows) {
        uint4 w[RPT];
#pragma unroll
        for (int i = 0; i < RPT; ++i) w[i] = dequant8<T>(q[0][i], zb[0], sc[0]);
        tc::mbar_wait(empty_bar(stage), phase ^ 1u);
#pragma unroll
        for (int j = 0; j < KS; ++j) {
#pragma unroll
          for (int i = 0; i < RPT; ++i)
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(dst + j * C::BSUB + RSTEP * i * 128),
                         "r"(w[i].x), "r"(w[i].y), "r"(w[i].z), "r"(w[i].w)
                         : "memory");
          if (j + 1 < KS) {
#pragma unroll
            for (int i = 0; i < RPT; ++i) w[i] = dequant8<T>(q[(j + 1) % KS][i], zb[(j + 1) % KS], sc[(j + 1) % KS]);
          }
        }
      } else {
        // groups smaller than / not aligned to a sub-block: look the group up per row
        const int c = (tile / m_blocks) * (BN / 8) + col;  // global int32 column
        const int cc = c < NC ? c : NC - 1;  // clamped: columns >= N are computed but never stored
        int g_cached = -1;
        uint4 zg = make_uint4(0, 0, 0, 0), sg = zg;
        tc::mbar_wait(empty_bar(stage), phase ^ 1u);
#pragma unroll
        for (int j = 0; j < KS; ++j) {
#pragma unroll
          for (int i = 0; i < RPT; ++i) {
            const int k = (kb * KS + j) * BK + row0 + RSTEP * i;
            uint4 w = make_uint4(0, 0, 0, 0);
            if (k < p.K) {  // rows past K must be exact zeros (x is zero-filled there, 0 * garbage != 0)
              const int g = k / p.group;
              if (g != g_cached) {
                g_cached = g;
                zg = __ldg(zeros + static_cast<int64_t>(g) * NC + cc);
                sg = __ldg(scales + static_cast<int64_t>(g) * NC + cc);
              }
              w = dequant8<T>(q[j][i], zg, sg);
            }
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(dst + j * C::BSUB + RSTEP * i * 128),
                         "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w)
                         : "memory");
          }
        }
      }
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(full_bar(stage));
      stage += DG;
      if (stage >= ST) {
        stage -= ST;
        phase ^= 1u;
      }
      kb += DG;
    }
  }

  if (SWAP && S > 1 && (warp < 2 || warp >= 6)) {  // everyone but the epilogue warps (done above)
    __syncwarp();
    cluster_sync_all();
    cluster_sync_all();
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

// Tile selection.  M <= 64: swap-AB (tokens are the UMMA N dimension).  Otherwise 128-row tiles, or
// 256-row tiles (half the dequant work per flop) once they still fill the machine.
// B200_W4_BM=32|64|129|257 (swap) |128|256 overrides (sweeps).
static int w4_pick_bm(int M, int N, int K, int sms) {
  static int forced = [] {
    const char* v = std::getenv("B200_W4_BM");
    return v && *v ? std::atoi(v) : 0;
  }();
  if ((forced == 32 || forced == 64) && M <= forced) return forced;
  if (forced == 129 && M <= 128) return forced;
  if (forced == 257 && M <= 256) return forced;
  if (forced == 128 || forced == 256) return forced;
  if (M <= 32) return 32;
  if (M <= 64) return 64;
  if (M <= 128) return ceil_div(N, 128) <= 64 ? 129 : 128;  // 129 = swap-AB with 128 tokens (split-K)
  // 257 = swap-AB with 256 tokens (split-K): only two operand stages fit, which pays for deep K only
  // (down_proj 5120 x 13824: 64.5 -> 50-52 us; 5120 x 2560: 17 -> 25-30 us, so not there)
  if (M <= 256 && ceil_div(N, 128) <= 40 && K >= 8192) return 257;
  return ceil_div(M, 256) * ceil_div(N, 128) >= (sms * 3) / 4 ? 256 : 128;
}

template <typename T, int BM, bool SWAP, int DQ, int DG>
static int launch_w4a16_gemm(const void* x, int64_t ldx, W4GemmParams p, cudaStream_t st) {
  using C = W4Cfg<BM, SWAP>;
  CUtensorMap tx, tq;
  int rc = make_tmap_2d(&tx, x, p.M, p.K, ldx * 2, SWAP ? BM : 128, 64, 2, false);
  if (rc) return rc;
  p.raw_tma = (p.N % 32 == 0 && aligned16(p.qweight)) ? 1 : 0;
  if (p.raw_tma) {
    rc = make_tmap_2d_plain(&tq, p.qweight, p.K, p.N / 8, static_cast<uint64_t>(p.N / 8) * 4,
                            C::KS * 64, 16, 4);
    if (rc) return rc;
  } else {
    tq = tx;
  }
  CUtensorMap tz = tx, ts = tx;
  if (p.raw_tma && p.group % 64 == 0) {
    const uint64_t G = static_cast<uint64_t>(p.K / p.group);
    rc = make_tmap_2d_plain(&tz, p.zeros, G, p.N, static_cast<uint64_t>(p.N) * 2, 1, 128, 2);
    if (rc) return rc;
    rc = make_tmap_2d_plain(&ts, p.scales, G, p.N, static_cast<uint64_t>(p.N) * 2, 1, 128, 2);
    if (rc) return rc;
  }
  auto kern = w4a16_gemm_kernel<T, BM, SWAP, DQ, DG>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    attr_done = true;
  }
  const int tiles = ceil_div(p.M, BM) * ceil_div(p.N, 128);
  const int sms = device_sm_count();
  // split-K for the swap-AB kernels when the tiles alone leave most SMs idle: the largest S <= 8
  // whose `tiles` clusters are co-resident in one wave with at least four stages per rank.
  // B200_W4_SPLITS forces it.
  int S = 1;
  if (SWAP) {
    static const int forced = [] {
      const char* v = std::getenv("B200_W4_SPLITS");
      return v && *v ? std::atoi(v) : 0;
    }();
    const int num_kb = ceil_div(p.K, 64 * C::KS);
    if (forced > 0) {
      S = forced;
    } else {
      // clusters are placed inside one GPC and the SMs per GPC differ between boards (the same
      // 40-tile x 3 launch ran in one wave on one B200 and two on another: 27.5 vs 39.7 us), so
      // the driver is asked how many clusters of each size are co-resident; answers are cached
      static int max_clusters[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int c = 8; c >= 2; --c) {
        if (num_kb < 4 * c || tiles > sms / c) continue;
        if (max_clusters[c] == 0) {
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(c * 8);
          cfg.blockDim = dim3((7 + DQ) * 32);
          cfg.dynamicSmemBytes = C::SMEM;
          cudaLaunchAttribute attr[1];
          attr[0].id = cudaLaunchAttributeClusterDimension;
          attr[0].val.clusterDim.x = c;
          attr[0].val.clusterDim.y = 1;
          attr[0].val.clusterDim.z = 1;
          cfg.attrs = attr;
          cfg.numAttrs = 1;
          int n = 0;
          if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) {
            (void)cudaGetLastError();
            n = -1;
          }
          max_clusters[c] = n;
        }
        if (max_clusters[c] >= tiles) {
          S = c;
          break;
        }
      }
    }
    while (S > 1 && (S - 1) * ceil_div(num_kb, S) >= num_kb) --S;
    p.kb_per_split = ceil_div(num_kb, S);
  }
  note_variant("w4a16_gemm bm=%d swap=%d splits=%d tiles=%d", BM, SWAP ? 1 : 0, S, tiles);
  {
    ScopedKernelTimer timer("w4a16_gemm", st);
    if (S > 1) {
      B200_CUDA_OK(launch_kernel(kern, dim3(tiles * S), dim3((7 + DQ) * 32), C::SMEM, st, S, 0, tx, tq, tz, ts, p));
    } else {
      kern<<<tiles < sms ? tiles : sms, (7 + DQ) * 32, C::SMEM, st>>>(tx, tq, tz, ts, p);
    }
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
  if (M <= w4a16_gemv_max_m(N) && w4a16_gemv_supported(M, N, K, group_size) && aligned16(qweight))
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
  p.raw_tma = 0;
  p.kb_per_split = 0;
  const int bm = w4_pick_bm(M, N, K, device_sm_count());
  // dequant warp layout: 8 warps in DG groups that own alternate stages.  B200_W4_CFG: 0 = one
  // group, 1 = two groups (default)
  static const int cfg = [] {
    const char* v = std::getenv("B200_W4_CFG");
    return v && *v ? std::atoi(v) : 1;
  }();
#define B200_W4_LAUNCH_C(TT, DQ_, DG_)                                                  \
  (bm == 256  ? launch_w4a16_gemm<TT, 256, false, DQ_, DG_>(x, ldx, p, st)              \
   : bm == 128 ? launch_w4a16_gemm<TT, 128, false, DQ_, DG_>(x, ldx, p, st)             \
   : bm == 129 ? launch_w4a16_gemm<TT, 128, true, DQ_, DG_>(x, ldx, p, st)              \
   : bm == 257 ? launch_w4a16_gemm<TT, 256, true, DQ_, DG_>(x, ldx, p, st)              \
   : bm == 64  ? launch_w4a16_gemm<TT, 64, true, DQ_, DG_>(x, ldx, p, st)               \
               : launch_w4a16_gemm<TT, 32, true, DQ_, DG_>(x, ldx, p, st))
#define B200_W4_LAUNCH(TT) (cfg == 0 ? B200_W4_LAUNCH_C(TT, 8, 1) : B200_W4_LAUNCH_C(TT, 8, 2))
  return dtype == B200_F16 ? B200_W4_LAUNCH(__half) : B200_W4_LAUNCH(__nv_bfloat16);
#undef B200_W4_LAUNCH_C
#undef B200_W4_LAUNCH
}

}  // extern "C"
