// Persistent warp-specialised tcgen05 GEMM for sm_100a:  C[M,N] = epilogue( A[M,K] . B[N,K]^T )
// Both operands are K-major ("TN"): A rows are tokens, B rows are output channels — exactly how the
// reference stores activations and (transposed-view) weights (fp8.py:349,391; fp8_utils.py:636-648).
//
//   warp 0   : TMA producer   — cp.async.bulk.tensor 128-byte-swizzled A/B tiles into a STAGES ring
//   warp 1   : MMA issuer     — one elected thread issues tcgen05.mma (M=128, N=BLOCK_N, K=32 B),
//                               accumulators live in TMEM (2 x BLOCK_N fp32 columns, double buffered)
//   warps 2-5: epilogue       — tcgen05.ld 32 lanes x 32 columns, scale / bias, cast, 16-byte stores
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue) and the
// static persistent tile schedule (tile = blockIdx.x + i * gridDim.x, M fastest so concurrently
// running CTAs share the same weight tile through L2).
#pragma once
#include <type_traits>

#include "tc_common.cuh"

namespace b200 {

enum GemmKind { kKindFp8 = 0, kKindBf16 = 1, kKindF16 = 2 };

struct GemmEpilogueParams {
  void* out;            // [M, N] OutT, row stride ldc
  int64_t ldc;
  const float* sa;      // per-row scale (stride sa_stride: 1 = per token, 0 = per tensor) or null
  const float* sb;      // per-column scale (stride sb_stride) or null
  int sa_stride, sb_stride;
  const void* bias;     // [N] OutT or null
  int M, N, K;
  // Split-K (mid-M regime: too few output tiles to occupy the SMs, see launch_gemm_tn): the
  // kernel is launched as thread-block clusters of `splits` CTAs, cluster rank = k-slice of ONE
  // output tile.  Every CTA parks its fp32 partial tile in its own (now idle) pipeline ring, then
  // rank r sums the `splits` partials of ITS share of the columns over distributed shared memory
  // in rank order — deterministic, no global workspace, no atomics — and runs the normal epilogue
  // on that share.  splits == 1: plain persistent path.
  int splits = 1;
  int kb_per_split = 0;
};

template <int KIND>
struct GemmKindTraits;
template <>
struct GemmKindTraits<kKindFp8> {
  static constexpr int ELT = 1;
  static constexpr uint32_t FMT = 0;  // E4M3
};
template <>
struct GemmKindTraits<kKindBf16> {
  static constexpr int ELT = 2;
  static constexpr uint32_t FMT = 1;  // BF16
};
template <>
struct GemmKindTraits<kKindF16> {
  static constexpr int ELT = 2;
  static constexpr uint32_t FMT = 0;  // F16
};

template <int KIND, int BLOCK_N, int STAGES>
struct GemmCfg {
  static constexpr int BLOCK_M = 128;
  static constexpr int ELT = GemmKindTraits<KIND>::ELT;
  static constexpr int BLOCK_K = 128 / ELT;        // elements per 128-byte swizzle row
  static constexpr int UMMA_K_BYTES = 32;
  static constexpr int A_BYTES = BLOCK_M * 128;
  static constexpr int B_BYTES = BLOCK_N * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = 2 * BLOCK_N;    // power of two for BLOCK_N in {64,128,256}
  static constexpr int THREADS = 192;
  static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // + alignment slack
  static_assert(BLOCK_N == 64 || BLOCK_N == 128 || BLOCK_N == 256, "BLOCK_N");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ void named_bar_sync_gemm(int id, int threads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(threads) : "memory");
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n"
               "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}

__device__ __forceinline__ float ld_cluster_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

template <typename OutT>
__device__ __forceinline__ uint32_t pack_out2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack_out2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack_out2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// MC (large M): CTA pairs (cluster of 2 along M, same N block) share the weight tile — each CTA
// loads HALF of the B rows and TMA-multicasts them into both CTAs, cutting the per-SM L2->SM
// ingest from 48 KiB to 32 KiB per K block (the measured limiter of the 128x256 tile).  A stage
// may be refilled only when BOTH CTAs' MMAs have consumed it: tcgen05.commit multicasts its arrive
// to both empty barriers (count 2).  tmap_b then describes BLOCK_N/2-row boxes.
template <int KIND, int BLOCK_N, int STAGES, typename OutT, bool MC = false>
__global__ void __launch_bounds__(192, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmEpilogueParams ep) {
  using C = GemmCfg<KIND, BLOCK_N, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  // per-column scale and bias of the tile being finished (double-buffered per TMEM accumulator):
  // staged BEFORE the accumulator wait.  Loading them inside the store loop serialised one L2/L1
  // round trip per 8 columns — loads may not be hoisted above the preceding output stores (possible
  // aliasing) — which the mid-M single-tile launches paid on their critical path.
  __shared__ float sb_s[2][BLOCK_N];
  __shared__ float bias_s[2][BLOCK_N];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B needs 1024 B
  const uint32_t bars = smem_base + STAGES * C::STAGE_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(
      smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const int m_blocks = ceil_div(ep.M, C::BLOCK_M);
  const int n_blocks = ceil_div(ep.N, BLOCK_N);
  const int num_tiles = m_blocks * n_blocks;
  const int num_kb = ceil_div(ep.K, C::BLOCK_K);
  const int S = ep.splits;
  // MC: items are PAIR tiles (two adjacent M blocks x one N block); this CTA takes M block
  // 2*pair + cluster rank.  Item stride is the number of clusters.
  const int mc_rank = MC ? static_cast<int>(cluster_ctarank()) : 0;
  const int num_items = MC ? (m_blocks / 2) * n_blocks : num_tiles * S;
  const int item0 = MC ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int item_step = MC ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmap_a);
    tc::prefetch_tmap(&tmap_b);
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(full_bar(s), 1);
      tc::mbar_init(empty_bar(s), MC ? 2 : 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(tfull_bar(a), 1);
      tc::mbar_init(tempty_bar(a), 4);  // one arrive per epilogue warp
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  if (MC) cluster_sync_all();  // the peer's barriers are initialised before anything targets them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // Programmatic dependent launch: everything a previous kernel may have written (activations,
  // scales, outputs) is touched only after griddep_wait().  The B operand is a WEIGHT in every use of
  // this kernel, so the producer thread puts the first ring stages of its first tile's B in flight
  // before it waits: the TMA fill then overlaps the previous kernel's tail (the mid-M launches are
  // a chain of fixed latencies, profiles/r02_ncu_fp8_smallm_m64_source.md).
  int b_prefetched = 0;
  if (warp == 0 && lane == 0) {
    if (!MC && item0 < num_items) {
      const int tile = item0 / S, ks = item0 - tile * S;
      const int n_blk = tile / m_blocks;
      const int kb0 = S > 1 ? ks * ep.kb_per_split : 0;
      const int kb1 = S > 1 ? min(num_kb, kb0 + ep.kb_per_split) : num_kb;
      b_prefetched = min(STAGES, kb1 - kb0);
      for (int i = 0; i < b_prefetched; ++i) {
        tc::mbar_arrive_expect_tx(full_bar(i), C::STAGE_BYTES);
        tc::tma_load_2d(smem_base + i * C::STAGE_BYTES + C::A_BYTES, &tmap_b, full_bar(i),
                        (kb0 + i) * C::BLOCK_K, n_blk * BLOCK_N);
      }
    }
  }
  griddep_wait();  // activations / scales come from the previous kernel
  griddep_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = item0; item < num_items; item += item_step) {
        const int tile = MC ? item : item / S, ks = MC ? 0 : item - tile * S;
        const int m_blk = MC ? 2 * (tile % (m_blocks / 2)) + mc_rank : tile % m_blocks;
        const int n_blk = MC ? tile / (m_blocks / 2) : tile / m_blocks;
        const int kb0 = S > 1 ? ks * ep.kb_per_split : 0;
        const int kb1 = S > 1 ? min(num_kb, kb0 + ep.kb_per_split) : num_kb;
        for (int kb = kb0; kb < kb1; ++kb) {
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          if (b_prefetched > 0) {  // first stages of the first tile: B is already on its way
            --b_prefetched;
            tc::tma_load_2d(sa, &tmap_a, full_bar(stage), kb * C::BLOCK_K, m_blk * C::BLOCK_M);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
            continue;
          }
          tc::mbar_wait(empty_bar(stage), phase ^ 1u);
          tc::mbar_arrive_expect_tx(full_bar(stage), C::STAGE_BYTES);
          tc::tma_load_2d(sa, &tmap_a, full_bar(stage), kb * C::BLOCK_K, m_blk * C::BLOCK_M);
          if (MC)
            tc::tma_load_2d_mcast(sa + C::A_BYTES + mc_rank * (C::B_BYTES / 2), &tmap_b,
                                  full_bar(stage), kb * C::BLOCK_K,
                                  n_blk * BLOCK_N + mc_rank * (BLOCK_N / 2), 0x3);
          else
          tc::tma_load_2d(sa + C::A_BYTES, &tmap_b, full_bar(stage), kb * C::BLOCK_K,
                          n_blk * BLOCK_N);
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
      constexpr uint32_t idesc = tc::make_idesc(GemmKindTraits<KIND>::FMT, GemmKindTraits<KIND>::FMT,
                                                C::BLOCK_M, BLOCK_N, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int item = item0; item < num_items; item += item_step, ++local) {
        const int ks = MC ? 0 : item % S;
        const int kb0 = S > 1 ? ks * ep.kb_per_split : 0;
        const int kb1 = S > 1 ? min(num_kb, kb0 + ep.kb_per_split) : num_kb;
        const int acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1u;
        tc::mbar_wait(tempty_bar(acc), acc_phase ^ 1u);  // epilogue drained this accumulator
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(full_bar(stage), phase);
          tc::tc_fence_after();
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sb = sa + C::A_BYTES;
#pragma unroll
          for (int k = 0; k < 128 / C::UMMA_K_BYTES; ++k) {
            const uint64_t adesc = tc::make_smem_desc_sw128(sa + k * C::UMMA_K_BYTES, 1024, 16);
            const uint64_t bdesc = tc::make_smem_desc_sw128(sb + k * C::UMMA_K_BYTES, 1024, 16);
            if (KIND == kKindFp8)
              tc::mma_f8f6f4(tmem_d, adesc, bdesc, idesc, ((kb - kb0) | k) != 0 ? 1u : 0u);
            else
              tc::mma_f16(tmem_d, adesc, bdesc, idesc, ((kb - kb0) | k) != 0 ? 1u : 0u);
          }
          if (MC) tc::mma_commit_mcast(empty_bar(stage), 0x3);  // frees the slot in BOTH CTAs
          else
          tc::mma_commit(empty_bar(stage));  // smem slot free once these MMAs have read it
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
        tc::mma_commit(tfull_bar(acc));  // accumulator complete
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    OutT* out = static_cast<OutT*>(ep.out);
    const OutT* bias = static_cast<const OutT*>(ep.bias);
    int local = 0;
    for (int item = item0; item < num_items; item += item_step, ++local) {
      const int tile = MC ? item : item / S, ks = MC ? 0 : item - tile * S;
      const int m_blk = MC ? 2 * (tile % (m_blocks / 2)) + mc_rank : tile % m_blocks;
      const int n_blk = MC ? tile / (m_blocks / 2) : tile / m_blocks;
      const int acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1u;
      for (int cidx = quarter * 32 + lane; cidx < BLOCK_N; cidx += 128) {
        const int col = n_blk * BLOCK_N + cidx;
        const bool ok = col < ep.N;
        sb_s[acc][cidx] = (ep.sb && ok) ? __ldg(ep.sb + static_cast<int64_t>(col) * ep.sb_stride) : 1.0f;
        bias_s[acc][cidx] = (bias && ok) ? to_float<OutT>(bias[col]) : 0.0f;
      }
      named_bar_sync_gemm(1, 128);  // the four epilogue warps; a thread gets here only after it has
                                    // finished the tile that used this buffer two tiles ago
      tc::mbar_wait(tfull_bar(acc), acc_phase);
      tc::tc_fence_after();
      const float* sbt = sb_s[acc];
      const float* bst = bias_s[acc];
      const int row = m_blk * C::BLOCK_M + quarter * 32 + lane;
      const bool row_ok = row < ep.M;
      const float sa = (ep.sa && row_ok) ? __ldg(ep.sa + static_cast<int64_t>(row) * ep.sa_stride)
                                         : 1.0f;
      OutT* orow = out + static_cast<int64_t>(row) * ep.ldc;
      if (S > 1) {
        // 1. park this k-slice's fp32 partial in the idle ring: part[col][row] (conflict-free)
        const uint32_t part = smem_base + (quarter * 32 + lane) * 4;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
          uint32_t r[32];
          tc::tmem_ld32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N +
                            c * 32, r);
          tc::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j)
            asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(part + (c * 32 + j) * 512), "r"(r[j])
                         : "memory");
        }
        tc::tc_fence_before();
        __syncwarp();
        cluster_sync_all();  // every partial of this tile is in place (all 192 threads arrive)
        // 2. my share of the columns, in groups of 8: ranks 0..rem-1 take one extra group
        constexpr int GROUPS = BLOCK_N / 8;
        const int base_g = GROUPS / S, rem = GROUPS % S;
        const int my_rank = static_cast<int>(cluster_ctarank());
        const int g0 = my_rank * base_g + min(my_rank, rem);
        const int g1 = g0 + base_g + (my_rank < rem ? 1 : 0);
        for (int g = g0; g < g1; ++g) {
          float sum[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) sum[j] = 0.f;
          for (int q = 0; q < S; ++q) {  // rank order
            const uint32_t src = map_to_rank(part + g * 8 * 512, q);
            float t[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) t[j] = ld_cluster_f32(src + j * 512);
#pragma unroll
            for (int j = 0; j < 8; ++j) sum[j] += t[j];
          }
          const int col0 = n_blk * BLOCK_N + g * 8;
          if (row_ok && col0 < ep.N) {
            if (col0 + 8 <= ep.N && (ep.ldc & 7) == 0) {
              uint32_t packed[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float s0 = sa * sbt[g * 8 + 2 * e], s1 = sa * sbt[g * 8 + 2 * e + 1];
                const float b0 = bst[g * 8 + 2 * e], b1 = bst[g * 8 + 2 * e + 1];
                packed[e] = pack_out2<OutT>(fmaf(sum[2 * e], s0, b0), fmaf(sum[2 * e + 1], s1, b1));
              }
              *reinterpret_cast<uint4*>(orow + col0) =
                  make_uint4(packed[0], packed[1], packed[2], packed[3]);
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (col0 + j >= ep.N) break;
                const float sc = sa * sbt[g * 8 + j];
                const float bb = bst[g * 8 + j];
                orow[col0 + j] = from_float<OutT>(fmaf(sum[j], sc, bb));
              }
            }
          }
        }
        __syncwarp();
        cluster_sync_all();  // peers have finished reading my partial
        continue;
      }
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t r[32];
        const uint32_t taddr =
            tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N + c * 32;
        tc::tmem_ld32(taddr, r);
        tc::tmem_ld_wait();
        const int col0 = n_blk * BLOCK_N + c * 32;
        if (row_ok && col0 < ep.N) {
          if (col0 + 32 <= ep.N && (ep.ldc & 7) == 0) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {  // 4 x 16-byte stores of 8 outputs
              uint32_t packed[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int j = v * 8 + e * 2;
                float x0 = __uint_as_float(r[j]), x1 = __uint_as_float(r[j + 1]);
                const float s0 = sa * sbt[c * 32 + j], s1 = sa * sbt[c * 32 + j + 1];
                const float b0 = bst[c * 32 + j], b1 = bst[c * 32 + j + 1];
                packed[e] = pack_out2<OutT>(fmaf(x0, s0, b0), fmaf(x1, s1, b1));
              }
              *reinterpret_cast<uint4*>(orow + col0 + v * 8) =
                  make_uint4(packed[0], packed[1], packed[2], packed[3]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (col0 + j >= ep.N) break;
              const float s = sa * sbt[c * 32 + j];
              const float b = bst[c * 32 + j];
              orow[col0 + j] = from_float<OutT>(fmaf(__uint_as_float(r[j]), s, b));
            }
          }
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tempty_bar(acc));
    }
  }
  if (S > 1 && warp < 2) {  // producer / MMA warps take part in the two cluster barriers
    __syncwarp();
    cluster_sync_all();
    cluster_sync_all();
  }
  if (MC) {  // the peer may still multicast into / arrive on this CTA's shared memory
    __syncwarp();
    cluster_sync_all();
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

// Split factor for `tiles` output tiles of `num_kb` K blocks on `sms` SMs (1 = no split).
int splitk_pick(int tiles, int num_kb, int sms);

// Host launcher: builds the two TMA descriptors and launches one CTA per SM (or per tile).
template <int KIND, int BLOCK_N, int STAGES, typename OutT, bool MC = false>
int launch_gemm_tn(const void* a, int64_t lda, const void* b, int64_t ldb,
                   const GemmEpilogueParams& ep_in, cudaStream_t st, const char* timer_name,
                   bool allow_split = false) {
  using C = GemmCfg<KIND, BLOCK_N, STAGES>;
  GemmEpilogueParams ep = ep_in;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d(&ta, a, ep.M, ep.K, lda * C::ELT, C::BLOCK_M, C::BLOCK_K, C::ELT, false);
  if (rc) return rc;
  rc = make_tmap_2d(&tb, b, ep.N, ep.K, ldb * C::ELT, MC ? BLOCK_N / 2 : BLOCK_N, C::BLOCK_K, C::ELT, true);
  if (rc) return rc;
  auto kern = gemm_tn_kernel<KIND, BLOCK_N, STAGES, OutT, MC>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  const int tiles = ceil_div(ep.M, C::BLOCK_M) * ceil_div(ep.N, BLOCK_N);
  const int sms = device_sm_count();
  ep.splits = 1;
  if (allow_split && BLOCK_N * 128 * 4 <= STAGES * C::STAGE_BYTES) {
    const int num_kb = ceil_div(ep.K, C::BLOCK_K);
    // largest split whose `tiles` clusters are all co-resident (one wave): cluster placement is
    // per GPC, so e.g. 48 clusters of 3 do not fit although 144 <= 148 SMs (measured: 24 us in two
    // waves vs 16 us unsplit).  Occupancy answers are cached per split factor.
    static int max_clusters[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int S = splitk_pick(tiles, num_kb, sms); S > 1; --S) {
      if ((S - 1) * ceil_div(num_kb, S) >= num_kb) continue;  // would leave an empty slice
      if (max_clusters[S] == 0) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(S * 8);
        cfg.blockDim = dim3(C::THREADS);
        cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = S;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) {
          (void)cudaGetLastError();
          n = -1;
        }
        max_clusters[S] = n;
      }
      if (max_clusters[S] >= tiles) {
        ep.splits = S;
        ep.kb_per_split = ceil_div(num_kb, S);
        break;
      }
    }
  }
  const int items = tiles * ep.splits;
  // split-K: one item per CTA, clusters of `splits` CTAs; else persistent over the tiles
  int grid = ep.splits > 1 ? items : (items < sms ? items : sms);
  if (MC) {  // CTA pairs: an even grid, one cluster per pair of adjacent M blocks
    const int pairs = (ceil_div(ep.M, C::BLOCK_M) / 2) * ceil_div(ep.N, BLOCK_N);
    grid = 2 * (pairs < sms / 2 ? pairs : sms / 2);
  }
  note_variant("gemm_tn kind=%d bn=%d stages=%d mc=%d splits=%d out=%s grid=%d", KIND, BLOCK_N, STAGES,
               MC ? 1 : 0, ep.splits, sizeof(OutT) == 2 && std::is_same<OutT, __half>::value ? "f16" : "bf16",
               grid);
  {
    ScopedKernelTimer timer(timer_name, st);
    B200_CUDA_OK(launch_kernel(kern, dim3(grid), dim3(C::THREADS), C::SMEM_BYTES, st, MC ? 2 : ep.splits, pdl_family(ep.M, kPdlGemmBig), ta, tb, ep));
  }
  B200_CHECK_LAUNCH("gemm_tn_kernel");
  return 0;
}

}  // namespace b200
