// Small-M (decode) FP8 W8A8 GEMM for sm_100a: the weight-streaming, HBM-bound regime M <= 128.
//
// Replaces the same reference ops as fp8_gemm.cu / fp8_block_gemm.cu (sgl_kernel.fp8_scaled_mm,
// fp8_gemm_kernel.cu:1071-1146; fp8_blockwise_scaled_mm / w8a8_block_fp8_matmul,
// fp8_kernel.py:774-891) for the decode batch sizes of BASELINE cfg 3 (M = 1..128).
//
// Why a second kernel: with M <= 128 there is exactly one 128-row token tile, so the big-tile
// kernel launches only N/128 CTAs (32..48 for the attention projections) and most SMs never touch
// HBM.  Here the problem is transposed ("swap-AB") and split along K:
//   * UMMA M = 128 WEIGHT rows (A operand, streamed once from HBM by TMA, 16 KiB per K block),
//     UMMA N = MT tokens (B operand, MT in {16,32,64,128}; rows >= M are TMA zero fill, no traffic),
//     accumulator in TMEM: lane = output channel, column = token.
//   * a thread-block CLUSTER of S CTAs shares one 128-channel tile; CTA r streams K blocks
//     [r*kbs, (r+1)*kbs).  n_tiles * S CTAs with 2 CTAs resident per SM keep every SM's TMA queue
//     full.  Partials are reduced through distributed shared memory in a FIXED order (rank 0..S-1),
//     so the result is deterministic and no global workspace or atomics are needed.
//   * epilogue: TMEM -> registers; token m of the partial is PUSHED (st.shared::cluster) into the
//     idle pipeline ring of rank m % S, then each rank sums the S slots of its own tokens from local
//     smem in rank order, applies scales/bias and stores (S == 1: straight from registers).
//   * block-scaled variant: each 128-wide K block goes to a fresh TMEM buffer and is promoted into
//     fp32 registers with As[m,kb] * Bs[n/128,kb] (exact fp32 scale semantics of
//     native_w8a8_block_fp8_matmul, test_block_fp8.py:276-325).
#pragma once

#include "comm.cuh"
#include "gemm_tc.cuh"

namespace b200 {

struct SmallMParams {
  void* out;
  int64_t ldc;
  const float* sa;   // per-token scale (rowcol) or null
  int sa_stride;
  const float* sb;   // per-channel scale (rowcol) or null
  int sb_stride;
  const float* as;   // block: [M, K/128] at as[m*as_row + kb*as_col]
  int64_t as_row, as_col;
  const float* bs;   // block: [N/128, K/128] row stride bs_row
  int64_t bs_row;
  const void* bias;
  int M, N, K;
  int kb_per_split;
  int pdl_late;  // trigger dependents after the main loop instead of right after the dependency wait
  // Row-parallel fusion (RowParallelLinear.forward, linear.py:1277-1299): when use_comm != 0 the
  // epilogue does not store its tile but pushes it (bf16/f16, already scaled, bias on rank 0 only)
  // into every tensor-parallel peer's receive slot over NVLink, then sums the `world` partials of
  // the same tile in rank order (comm.cuh protocol) and stores the all-reduced result.
  int use_comm;
  CommDev comm;
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(threads) : "memory");
}

template <int MT, bool BLK, int KSUB = 1>
struct SmallMCfg {
  // A ring slot ("block") spans KS sub-blocks of 128 K: every barrier hand-off of the pipeline costs
  // a fixed ~0.3 us round trip per role (profiles/r01_w4a16_notes.md), so the decode instances
  // can move 32 KiB of weights per hand-off (KSUB = 2): measured +5-8 % on the wide weights
  // (gate_up M = 1: 20.7 -> 19.6 us = 6.0 TB/s, 92 % of the HBM peak; down 11.1 us) but -7 % on
  // qkv / o, whose short K slices then hold too few blocks, so the launcher picks it by weight size.
  // The block-scaled variant needs one TMEM buffer per 128 K and keeps KS = 1.
  // MT = 16 only: the MT = 32 instance with 32 KiB ring blocks sits at the two-CTAs-per-SM shared
  // memory limit and fell to one CTA per SM once the epilogue's scale staging was added (gate_up
  // M = 32: 21.5 -> 26.8 us, down 13.3 -> 25.7 us, profiles/r02_bench_gemm_midm.json)
  static constexpr int KS = (!BLK && MT <= 16) ? KSUB : 1;
  static constexpr int W_SUB = 128 * 128, X_SUB = MT * 128;
  static constexpr int W_BYTES = KS * W_SUB;
  static constexpr int X_BYTES = KS * X_SUB;
  // Two CTAs per SM (one for the 128-token block-scaled variant whose accumulators need all
  // registers).  The weight ring is as deep as shared memory allows — the kernel is bound by bytes
  // in flight (measured: 3/4/5 stages -> 3.2/4.4/5.7 TB/s) — while the activation ring, fed from
  // L2 and identical for every CTA, is only NX deep and lags the weight ring.
  // MT = 64: 5 weight + 3 activation slots = 104 KiB, the footprint of the MT <= 32 instances that
  // are known to run two CTAs per SM (ncu showed one CTA per SM for the 112 KiB ring)
  static constexpr int BUDGET = (BLK && MT == 128) ? 208 * 1024 : (MT == 64 ? 108 * 1024 : 112 * 1024);
  // MT = 64: three activation slots like the smaller instances.  With two, the x tile of block i is only requested when the
  // MMA of block i - 2 retires, so every other K block exposed an L2 round trip (measured: the same
  // 25 MB weight took 9.4 us at MT = 32 and 14-16 us at MT = 64, profiles/r02_bench_gemm_midm_mt64.json)
  static constexpr int NX = (MT <= 32 && KS == 1) ? 3 : (MT == 64 ? 3 : 2);
  static constexpr int NW = (BUDGET - NX * X_BYTES) / W_BYTES;
  static constexpr int NBUF = BLK ? (MT <= 64 ? 4 : 2) : 1;
  static constexpr int TCOLS = (NBUF * MT) < 32 ? 32 : NBUF * MT;
  static constexpr int PART_BYTES = (MT + 8) * 128 * 4;  // S * ceil(M/S) <= M + S - 1 token slots
  static constexpr int RING_BYTES = NW * W_BYTES + NX * X_BYTES;
  static constexpr int BAR_BYTES = (2 * NW + 2 * NBUF) * 8 + 16;
  static constexpr int SMEM_BYTES = RING_BYTES + BAR_BYTES;
  static_assert(PART_BYTES <= NW * W_BYTES, "partial tile aliases the weight ring");
  static_assert(NW > NX && NX >= 2, "ring depths");
};

// AR: the epilogue reduces across the tensor-parallel group (p.comm); a separate instantiation so
// the plain kernel's unrolled epilogue carries none of that code (it cost 12-26 % at M = 16-32).
template <int MT, bool BLK, typename OutT, bool AR, int KSUB>
__global__ void __launch_bounds__(192, (BLK && MT == 128) ? 1 : 2)
fp8_gemm_smallm_kernel(const __grid_constant__ CUtensorMap tmap_w,
                       const __grid_constant__ CUtensorMap tmap_x, const SmallMParams p) {
  using C = SmallMCfg<MT, BLK, KSUB>;
  constexpr int NW = C::NW, NX = C::NX, NBUF = C::NBUF;
  extern __shared__ __align__(1024) uint8_t smem_raw[];  // SWIZZLE_128B tiles need 1024-byte alignment
  const uint32_t smem_base = smem_u32(smem_raw);
  const uint32_t ring_x = smem_base + NW * C::W_BYTES;
  const uint32_t bars = smem_base + C::RING_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (NW + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * NW + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * NW + NBUF + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * NW + 2 * NBUF);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_base));
  float* part = reinterpret_cast<float*>(smem_raw);  // [S * ceil(M/S)][128], aliases the W ring
  // per-token activation scales, staged once by the epilogue warps: read inside the store loop
  // they cost one serialised L2 round trip PER TOKEN (a load may not be hoisted above the previous
  // token's output store) — 0.1-0.2 us per token, the whole M dependence of this kernel
  // (profiles/r02_ncu_fp8_smallm_m64_source.md)
  __shared__ float sa_s[MT];

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t S = cluster_nctarank();
  const int n_tile = blockIdx.x / S;
  const int num_kb = ceil_div(p.K, 128 * C::KS);
  const int kb0 = min(num_kb, static_cast<int>(rank) * p.kb_per_split);
  const int kb1 = min(num_kb, kb0 + p.kb_per_split);
  const int nkb = kb1 - kb0;
  const bool has_work = nkb > 0;

  // K block i (local index) uses weight slot i % NW and activation slot i % NX; both loads credit
  // full_bar(i % NW).  The MMA of block i frees both slots with one commit on empty_bar(i % NW).
  auto issue_w = [&](int i) {
    const int s = i % NW;
    tc::mbar_arrive_expect_tx(full_bar(s), C::W_BYTES + C::X_BYTES);
    const uint32_t sw = smem_base + s * C::W_BYTES;
#pragma unroll
    for (int u = 0; u < C::KS; ++u)
      tc::tma_load_2d_hint(sw + u * C::W_SUB, &tmap_w, full_bar(s), ((kb0 + i) * C::KS + u) * 128,
                           n_tile * 128, tc::kEvictFirst);
  };
  auto issue_x = [&](int i) {
#pragma unroll
    for (int u = 0; u < C::KS; ++u)
      tc::tma_load_2d_hint(ring_x + (i % NX) * C::X_BYTES + u * C::X_SUB, &tmap_x, full_bar(i % NW),
                           ((kb0 + i) * C::KS + u) * 128, 0, tc::kEvictLast);
  };
  auto wait_mma_done = [&](int i) { tc::mbar_wait(empty_bar(i % NW), (i / NW) & 1u); };

  if (warp == 0 && lane == 0) {
    if ((smem_base & 1023u) != 0) __trap();
    tc::prefetch_tmap(&tmap_w);
    tc::prefetch_tmap(&tmap_x);
    for (int s = 0; s < NW; ++s) {
      tc::mbar_init(full_bar(s), 1);
      tc::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < NBUF; ++a) {
      tc::mbar_init(tfull_bar(a), 1);
      tc::mbar_init(tempty_bar(a), 4);
    }
    tc::fence_barrier_init();
    // Fill the weight ring BEFORE the CTA-wide setup barrier AND before the programmatic-launch
    // dependency wait: weights are never written during a step, so their first NW x 16 KiB are in
    // flight while the previous kernel of the stream / graph is still draining and the rest of
    // this CTA is being set up.  Activations (written by the previous kernel) only after the wait.
    for (int i = 0; i < NW && i < nkb; ++i) issue_w(i);
    griddep_wait();
    if (!p.pdl_late) griddep_launch_dependents();
    for (int i = 0; i < NX && i < nkb; ++i) issue_x(i);
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_slot, C::TCOLS);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer: stream this rank's K slice of the weight tile ==========
    if (lane == 0) {
      for (int i = NW; i < nkb; ++i) {  // W runs NW blocks ahead of the MMA, X only NX blocks
        wait_mma_done(i - NW);
        issue_w(i);
        issue_x(i - NW + NX);
      }
      for (int j = NX + (nkb > NW ? nkb - NW : 0); j < nkb; ++j) {
        wait_mma_done(j - NX);
        issue_x(j);
      }
      if (p.pdl_late) griddep_launch_dependents();  // every load of this CTA has been issued
    }
  } else if (warp == 1) {
    // ===================== MMA issuer ==========================================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc(0, 0, 128, MT, 0, 0);
      for (int i = 0; i < nkb; ++i) {
        const int buf = BLK ? i % NBUF : 0;
        if (BLK) tc::mbar_wait(tempty_bar(buf), ((i / NBUF) & 1u) ^ 1u);
        tc::mbar_wait(full_bar(i % NW), (i / NW) & 1u);
        tc::tc_fence_after();
        const uint32_t sw = smem_base + (i % NW) * C::W_BYTES;
        const uint32_t sx = ring_x + (i % NX) * C::X_BYTES;
        const uint32_t tmem_d = tmem_base + buf * MT;
#pragma unroll
        for (int u = 0; u < C::KS; ++u) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adesc = tc::make_smem_desc_sw128(sw + u * C::W_SUB + k * 32, 1024, 16);
            const uint64_t bdesc = tc::make_smem_desc_sw128(sx + u * C::X_SUB + k * 32, 1024, 16);
            const uint32_t accum = BLK ? (k != 0) : ((i != 0) | (u != 0) | (k != 0));
            tc::mma_f8f6f4(tmem_d, adesc, bdesc, idesc, accum ? 1u : 0u);
          }
        }
        tc::mma_commit(empty_bar(i % NW));
        if (BLK) tc::mma_commit(tfull_bar(buf));
      }
      if (!BLK && has_work) tc::mma_commit(tfull_bar(0));
    }
  } else {
    // ===================== epilogue ===========================================================
    // S == 1: TMEM -> registers -> scales/bias -> global (a warp writes 32 consecutive channels).
    // S  > 1: fixed-order cluster reduction in two phases separated by cluster barriers:
    //   barrier 1 (after this CTA's last MMA): every peer's pipeline ring is idle and may be
    //              overwritten;
    //   push:      token m of my partial goes to rank (m % S), slot [my rank][m / S][channel] of
    //              ITS ring with st.shared::cluster (fire-and-forget, no round trip);
    //   barrier 2: all pushes have landed;
    //   reduce:    each rank sums the S slots of its own tokens from LOCAL smem in rank order
    //              (deterministic), applies scales/bias and stores.
    griddep_wait();  // scales / bias / output buffers may belong to the previous kernel
    const int quarter = warp & 3;
    const int ch = quarter * 32 + lane;  // output channel within the tile (TMEM lane)
    if (!BLK) {
      if (ch < MT) sa_s[ch] = (p.sa && ch < p.M) ? __ldg(p.sa + static_cast<int64_t>(ch) * p.sa_stride) : 1.f;
      named_bar_sync(2, 128);
    }
    const int n = n_tile * 128 + ch;
    const bool n_ok = n < p.N;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    OutT* out = static_cast<OutT*>(p.out);
    float sbn = 1.f, bn = 0.f;
    if (!BLK && p.sb && n_ok) sbn = __ldg(p.sb + static_cast<int64_t>(n) * p.sb_stride);
    if (p.bias && n_ok) bn = to_float<OutT>(static_cast<const OutT*>(p.bias)[n]);
    const int tpr = ceil_div(p.M, static_cast<int>(S));  // tokens per rank (slots per source rank)
    const uint32_t push_base = smem_base + (rank * tpr * 128 + ch) * 4;
    const uint32_t ar_epoch = AR ? comm_begin(p.comm) : 0u;
    // Fused tensor-parallel reduce (LL protocol, comm.cuh): the thread's outputs come in a fixed
    // sequence (tokens 0,1,2,.. if S == 1, else rank, rank+S, ..); two consecutive ones are packed
    // into one 32-bit payload and pushed as an 8-byte {payload, epoch} word to word index
    // (pair * S + cluster rank) * N + n of slot [my tp rank] on EVERY tp rank (mine included).
    uint32_t ll_stash = 0;
    int ll_cnt = 0;
    auto ll_push = [&](uint32_t payload) {
      const int64_t widx = (static_cast<int64_t>(ll_cnt >> 1) * S + rank) * p.N + n;
      for (int r = 0; r < p.comm.world; ++r)
        ll_store1(comm_ll_slot(p.comm, r, p.comm.rank, ar_epoch) + widx, payload, ar_epoch);
    };

    auto finish = [&](int m, float v) {  // scales + bias + cast + store of one output element
      float sc = 1.f;
      if (!BLK) sc = sa_s[m] * sbn;
      if (!n_ok) return;
      const OutT o = from_float<OutT>(fmaf(v, sc, bn));
      if constexpr (AR) {
        const uint32_t bits = *reinterpret_cast<const unsigned short*>(&o);
        if (ll_cnt & 1) ll_push(ll_stash | (bits << 16));
        else ll_stash = bits;
        ++ll_cnt;
      } else {
        out[static_cast<int64_t>(m) * p.ldc + n] = o;
      }
    };
    // running (rank, slot) of token m = c*16 + j while the unrolled loops walk m upwards
    uint32_t dst_rank = 0, dst_slot = 0;
    auto emit = [&](int m, float v) {
      if (S == 1) {
        finish(m, v);
      } else {
        const uint32_t a = map_to_rank(push_base + dst_slot * 512, dst_rank);
        asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(a), "f"(v) : "memory");
        if (++dst_rank == S) {
          dst_rank = 0;
          ++dst_slot;
        }
      }
    };

    if (BLK) {
      float acc[MT];
#pragma unroll
      for (int m = 0; m < MT; ++m) acc[m] = 0.f;
      const float* bs_row = p.bs + static_cast<int64_t>(n_tile) * p.bs_row;
      for (int it = 0; it < nkb; ++it) {
        const int kb = kb0 + it;
        const int buf = it % NBUF;
        const float sb = __ldg(bs_row + kb);
        const float* as_kb = p.as + static_cast<int64_t>(kb) * p.as_col;
        tc::mbar_wait(tfull_bar(buf), (it / NBUF) & 1u);
        tc::tc_fence_after();
#pragma unroll
        for (int c = 0; c < MT / 16; ++c) {
          if (c * 16 >= p.M) break;
          uint32_t r[16];
          tc::tmem_ld16(lane_addr + buf * MT + c * 16, r);
          float sc[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int m = c * 16 + j;
            sc[j] = __ldg(as_kb + static_cast<int64_t>(m < p.M ? m : p.M - 1) * p.as_row) * sb;
          }
          tc::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            acc[c * 16 + j] = fmaf(__uint_as_float(r[j]), sc[j], acc[c * 16 + j]);
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(tempty_bar(buf));
      }
      if (S > 1) {
        __syncwarp();
        cluster_sync_all();  // barrier 1
      }
#pragma unroll
      for (int m = 0; m < MT; ++m)
        if (m < p.M) emit(m, acc[m]);
    } else {
      if (has_work) {
        tc::mbar_wait(tfull_bar(0), 0);
        tc::tc_fence_after();
      }
      if (S > 1) {
        __syncwarp();
        cluster_sync_all();  // barrier 1
      }
#pragma unroll
      for (int c = 0; c < MT / 16; ++c) {
        if (c * 16 >= p.M) break;
        uint32_t r[16];
        if (has_work) {
          tc::tmem_ld16(lane_addr + c * 16, r);
          tc::tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = 0u;
        }
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (c * 16 + j < p.M) emit(c * 16 + j, __uint_as_float(r[j]));
      }
    }
    if (S > 1) {
      __syncwarp();
      cluster_sync_all();  // barrier 2: every peer's pushes into my ring are visible
      const float* mine = part + ch;
      int m = rank;
#pragma unroll 4
      for (int j = 0; j < tpr; ++j, m += S) {
        if (m >= p.M) break;
        float sum = 0.f;
        for (uint32_t q = 0; q < S; ++q) sum += mine[(q * tpr + j) * 128];
        finish(m, sum);
      }
    }
    if constexpr (AR) {
      if (ll_cnt & 1) ll_push(ll_stash);  // odd tail: upper half is padding
      const int pairs = (ll_cnt + 1) >> 1;
      for (int pi = 0; pi < pairs; ++pi) {
        const int64_t widx = (static_cast<int64_t>(pi) * S + rank) * p.N + n;
        float s0 = 0.f, s1 = 0.f;
        for (int r = 0; r < p.comm.world; ++r) {  // rank order: bit-identical on every tp rank
          const uint32_t w = ll_load1(comm_ll_slot(p.comm, p.comm.rank, r, ar_epoch) + widx, ar_epoch);
          const float2 f = unpack2<OutT>(w);
          s0 += f.x;
          s1 += f.y;
        }
        const int q0 = 2 * pi, q1 = 2 * pi + 1;
        const int m0 = S == 1 ? q0 : static_cast<int>(rank) + q0 * static_cast<int>(S);
        const int m1 = S == 1 ? q1 : static_cast<int>(rank) + q1 * static_cast<int>(S);
        out[static_cast<int64_t>(m0) * p.ldc + n] = from_float<OutT>(s0);
        if (q1 < ll_cnt) out[static_cast<int64_t>(m1) * p.ldc + n] = from_float<OutT>(s1);
      }
      named_bar_sync(1, 128);
      if (ch == 0) comm_end(p.comm, ar_epoch, gridDim.x);
    }
  }
  if (warp < 2 && S > 1) {  // producer / MMA warps take part in the two cluster barriers
    __syncwarp();
    cluster_sync_all();
    cluster_sync_all();
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, C::TCOLS);
}

// Host side --------------------------------------------------------------------------------------
int smallm_max_m();                       // dispatch threshold (B200_SMALLM_MAX_M, default 128)
int smallm_pick_splits(int n_tiles, int num_kb);
int smallm_force_mt();
int pdl_late();
bool smallm_forced();                      // threshold set explicitly through the test hook
bool comm_can_fuse(const void* comm, int M, int N);
int comm_all_reduce_rows(void* comm, void* out, int64_t ldc, int M, int N, int dtype, void* stream);

template <int MT, bool BLK, typename OutT, bool AR, int KSUB>
int launch_smallm_ks(const void* x, int64_t ldx, const void* w, int64_t ldw, SmallMParams p,
                     cudaStream_t st, const char* timer_name) {
  using C = SmallMCfg<MT, BLK, KSUB>;
  CUtensorMap tw, tx;
  int rc = make_tmap_2d(&tw, w, p.N, p.K, ldw, 128, 128, 1, true);
  if (rc) return rc;
  rc = make_tmap_2d(&tx, x, p.M, p.K, ldx, MT, 128, 1, false);
  if (rc) return rc;
  auto kern = fp8_gemm_smallm_kernel<MT, BLK, OutT, AR, KSUB>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  const int n_tiles = ceil_div(p.N, 128);
  // the split factor is chosen on 128-wide K blocks as before; slices are whole ring blocks
  int S = smallm_pick_splits(n_tiles, ceil_div(p.K, 128));
  const int num_kb = ceil_div(p.K, 128 * C::KS);
  if (S > num_kb) S = num_kb;
  p.kb_per_split = ceil_div(num_kb, S);
  p.pdl_late = pdl_late();
  static int ctas_per_sm = -1;  // reported so that sweeps / tests can see the residency
  if (ctas_per_sm < 0 &&
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kern, 192, C::SMEM_BYTES) != cudaSuccess) {
    (void)cudaGetLastError();
    ctas_per_sm = 0;
  }
  note_variant("fp8_smallm mt=%d blk=%d ar=%d ksub=%d splits=%d tiles=%d ctas_per_sm=%d", MT,
               BLK ? 1 : 0, AR ? 1 : 0, KSUB, S, n_tiles, ctas_per_sm);
  {
    ScopedKernelTimer timer(timer_name, st);
    B200_CUDA_OK(launch_kernel(kern, dim3(n_tiles * S), dim3(192), C::SMEM_BYTES, st, S, pdl_family(p.M, kPdlGemmSmall), tw, tx, p));
  }
  B200_CHECK_LAUNCH("fp8_gemm_smallm_kernel");
  return 0;
}

template <int MT, bool BLK, typename OutT, bool AR>
int launch_smallm(const void* x, int64_t ldx, const void* w, int64_t ldw, const SmallMParams& p,
                  cudaStream_t st, const char* timer_name) {
  if constexpr (!BLK && MT <= 16 && !AR) {
    // 256-K ring blocks for weights of 48 MB and more (gate_up, down_proj of the named models)
    if (static_cast<int64_t>(p.N) * p.K >= (48ll << 20))
      return launch_smallm_ks<MT, BLK, OutT, AR, 2>(x, ldx, w, ldw, p, st, timer_name);
  }
  return launch_smallm_ks<MT, BLK, OutT, AR, 1>(x, ldx, w, ldw, p, st, timer_name);
}

template <bool BLK>
int dispatch_smallm(const void* x, int64_t ldx, const void* w, int64_t ldw, const SmallMParams& p,
                    int out_dtype, cudaStream_t st, const char* timer_name) {
#define B200_SMALLM_CASE(MT_)                                                                    \
  if (p.use_comm)                                                                                \
    return out_dtype == B200_BF16                                                                \
               ? launch_smallm<MT_, BLK, __nv_bfloat16, true>(x, ldx, w, ldw, p, st, timer_name) \
               : launch_smallm<MT_, BLK, __half, true>(x, ldx, w, ldw, p, st, timer_name);       \
  return out_dtype == B200_BF16                                                                  \
             ? launch_smallm<MT_, BLK, __nv_bfloat16, false>(x, ldx, w, ldw, p, st, timer_name)  \
             : launch_smallm<MT_, BLK, __half, false>(x, ldx, w, ldw, p, st, timer_name)
  const int mt_floor = smallm_force_mt();  // experiments only (B200_SMALLM_FORCE_MT)
  if (p.M <= 16 && mt_floor <= 16) { B200_SMALLM_CASE(16); }
  if (p.M <= 32 && mt_floor <= 32) { B200_SMALLM_CASE(32); }
  if (p.M <= 64 && mt_floor <= 64) { B200_SMALLM_CASE(64); }
  B200_SMALLM_CASE(128);
#undef B200_SMALLM_CASE
}

}  // namespace b200
