// Block-scaled FP8 GEMM (DeepSeek-style 128x128 weight blocks, 1x128 activation groups) on tcgen05.
// Replaces sgl-kernel fp8_blockwise_scaled_mm (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:
// 345-423) and the Triton _w8a8_block_fp8_matmul (python/sglang/srt/layers/quantization/
// fp8_kernel.py:435-517):
//   C[m,n] = sum_kb ( A[m, kb] . B[n, kb]^T ) * As[m, kb] * Bs[n/128, kb]      (fp32), then cast
//
// The fp32 scales differ per 128-wide K block, so they cannot ride inside one tensor-core
// accumulation.  Each K block is therefore multiplied into a FRESH TMEM accumulator (2 rotating
// 256-column buffers); the eight epilogue warps pull the finished partial tile out of TMEM and
// "promote" it into fp32 registers with one packed FMA per element pair:
//   acc[m, :] += (As[m,kb] * Bs[nb,kb]) * partial[m, :]
// (a warpgroup owns 128 columns = one weight scale block, so the combined scale is a per-thread scalar) while
// the MMA warp already works on the next K blocks.  Same TMA / mbarrier ring as gemm_tc.cuh.
#include "fp8_gemm_smallm.cuh"

namespace b200 {

struct BlockGemmParams {
  void* out;
  int64_t ldc;
  const float* as;      // [M, K/128] element (m, kb) at as[m * as_row + kb * as_col]
  int64_t as_row, as_col;
  const float* bs;      // [ceil(N/128), ceil(K/128)] row-major, row stride bs_row
  int64_t bs_row;
  const void* bias;
  int M, N, K;
};

// Tile 128 x 256: a 128 x 128 tile moves 64 KiB through shared memory (TMA in + UMMA out) per 256
// tensor-core clocks, twice what an SM's shared memory delivers, and plateaued at 1.59 PFLOP/s
// whatever the epilogue did (measured).  With N = 256 the tile spans TWO weight scale blocks: two
// epilogue warpgroups take one 128-column block each, so the combined scale stays a per-thread
// scalar.  One K block (512 tensor-core clocks) per TMEM hand-off, two 256-column accumulators.
// BN = 128 (twice the tiles, each warpgroup takes 64 columns of the one scale block) is kept for
// problems whose 128 x 256 tiles would leave SMs idle.
constexpr int kBlkThreads = 320;  // TMA, MMA, 2 x 4 epilogue warps
template <int BN>
struct BlkCfg {
  static constexpr int STAGES = BN == 256 ? 4 : 6;   // 48 / 32 KiB each: A 128 x 128, B BN x 128
  static constexpr int NBUF = 512 / BN;              // TMEM accumulator buffers of BN columns
  static constexpr int SMEM = STAGES * (128 + BN) * 128 + (2 * STAGES + 2 * NBUF) * 8 + 16 + 1024;
};

template <typename OutT, int BN>
__global__ void __launch_bounds__(kBlkThreads, 1)
fp8_block_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a,
                      const __grid_constant__ CUtensorMap tmap_b, const BlockGemmParams p) {
  constexpr int BLOCK = 128, STAGES = BlkCfg<BN>::STAGES, NBUF = BlkCfg<BN>::NBUF;
  constexpr int COLS = BN / 2;  // columns per epilogue warpgroup
  constexpr int A_BYTES = BLOCK * 128, B_BYTES = BN * 128, STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES * STAGE_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + NBUF + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 2 * NBUF);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int m_blocks = ceil_div(p.M, BLOCK), n_blocks = ceil_div(p.N, BN);
  const int num_tiles = m_blocks * n_blocks;
  const int num_kb = ceil_div(p.K, BLOCK);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmap_a);
    tc::prefetch_tmap(&tmap_b);
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(full_bar(s), 1);
      tc::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < NBUF; ++a) {
      tc::mbar_init(tfull_bar(a), 1);
      tc::mbar_init(tempty_bar(a), 8);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_slot, NBUF * BN);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  griddep_wait();  // activations / scales come from the previous kernel
  griddep_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(empty_bar(stage), phase ^ 1u);
          tc::mbar_arrive_expect_tx(full_bar(stage), STAGE_BYTES);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          tc::tma_load_2d(sa, &tmap_a, full_bar(stage), kb * BLOCK, m_blk * BLOCK);
          tc::tma_load_2d(sa + A_BYTES, &tmap_b, full_bar(stage), kb * BLOCK, n_blk * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc(0, 0, BLOCK, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;  // running K-block counter across tiles -> TMEM buffer ring
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int buf = it % NBUF;
          const uint32_t buf_phase = (it / NBUF) & 1u;
          tc::mbar_wait(tempty_bar(buf), buf_phase ^ 1u);
          tc::mbar_wait(full_bar(stage), phase);
          tc::tc_fence_after();
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_BYTES;
          const uint32_t tmem_d = tmem_base + buf * BN;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adesc = tc::make_smem_desc_sw128(sa + k * 32, 1024, 16);
            const uint64_t bdesc = tc::make_smem_desc_sw128(sb + k * 32, 1024, 16);
            tc::mma_f8f6f4(tmem_d, adesc, bdesc, idesc, k != 0 ? 1u : 0u);
          }
          tc::mma_commit(empty_bar(stage));
          tc::mma_commit(tfull_bar(buf));
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else {
    const int quarter = warp & 3;          // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;      // which half of the tile's columns
    OutT* out = static_cast<OutT*>(p.out);
    const OutT* bias = static_cast<const OutT*>(p.bias);
    const int nsb = ceil_div(p.N, BLOCK);  // weight scale blocks along N
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
      const int row = m_blk * BLOCK + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      const int col_base = n_blk * BN + half * COLS;
      const int sblk = col_base / BLOCK;
      const bool blk_ok = sblk < nsb;      // the tile's second half may lie past N
      const float* as_row = p.as + static_cast<int64_t>(row_ok ? row : 0) * p.as_row;
      const float* bs_row = p.bs + static_cast<int64_t>(blk_ok ? sblk : 0) * p.bs_row;
      float2 acc[COLS / 2];
#pragma unroll
      for (int j = 0; j < COLS / 2; ++j) acc[j] = make_float2(0.f, 0.f);
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int buf = it % NBUF;
        const uint32_t buf_phase = (it / NBUF) & 1u;
        const float s = __ldg(as_row + kb * p.as_col) * __ldg(bs_row + kb);
        const float2 s2 = make_float2(s, s);
        tc::mbar_wait(tfull_bar(buf), buf_phase);
        tc::tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + buf * BN +
                               half * COLS;
#pragma unroll
        for (int c = 0; c < COLS / 32; ++c) {
          uint32_t r[32];
          tc::tmem_ld32(taddr + c * 32, r);
          tc::tmem_ld_wait();
          if (c == COLS / 32 - 1) {  // the whole partial is in registers: hand the buffer back before the math
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tempty_bar(buf));
          }
#pragma unroll
          for (int j = 0; j < 16; ++j)
            acc[c * 16 + j] = ffma2(
                make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])), s2,
                acc[c * 16 + j]);
        }
      }
      if (row_ok && blk_ok) {
        OutT* orow = out + static_cast<int64_t>(row) * p.ldc;
        const int col0 = col_base;
        if (col0 + COLS <= p.N && (p.ldc & 7) == 0) {
#pragma unroll
          for (int v = 0; v < COLS / 8; ++v) {
            uint32_t packed[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float2 x = acc[v * 4 + e];
              if (bias) {
                x.x += to_float<OutT>(bias[col0 + v * 8 + e * 2]);
                x.y += to_float<OutT>(bias[col0 + v * 8 + e * 2 + 1]);
              }
              packed[e] = pack_out2<OutT>(x.x, x.y);
            }
            *reinterpret_cast<uint4*>(orow + col0 + v * 8) =
                make_uint4(packed[0], packed[1], packed[2], packed[3]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < COLS / 2; ++j) {
            const int c = col0 + 2 * j;
            if (c < p.N) orow[c] = from_float<OutT>(acc[j].x + (bias ? to_float<OutT>(bias[c]) : 0.f));
            if (c + 1 < p.N)
              orow[c + 1] = from_float<OutT>(acc[j].y + (bias ? to_float<OutT>(bias[c + 1]) : 0.f));
          }
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, NBUF * BN);
}

template <typename OutT, int BN>
static int launch_block_gemm_bn(const void* a, int64_t lda, const void* b, int64_t ldb,
                                const BlockGemmParams& p, cudaStream_t st) {
  constexpr int SMEM = BlkCfg<BN>::SMEM;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d(&ta, a, p.M, p.K, lda, 128, 128, 1, false);
  if (rc) return rc;
  rc = make_tmap_2d(&tb, b, p.N, p.K, ldb, BN, 128, 1, true);
  if (rc) return rc;
  auto kern = fp8_block_gemm_kernel<OutT, BN>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_done = true;
  }
  const int tiles = ceil_div(p.M, 128) * ceil_div(p.N, BN);
  const int sms = device_sm_count();
  note_variant("fp8_block_gemm bn=%d tiles=%d", BN, tiles);
  {
    ScopedKernelTimer timer("fp8_block_gemm", st);
    B200_CUDA_OK(launch_kernel(kern, dim3(tiles < sms ? tiles : sms), dim3(kBlkThreads), SMEM, st, 1, pdl_family(p.M, kPdlGemmBig), ta, tb, p));
  }
  B200_CHECK_LAUNCH("fp8_block_gemm_kernel");
  return 0;
}

template <typename OutT>
static int launch_block_gemm(const void* a, int64_t lda, const void* b, int64_t ldb,
                             const BlockGemmParams& p, cudaStream_t st) {
  // 128 x 256 tiles once they give every SM at least ~1.5 tiles, else 128 x 128
  const int tiles256 = ceil_div(p.M, 128) * ceil_div(p.N, 256);
  if (2 * tiles256 >= 3 * device_sm_count()) return launch_block_gemm_bn<OutT, 256>(a, lda, b, ldb, p, st);
  return launch_block_gemm_bn<OutT, 128>(a, lda, b, ldb, p, st);
}

}  // namespace b200

extern "C" {

static int fp8_gemm_block128_impl(const void* a, int64_t lda, const void* b, int64_t ldb,
                                  const float* as, int64_t as_row_stride, int64_t as_col_stride,
                                  const float* bs, int64_t bs_row_stride, const void* bias,
                                  void* out, int64_t ldc, int M, int N, int K, int out_dtype,
                                  void* stream, void* comm) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(a && b && as && bs && out, "b200_fp8_gemm_block128: null pointer");
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "b200_fp8_gemm_block128: bad shape %dx%dx%d", M, N, K);
  B200_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && aligned16(a) && aligned16(b),
               "b200_fp8_gemm_block128: A/B must be 16-byte aligned with lda/ldb %% 16 == 0");
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F16,
               "b200_fp8_gemm_block128: out dtype %d", out_dtype);
  cudaStream_t st0 = static_cast<cudaStream_t>(stream);
  // (block-scaled: up to 64 tokens; the 128-token promotion loop is slower than the big tile)
  const bool fuse_ar = M <= 64 && comm_can_fuse(comm, M, N);
  if (fuse_ar || (M <= smallm_max_m() && (M <= 64 || smallm_forced()))) {  // decode regime: weight streaming, swap-AB + cluster split-K
    SmallMParams sp = {};
    if (fuse_ar) {
      sp.use_comm = 1;
      sp.comm = static_cast<const Comm*>(comm)->dev;
    }
    sp.out = out;
    sp.ldc = ldc;
    sp.as = as;
    sp.as_row = as_row_stride;
    sp.as_col = as_col_stride;
    sp.bs = bs;
    sp.bs_row = bs_row_stride;
    sp.bias = bias;
    sp.M = M;
    sp.N = N;
    sp.K = K;
    const int rc = dispatch_smallm<true>(a, lda, b, ldb, sp, out_dtype, st0, "fp8_block_gemm_smallm");
    if (rc || !comm || fuse_ar) return rc;
    return comm_all_reduce_rows(comm, out, ldc, M, N, out_dtype, stream);
  }
  BlockGemmParams p;
  p.out = out;
  p.ldc = ldc;
  p.as = as;
  p.as_row = as_row_stride;
  p.as_col = as_col_stride;
  p.bs = bs;
  p.bs_row = bs_row_stride;
  p.bias = bias;
  p.M = M;
  p.N = N;
  p.K = K;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rc = out_dtype == B200_BF16 ? launch_block_gemm<__nv_bfloat16>(a, lda, b, ldb, p, st)
                                        : launch_block_gemm<__half>(a, lda, b, ldb, p, st);
  if (rc || !comm) return rc;
  return comm_all_reduce_rows(comm, out, ldc, M, N, out_dtype, stream);
}

int b200_fp8_gemm_block128(const void* a, int64_t lda, const void* b, int64_t ldb, const float* as,
                           int64_t as_row_stride, int64_t as_col_stride, const float* bs,
                           int64_t bs_row_stride, const void* bias, void* out, int64_t ldc, int M,
                           int N, int K, int out_dtype, void* stream) {
  return fp8_gemm_block128_impl(a, lda, b, ldb, as, as_row_stride, as_col_stride, bs, bs_row_stride,
                                bias, out, ldc, M, N, K, out_dtype, stream, nullptr);
}

int b200_fp8_gemm_block128_ar(void* comm, const void* a, int64_t lda, const void* b, int64_t ldb,
                              const float* as, int64_t as_row_stride, int64_t as_col_stride,
                              const float* bs, int64_t bs_row_stride, const void* bias, void* out,
                              int64_t ldc, int M, int N, int K, int out_dtype, void* stream) {
  B200_REQUIRE(comm, "b200_fp8_gemm_block128_ar: null communicator");
  return fp8_gemm_block128_impl(a, lda, b, ldb, as, as_row_stride, as_col_stride, bs, bs_row_stride,
                                bias, out, ldc, M, N, K, out_dtype, stream, comm);
}

}  // extern "C"
