// FP8 W8A8 GEMM with per-token x per-channel scales on tcgen05 (replaces sgl-kernel
// fp8_scaled_mm, sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1071-1146, called from
// apply_fp8_linear, python/sglang/srt/layers/quantization/fp8_utils.py:636-648):
//   out[M,N] = (A_e4m3[M,K] . B_e4m3[N,K]^T) * sa[m] * sb[n] (+ bias[n])  -> bf16 / f16
// plus the TMA-descriptor helper shared by every tcgen05 kernel of the library.
#include "fp8_gemm_smallm.cuh"

#include <cstdlib>
#include <mutex>
#include <vector>

namespace b200 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                 uint64_t pitch_bytes, uint32_t box_rows, uint32_t box_cols, int elt_bytes,
                 bool is_weight) {
  EncodeTiledFn enc = get_encode_tiled();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15u) == 0 && pitch_bytes % 16 == 0,
               "TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch "
               "(pitch=%llu)", static_cast<unsigned long long>(pitch_bytes));
  B200_REQUIRE(box_cols * elt_bytes == 128 && box_rows <= 256, "bad TMA box %ux%u", box_rows,
               box_cols);
  const CUtensorMapDataType dt = elt_bytes == 1   ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                 : elt_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                                  : CU_TENSOR_MAP_DATA_TYPE_UINT32;
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         is_weight ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                   : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// Same, without swizzle: the box lands as dense rows of box_cols * elt_bytes bytes (a multiple of 16).
int make_tmap_2d_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                       uint64_t pitch_bytes, uint32_t box_rows, uint32_t box_cols, int elt_bytes) {
  EncodeTiledFn enc = get_encode_tiled();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15u) == 0 && pitch_bytes % 16 == 0,
               "TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch "
               "(pitch=%llu)", static_cast<unsigned long long>(pitch_bytes));
  B200_REQUIRE((box_cols * elt_bytes) % 16 == 0 && box_rows <= 256 && box_cols <= 256,
               "bad TMA box %ux%u", box_rows, box_cols);
  const CUtensorMapDataType dt = elt_bytes == 1   ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                 : elt_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                                  : CU_TENSOR_MAP_DATA_TYPE_UINT32;
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// ---- small-M dispatch knobs (env overrides are for sweeps/tests; defaults are the tuned values) --
static int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return v && *v ? std::atoi(v) : dflt;
}
static std::atomic<int> g_smallm_max_m{-1};
static std::atomic<int> g_smallm_splits{-1};
static std::atomic<bool> g_smallm_forced{false};
bool smallm_forced() { return g_smallm_forced.load(std::memory_order_relaxed); }

int pdl_late() {
  static int v = env_int("B200_PDL_LATE", 1);
  return v;
}

int smallm_force_mt() {
  static int v = env_int("B200_SMALLM_FORCE_MT", 0);
  return v;
}

int smallm_max_m() {
  int v = g_smallm_max_m.load(std::memory_order_relaxed);
  if (v < 0) {
    v = env_int("B200_SMALLM_MAX_M", 128);
    if (v > 128) v = 128;
    g_smallm_max_m.store(v, std::memory_order_relaxed);
  }
  return v;
}

// K-split (= cluster size) of the small-M kernel: the smallest power of two that yields at least
// one CTA per SM (measured on B200: 128..256 CTAs stream fastest, more only adds reduce work), at
// least 4 K blocks (64 KiB of weights) per CTA, at most 8 (portable cluster size).
int smallm_pick_splits(int n_tiles, int num_kb) {
  int forced = g_smallm_splits.load(std::memory_order_relaxed);
  if (forced < 0) {
    forced = env_int("B200_SMALLM_SPLITS", 0);
    g_smallm_splits.store(forced, std::memory_order_relaxed);
  }
  int s;
  if (forced > 0) {
    s = forced;
  } else {
    const int want = ceil_div(device_sm_count(), n_tiles);
    const int by_k = num_kb / 4 > 0 ? num_kb / 4 : 1;
    s = 1;
    while (s < want && s * 2 <= by_k && s < 8) s *= 2;
  }
  if (s > 8) s = 8;
  if (s > num_kb) s = num_kb;
  if (s < 1) s = 1;
  return s;
}

// Mid-M regime (one or two 128-row token blocks): the output has far fewer 128x128 tiles than the
// chip has SMs and a single CTA ingests only ~55 GB/s (measured: 48 CTAs -> 16 us for a 25 MB
// weight, whatever M), so K is split until roughly every SM streams (B200_GEMM_SPLITK=0 disables,
// =n forces n).  At least 4 K blocks per slice, at most 8 slices.
int splitk_pick(int tiles, int num_kb, int sms) {
  static int forced = env_int("B200_GEMM_SPLITK", -1);
  if (forced == 0) return 1;
  int s = forced > 0 ? forced : sms / (tiles > 0 ? tiles : 1);
  if (s > 8) s = 8;
  if (s > num_kb / 4) s = num_kb / 4;
  if (s < 1) s = 1;
  while (s > 1 && (s - 1) * ceil_div(num_kb, s) >= num_kb) --s;  // no empty slice
  return s;
}

// Row-parallel fusion: the small-M kernel can reduce across the TP group in its epilogue when the
// dense [M, N] partial fits one receive slot and the grid fits the flag table.
// Policy measured on B200: at TP = 2 (profiles/r01_bench_allreduce_tp2.json) fusing wins while the
// swap-AB kernel itself is the right GEMM (M <= 32: 7.7 vs 8.2 us at M = 1, 13.1 vs 15.4 us at
// M = 32, against GEMM + stand-alone LL all-reduce).  At TP = 8 (profiles/r01_bench_allreduce_tp8.json)
// every tile pushes to 7 peers and the fused epilogue only wins at the smallest batches (o_proj:
// 8.4 vs 10.8 us at M = 1 but 16.1 vs 12.7 at M = 8 and 51.6 vs 20.2 at M = 32), so the limit
// shrinks with the group: 32 / 8 / 4 tokens at world 2 / 4 / 8 (world 4 interpolated).
// b200_comm_fuse_config / B200_COMM_FUSE_MAX_M override.
static std::atomic<int> g_fuse_max_m{-1};
static int fuse_max_m(int world) {
  int v = g_fuse_max_m.load(std::memory_order_relaxed);
  if (v < 0) {
    static const int env = env_int("B200_COMM_FUSE_MAX_M", -1);
    v = env >= 0 ? env : (world <= 2 ? 32 : world <= 4 ? 8 : 4);
  }
  return v > 128 ? 128 : v;
}

bool comm_can_fuse(const void* comm, int M, int N) {
  if (!comm) return false;
  const Comm* c = static_cast<const Comm*>(comm);
  if (M > fuse_max_m(c->dev.world)) return false;
  // LL words: one 8-byte word per (token pair, channel), pairs padded per cluster rank (S <= 8)
  return static_cast<uint64_t>(M / 2 + 8) * N * 8 <= c->dev.slot_bytes &&
         ceil_div(N, 128) * 8 <= kCommMaxBlocks;
}

// Unfused tail for the big-tile kernels: in-place one-shot all-reduce of out[M, N] (ldc == N) or
// an error asking the caller to use NCCL when the message exceeds the communicator's slot.
int comm_all_reduce_rows(void* comm, void* out, int64_t ldc, int M, int N, int dtype,
                         void* stream) {
  B200_REQUIRE(ldc == N, "fused all-reduce needs a contiguous output (ldc %lld != N %d)",
               static_cast<long long>(ldc), N);
  return b200_comm_all_reduce(comm, out, out, static_cast<int64_t>(M) * N, dtype, stream);
}

}  // namespace b200

extern "C" {

/* Largest M for which the *_ar entry points fuse the reduce into the GEMM epilogue (< 0 restores
 * the default). */
int b200_comm_fuse_config(int max_m) {
  b200::g_fuse_max_m.store(max_m, std::memory_order_relaxed);
  return 0;
}

/* Test/sweep hook: max_m < 0 / splits < 0 re-read the environment defaults. */
int b200_fp8_gemm_smallm_config(int max_m, int splits) {
  b200::g_smallm_max_m.store(max_m > 128 ? 128 : max_m, std::memory_order_relaxed);
  b200::g_smallm_splits.store(splits, std::memory_order_relaxed);
  b200::g_smallm_forced.store(max_m > 0, std::memory_order_relaxed);
  return 0;
}

static int fp8_gemm_rowcol_impl(const void* a, int64_t lda, const void* b, int64_t ldb,
                               const float* sa, int sa_stride, const float* sb, int sb_stride,
                               const void* bias, void* out, int64_t ldc, int M, int N, int K,
                               int out_dtype, void* stream, void* comm) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(a && b && out, "b200_fp8_gemm_rowcol: null pointer");
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "b200_fp8_gemm_rowcol: bad shape %dx%dx%d", M, N, K);
  B200_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && aligned16(a) && aligned16(b),
               "b200_fp8_gemm_rowcol: A/B must be 16-byte aligned with lda/ldb %% 16 == 0 "
               "(same rule as the reference kernel, fp8_gemm_kernel.cu:1090-1097)");
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F16,
               "b200_fp8_gemm_rowcol: out dtype %d", out_dtype);
  GemmEpilogueParams ep;
  ep.out = out;
  ep.ldc = ldc;
  ep.sa = sa;
  ep.sb = sb;
  ep.sa_stride = sa_stride;
  ep.sb_stride = sb_stride;
  ep.bias = bias;
  ep.M = M;
  ep.N = N;
  ep.K = K;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Decode regime: weight streaming, swap-AB + cluster split-K.  Measured crossover against the
  // big-tile kernel with cluster split-K (profiles/r01_bench_gemm_midm_splitk.json): the swap-AB
  // kernel wins up to M = 32; from M = 33 the 128x128 tile split over K is faster on every
  // Llama-3.1-8B shape (10.8 vs 15.4-20.6 us at N = K = 4096, 17.2 vs 21.7-27 us at K = 14336).
  const bool fuse_ar = comm_can_fuse(comm, M, N);
  // Exception (profiles/r01_bench_gemm_midm_swap.json): a weight with more 128-row tiles than SMs
  // (gate_up 28672x4096) cannot be split and runs the big tiles in two uneven waves; there the
  // swap-AB kernel stays ahead up to M = 64 (27.4 vs 31.5 us at M = 48, 29.1 vs 31.7 at M = 64).
  const bool wide = ceil_div(N, 128) > device_sm_count();
  if (fuse_ar || (M <= smallm_max_m() && (M <= 32 || (wide && M <= 64) || smallm_forced()))) {
    SmallMParams sp = {};
    if (fuse_ar) {
      sp.use_comm = 1;
      sp.comm = static_cast<const Comm*>(comm)->dev;
    }
    sp.out = out;
    sp.ldc = ldc;
    sp.sa = sa;
    sp.sa_stride = sa_stride;
    sp.sb = sb;
    sp.sb_stride = sb_stride;
    sp.bias = bias;
    sp.M = M;
    sp.N = N;
    sp.K = K;
    const int rc = dispatch_smallm<false>(a, lda, b, ldb, sp, out_dtype, st, "fp8_gemm_smallm");
    if (rc || !comm || fuse_ar) return rc;
    return comm_all_reduce_rows(comm, out, ldc, M, N, out_dtype, stream);
  }
  // Tile choice: 128x256 when there is enough work to fill the SMs with it, else 128x128.
  const int sms = device_sm_count();
  const int64_t tiles256 = static_cast<int64_t>(ceil_div(M, 128)) * ceil_div(N, 256);
  bool big = tiles256 >= sms;
  // Measured (profiles/r01_bench_gemm_tilewidth.json): in the weight-streaming regime below that
  // the 256-wide tile does not pay even when it saves a wave (28672x4096, M <= 128: 34 vs 31.6 us;
  // M = 192-256: 44-46 vs 46-47 us), so the rule stays.  B200_GEMM_BN = 128 / 256 forces the choice.
  static const int force_bn = env_int("B200_GEMM_BN", 0);
  if (force_bn == 128) big = false;
  if (force_bn == 256) big = true;
  int rc;
  // pairs of M blocks share the weight tile by TMA multicast (B200_GEMM_MCAST=0 turns it off)
  static const int mcast_on = env_int("B200_GEMM_MCAST", 1);
  if (big && mcast_on && ceil_div(M, 128) % 2 == 0) {
    rc = out_dtype == B200_BF16
             ? launch_gemm_tn<kKindFp8, 256, 4, __nv_bfloat16, true>(a, lda, b, ldb, ep, st, "fp8_gemm")
             : launch_gemm_tn<kKindFp8, 256, 4, __half, true>(a, lda, b, ldb, ep, st, "fp8_gemm");
  } else
  if (out_dtype == B200_BF16) {
    rc = big ? launch_gemm_tn<kKindFp8, 256, 4, __nv_bfloat16>(a, lda, b, ldb, ep, st, "fp8_gemm")
             : launch_gemm_tn<kKindFp8, 128, 6, __nv_bfloat16>(a, lda, b, ldb, ep, st, "fp8_gemm", true);
  } else {
    rc = big ? launch_gemm_tn<kKindFp8, 256, 4, __half>(a, lda, b, ldb, ep, st, "fp8_gemm")
             : launch_gemm_tn<kKindFp8, 128, 6, __half>(a, lda, b, ldb, ep, st, "fp8_gemm", true);
  }
  if (rc || !comm) return rc;
  return comm_all_reduce_rows(comm, out, ldc, M, N, out_dtype, stream);
}

int b200_fp8_gemm_rowcol(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                         int sa_stride, const float* sb, int sb_stride, const void* bias,
                         void* out, int64_t ldc, int M, int N, int K, int out_dtype,
                         void* stream) {
  return fp8_gemm_rowcol_impl(a, lda, b, ldb, sa, sa_stride, sb, sb_stride, bias, out, ldc, M, N, K,
                              out_dtype, stream, nullptr);
}

int b200_fp8_gemm_rowcol_ar(void* comm, const void* a, int64_t lda, const void* b, int64_t ldb,
                            const float* sa, int sa_stride, const float* sb, int sb_stride,
                            const void* bias, void* out, int64_t ldc, int M, int N, int K,
                            int out_dtype, void* stream) {
  B200_REQUIRE(comm, "b200_fp8_gemm_rowcol_ar: null communicator");
  return fp8_gemm_rowcol_impl(a, lda, b, ldb, sa, sa_stride, sb, sb_stride, bias, out, ldc, M, N, K,
                              out_dtype, stream, comm);
}

}  // extern "C"
