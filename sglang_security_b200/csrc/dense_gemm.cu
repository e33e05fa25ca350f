// Dense 16-bit GEMM on tcgen05 for the layers the quantisation configs leave unquantised
// (UnquantizedLinearMethod.apply = F.linear, python/sglang/srt/layers/linear.py:170-176; the
// lm_head of the decode step, vocab_parallel_embedding.py / logits_processor.py:_get_logits):
//   out[M, N] = x[M, K] . w[N, K]^T (+ bias[N]),  bf16 / f16 operands, fp32 accumulation in TMEM.
// Same persistent TMA + tcgen05 kernel as the FP8 path (gemm_tc.cuh) with kind::f16; K-major
// operands, so the reference's [N, K] weight is used as stored.
#include "gemm_tc.cuh"

namespace b200 {

template <int KIND, typename OutT>
static int dense_gemm_launch(const void* x, int64_t ldx, const void* w, int64_t ldw,
                             const GemmEpilogueParams& ep, cudaStream_t st) {
  // 128 x 256 tiles once they fill the machine, else 128 x 128 (with cluster split-K when even
  // those leave SMs idle) — the rule of the FP8 dispatch (fp8_gemm.cu)
  const int64_t tiles256 = static_cast<int64_t>(ceil_div(ep.M, 128)) * ceil_div(ep.N, 256);
  if (tiles256 >= device_sm_count())
    return launch_gemm_tn<KIND, 256, 4, OutT>(x, ldx, w, ldw, ep, st, "dense_gemm");
  return launch_gemm_tn<KIND, 128, 6, OutT>(x, ldx, w, ldw, ep, st, "dense_gemm", true);
}

}  // namespace b200

extern "C" int b200_gemm_16bit(const void* x, int64_t ldx, const void* w, int64_t ldw,
                               const void* bias, void* out, int64_t ldc, int M, int N, int K,
                               int dtype, void* stream) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(x && w && out, "b200_gemm_16bit: null pointer");
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "b200_gemm_16bit: bad shape %dx%dx%d", M, N, K);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_gemm_16bit: dtype %d", dtype);
  B200_REQUIRE(aligned16(x) && aligned16(w) && ldx % 8 == 0 && ldw % 8 == 0,
               "b200_gemm_16bit: x / w must be 16-byte aligned with row strides %% 8 == 0 (TMA)");
  GemmEpilogueParams ep;
  ep.out = out;
  ep.ldc = ldc;
  ep.sa = nullptr;
  ep.sb = nullptr;
  ep.sa_stride = 0;
  ep.sb_stride = 0;
  ep.bias = bias;
  ep.M = M;
  ep.N = N;
  ep.K = K;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return dtype == B200_BF16 ? dense_gemm_launch<kKindBf16, __nv_bfloat16>(x, ldx, w, ldw, ep, st)
                            : dense_gemm_launch<kKindF16, __half>(x, ldx, w, ldw, ep, st);
}
