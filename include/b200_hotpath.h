/*
 * b200_hotpath.h — C-ABI of the B200-native (sm_100a) serving hot path.
 *
 * Every entry point takes raw DEVICE pointers, sizes/strides and a CUDA stream
 * (cudaStream_t passed as void*), returns an int status (0 = ok, non-zero =
 * error; text via b200_last_error()) and launches asynchronously on the given
 * stream.  No torch types cross this boundary.  The Python host side
 * (sglang_security_b200/*.py) binds it with ctypes and raises on non-zero
 * status, mirroring the reference convention that kernels raise
 * (TORCH_CHECK -> RuntimeError, e.g. sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1078-1108).
 *
 * Each function cites the reference interface (file:line under the reference
 * tree) it replaces.  Strides are in ELEMENTS of the tensor's dtype unless the
 * name says bytes.
 *
 * Process model: ONE GPU PER PROCESS (the reference's own model: one scheduler / model-runner
 * process per TP rank, managers/scheduler.py run_scheduler_process).  Kernel attributes, the SM
 * count and the communicator state are cached once per process for the first device used; driving
 * a second device from the same process is not supported.  b200_attn_decode_noalloc alone keeps
 * its state per device.
 */
#ifndef B200_HOTPATH_H_
#define B200_HOTPATH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes used by the *_dtype arguments */
enum b200_dtype {
  B200_BF16 = 0,
  B200_F16 = 1,
  B200_F32 = 2,
  B200_FP8_E4M3 = 3
};
/* Flag OR'ed into the `dtype` argument of the FP8-KV-pool entry points (b200_kv_write_fp8,
 * b200_rope_kv_write_fp8, b200_attn_decode_fp8kv[_fused], b200_kv_dequant_gather): the pool holds
 * float8_e5m2 bytes (`--kv-cache-dtype fp8_e5m2`, model_runner.py:845-846) instead of e4m3.
 * Writers then follow torch's `.to(torch.float8_e5m2)`: round-to-nearest-even, |x| >= 61440 -> Inf
 * byte (sign | 0x7c), NaN -> sign | 0x7f.  The e5m2 decode kernel has no in-kernel tail: a ticket
 * buffer passed to the _fused entry point is ignored (one merge / merge+quantise launch).        */
#define B200_KV_FP8_E5M2 0x100

/* ---- library ---------------------------------------------------------- */
int b200_version(void);
/* Thread-local text of the last error raised by a b200_* call in this thread. */
const char* b200_last_error(void);
/* Test hook: thread-local description of the kernel INSTANCE the last GEMM-family call of this
 * thread dispatched to (e.g. "gemm_tn kind=0 bn=256 stages=4 mc=1 splits=1 out=bf16 grid=148",
 * "fp8_block_gemm bn=256 tiles=768", "w4a16_gemm bm=256 swap=0 splits=1 tiles=...") so that parity
 * tests can assert that a shape reached the instance it is meant to cover. */
const char* b200_last_variant(void);
/* Number of kernels this library has launched so far in this process
 * (bench.py's gpu_launches counter). */
int64_t b200_launch_count(void);

/* Programmatic dependent launch (default on; env B200_PDL=0 turns it off): every kernel of the
 * library is launched with cudaLaunchAttributeProgrammaticStreamSerialization and orders its own
 * dependent accesses with griddepcontrol.wait, so consecutive kernels of a decode step overlap
 * launch latency, prologue and weight prefetch with the tail of their predecessor (also inside a
 * captured CUDA graph).  Returns the previous setting. */
int b200_set_pdl(int on);

/* Per-kernel device timing for bench.py's roofline leg: while enabled, the dominant kernels are
 * bracketed by CUDA events on the launching stream (not during graph capture).
 * b200_timing_read sums and clears the records of kernel `name` (NULL/"" = all). Names:
 * "attn_decode_split", "attn_merge", "kv_write", "attn_extend", "fp8_gemm", "w4a16_gemm", ... */
int b200_timing_enable(int on);
int b200_timing_read(const char* name, int64_t* launches, double* total_ms);

/* ---- KV pool write ---------------------------------------------------- */
/* Replaces MHATokenToKVPool.set_kv_buffer (python/sglang/srt/mem_cache/memory_pool.py:378-412):
 *   k_cache[loc[t], :, :] = k[t, :, :]; v_cache[loc[t], :, :] = v[t, :, :]   (bit-exact byte copy)
 * k, v          [num_tokens, Hkv, D] rows of row_bytes contiguous bytes, token stride given in bytes
 *               (k/v are strided slices of the fused qkv output, models/llama.py:186-189)
 * k_cache/v_cache  [size+page, Hkv, D] contiguous rows of row_bytes
 * loc           int64 [num_tokens] slot ids (out_cache_loc)                       */
int b200_kv_write(const void* k, const void* v, int64_t k_token_stride_bytes,
                  int64_t v_token_stride_bytes, void* k_cache, void* v_cache,
                  const int64_t* loc, int num_tokens, int k_row_bytes, int v_row_bytes,
                  void* stream);

/* Quantising variant for an FP8 (e4m3) pool, `--kv-cache-dtype fp8_e4m3` (model_runner.py:845-858):
 * replaces the cast branch of set_kv_buffer (memory_pool.py:390-400) in one kernel, bit-exact with
 * torch: has_scale != 0 -> x = round_to_src_dtype(x / scale) (the in-place div_), then
 * `.to(float8_e4m3fn)` (round-to-nearest-even, NO saturation: |x| > 464 and NaN -> NaN byte).
 * k, v  [num_tokens, row_elems] bf16/f16, token strides in ELEMENTS; caches [slots, row_elems] bytes.
 * Unlike the reference the source k/v tensors are left untouched.                               */
int b200_kv_write_fp8(const void* k, const void* v, int64_t k_token_stride,
                      int64_t v_token_stride, void* k_cache, void* v_cache, const int64_t* loc,
                      int num_tokens, int row_elems, float k_scale, float v_scale, int has_scale,
                      int dtype, void* stream);

/* ---- request -> token index gather ------------------------------------ */
/* Replaces create_flashinfer_kv_indices_triton (python/sglang/srt/layers/attention/utils.py:5-41):
 *   kv_indices[kv_indptr[b] + i] = req_to_token[req_pool_indices[b], kv_start[b] + i], i < lens[b]
 * req_pool_indices / lens are int64 (ForwardBatch fields) ; kv_start may be NULL.   */
int b200_kv_indices(const int32_t* req_to_token, int64_t req_to_token_stride,
                    const int64_t* req_pool_indices, const int64_t* lens,
                    const int32_t* kv_indptr, const int32_t* kv_start, int32_t* kv_indices,
                    int bs, void* stream);

/* ---- paged decode attention ------------------------------------------- */
/* Replaces AttentionBackend.forward_decode's kernel (torch_native_backend.py:112-180 as the
 * numerical oracle; triton_ops/decode_attention.py:682-733 as the GPU implementation):
 * for every request b and query head h
 *   o[b,h,:] = softmax_t( scale * q[b,h,:] . K[slot(b,t), h/g, :] ) @ V[slot(b,t), h/g, :],
 *   slot(b,t) = req_to_token[req_pool_indices[b], t],  t < seq_lens[b]
 * Split-KV: num_splits partial results go to `workspace`, merged by a second kernel.
 * q   [bs, Hq, D]   token stride q_stride elements; head stride D
 * o   [bs, Hq, Dv]  token stride o_stride elements
 * k_cache/v_cache [slots, Hkv, D] row stride kv_row_stride elements (normally Hkv*D)
 * dtype: B200_BF16 or B200_F16 for q/k/v/o.  logit_cap <= 0 disables tanh capping.
 * workspace must hold b200_attn_decode_workspace_bytes(bs, Hq, Dv, num_splits) bytes. */
size_t b200_attn_decode_workspace_bytes(int bs, int Hq, int Dv, int num_splits);
/* Split-count heuristic (replaces get_num_kv_splits_triton, triton_backend.py:23-72, with a
 * host-side choice that needs no device sync): avg_len = expected tokens per request.       */
int b200_attn_decode_num_splits(int bs, int Hkv, int group, int D, int64_t avg_len, int num_sms);
int b200_attn_decode(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                     int64_t kv_row_stride, void* o, int64_t o_stride,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const int64_t* req_pool_indices, const int64_t* seq_lens, int bs, int Hq,
                     int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                     void* workspace, size_t workspace_bytes, int num_splits, void* stream);

/* b200_attn_decode for callers that cannot allocate device memory (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same operands, math and kernels as b200_attn_decode, minus the workspace:
 * the split-KV partials go to a scratch array built into the library (one per GPU that loads it,
 * held until exit).  The call allocates nothing (no cudaMalloc / cudaFree / allocating library call).
 * Limit: b200_attn_decode_workspace_bytes(bs, Hq, Dv, num_splits) must not exceed
 * B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES (8 MiB: e.g. 64 requests x 32 heads x 7 splits at Dv 128);
 * larger requests are rejected with an error (status 1, b200_last_error()), nothing is launched.
 * The device that owns `stream` must be current.  Every call on a device shares the scratch and
 * finds the previous call's contents in it (none of it has to be zero); calls that overlap on the
 * GPU (different streams, concurrent graph launches) race on it - serialise them on one stream.  */
#define B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES (8u << 20)
int b200_attn_decode_noalloc(const void* q, int64_t q_stride, const void* k_cache,
                             const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                             const int32_t* req_to_token, int64_t req_to_token_stride,
                             const int64_t* req_pool_indices, const int64_t* seq_lens, int bs,
                             int Hq, int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                             int num_splits, void* stream);

/* Same contract over an FP8 (e4m3) KV pool (k_cache/v_cache rows of D bytes, kv_row_stride in
 * bytes): o = v_scale * softmax(scale * k_scale * q.K8^T) @ V8 — FlashInfer-backend semantics
 * (flashinfer_backend.py:499-509); pass 1.0 for the Triton backend's scale-free reading
 * (triton_backend.py:808-833).  q/o dtype bf16 or f16.  Splits are aligned to
 * b200_attn_decode_fp8kv_split_align() tokens.  Half the HBM bytes of the 16-bit pool.          */
int b200_attn_decode_fp8kv_split_align(void);
int b200_attn_decode_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                           const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens, int bs,
                           int Hq, int Hkv, int D, int Dv, float scale, float logit_cap,
                           float k_scale, float v_scale, int dtype, void* workspace,
                           size_t workspace_bytes, int num_splits, void* stream);

/* Fused decode tail: same operands and math as b200_attn_decode / b200_attn_decode_fp8kv, but the
 * split partials are merged inside the kernel — the last split CTA of every (request, kv-head tile)
 * to finish merges its own heads (ticket counters, no b200_attn_merge launch) — and, when
 * q_out != NULL, the last tile of a request also quantises the output row per token to e4m3 exactly
 * as b200_fp8_quant_per_token would from the rounded 16-bit row (q_out [bs, q_out_stride] bytes,
 * q_scales [bs] fp32): the o_proj input quantiser of Fp8LinearMethod.apply (fp8_utils.py:454-648)
 * disappears from the decode step.  `tickets`: b200_attn_decode_fused_ticket_bytes(bs, Hq, Hkv) bytes
 * of device memory, zero before the first call; the counters re-arm themselves, so CUDA-graph
 * replays need no memset.  Requires num_splits > 1 or q_out != NULL.
 * tickets == NULL selects the two-launch form of the same result: the split kernel followed by ONE
 * merge(+quantise) kernel (one CTA per request) instead of b200_attn_merge + the quantiser —
 * the better choice once the split kernel is bandwidth-bound (large batch x context).           */
int b200_attn_decode_fused(const void* q, int64_t q_stride, const void* k_cache,
                           const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens, int bs, int Hq,
                           int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                           void* workspace, size_t workspace_bytes, int num_splits, void* tickets,
                           size_t ticket_bytes, void* q_out, int64_t q_out_stride, float* q_scales,
                           void* stream);
int b200_attn_decode_fp8kv_fused(const void* q, int64_t q_stride, const void* k_cache,
                                 const void* v_cache, int64_t kv_row_stride, void* o,
                                 int64_t o_stride, const int32_t* req_to_token,
                                 int64_t req_to_token_stride, const int64_t* req_pool_indices,
                                 const int64_t* seq_lens, int bs, int Hq, int Hkv, int D, int Dv,
                                 float scale, float logit_cap, float k_scale, float v_scale,
                                 int dtype, void* workspace, size_t workspace_bytes, int num_splits,
                                 void* tickets, size_t ticket_bytes, void* q_out,
                                 int64_t q_out_stride, float* q_scales, void* stream);
size_t b200_attn_decode_fused_ticket_bytes(int bs, int Hq, int Hkv);

/* Merge of split-KV partial states (replaces _fwd_kernel_stage2, decode_attention.py:497-557,
 * and merge_state_v2, sgl-kernel/csrc/attention/merge_attn_states.cu): exposed for tests.
 *   part_o   fp32 [bs, Hq, num_splits, Dv] ; part_lse fp32 [bs, Hq, num_splits] (log2 domain)  */
/* Request b owns splits of split_chunk = round_up(ceil(len_b / num_splits), split_align) tokens;
 * only the first ceil(len_b / split_chunk) partials are read.                                  */
int b200_attn_decode_split_align(int D, int group);
int b200_attn_merge(const float* part_o, const float* part_lse, void* o, int64_t o_stride,
                    const int64_t* seq_lens, int bs, int Hq, int Dv, int num_splits,
                    int split_align, int dtype, void* stream);

/* ---- extend (prefill with cached radix prefix) attention, tcgen05 -------------- */
/* Replaces AttentionBackend.forward_extend's kernel (oracle torch_native_backend.py:27-110; GPU
 * reference extend_attention_fwd, triton_ops/extend_attention.py:306-438).  For request b the
 * extend_seq_lens[b] new tokens (rows extend_start_loc[b].. of q / o) attend to keys
 *   slot(b,t) = req_to_token[req_pool_indices[b], t],  t <= extend_prefix_lens[b] + i   (causal)
 * all gathered from the pool (the new tokens' K/V must already be written, as torch_native does).
 * q [T, Hq, D] token stride q_stride; o [T, Hq, Dv] token stride o_stride; D == Dv in {64, 128}.
 * max_extend_len = max_b extend_seq_lens[b] (host value, sizes the grid).                       */
int b200_attn_extend(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                     int64_t kv_row_stride, void* o, int64_t o_stride,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const int64_t* req_pool_indices, const int64_t* seq_lens,
                     const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens,
                     const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv,
                     int D, int Dv, float scale, float logit_cap, int causal, int dtype,
                     void* stream);

/* Same kernel with an output multiplier: extend over an FP8 KV pool runs on an exact 16-bit copy
 * of the requests' rows (b200_kv_dequant_gather) with scale * k_scale and out_scale = v_scale.    */
int b200_attn_extend_scaled(const void* q, int64_t q_stride, const void* k_cache,
                            const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                            const int32_t* req_to_token, int64_t req_to_token_stride,
                            const int64_t* req_pool_indices, const int64_t* seq_lens,
                            const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens,
                            const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq,
                            int Hkv, int D, int Dv, float scale, float logit_cap, float out_scale,
                            int causal, int dtype, void* stream);
/* FP8 (e4m3) pool -> dense 16-bit rows for the extend kernel (prefill is compute-bound; the extra
 * pass costs 3 bytes/element once per layer): for request b and t < seq_lens[b]
 *   k_out[kv_indptr[b] + t, :] = exact(k_cache[slot(b,t), :]),  dense_rtt[b, t] = kv_indptr[b] + t
 * k_cache/v_cache [slots, row_elems] bytes; k_out/v_out [sum seq_lens, row_elems] bf16/f16;
 * kv_indptr int32 [bs+1] exclusive prefix sum of seq_lens; max_len >= max_b seq_lens[b].          */
int b200_kv_dequant_gather(const void* k_cache, const void* v_cache, int row_elems,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens,
                           const int32_t* kv_indptr, void* k_out, void* v_out, int32_t* dense_rtt,
                           int64_t dense_rtt_stride, int bs, int64_t max_len, int dtype,
                           void* stream);

/* ---- FP8 activation quantisation ----------------------------------------- */
/* Per-token dynamic (static_scale == NULL): scales[m] = absmax(x[m,:]) / 448, q = sat(x * (1/scale))
 *   replaces sgl_per_token_quant_fp8 (sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:9-106).
 * use_division != 0: scales[m] = max(absmax, eps) / 448, q = sat(x / scale) — the group_size == K
 *   case of per_token_group_quant_fp8 used for load-time per-channel weight quant (fp8.py:340-343).
 * static_scale != NULL: q = sat(x * (1 / *static_scale)), scales untouched
 *   (replaces _static_quant_fp8, python/sglang/srt/layers/quantization/fp8_kernel.py:336-376).
 * x [M,K] bf16/f16 row stride ldx; q [M,K] e4m3 row stride ldq; K % 8 == 0.                    */
int b200_fp8_quant_per_token(const void* x, int64_t ldx, void* q, int64_t ldq, float* scales,
                             const float* static_scale, int M, int K, int use_division, float eps,
                             int dtype, void* stream);
/* Per-(token, 128-group): scale = max(absmax, eps) / 448, q = sat(x / scale); replaces
 * sgl_per_token_group_quant_fp8 (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:19-197).
 * scales[m * scale_row_stride + g * scale_col_stride] (row-major, column-major or TMA-aligned
 * layouts of fp8_kernel.py:218-238).                                                          */
int b200_fp8_quant_group128(const void* x, int64_t ldx, void* q, int64_t ldq, float* scales,
                            int64_t scale_row_stride, int64_t scale_col_stride, int M, int K,
                            float eps, int dtype, void* stream);

/* ---- FP8 W8A8 GEMM, per-token x per-channel scales (tcgen05) ---------------- */
/* Replaces sgl_kernel.fp8_scaled_mm (sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1071-1146):
 *   out[M,N] = (A[M,K] . B[N,K]^T) * sa[m*sa_stride] * sb[n*sb_stride] (+ bias[n])
 * A e4m3 row-major (lda), B e4m3 stored [N,K] K-contiguous (ldb) — i.e. the reference's
 * column-major [K,N] weight view; sa/sb fp32 (stride 0 = per-tensor); bias/out bf16 or f16.
 * B is treated as a WEIGHT: with programmatic dependent launch (b200_set_pdl, on by default) the
 * kernels start fetching B before the previous kernel of the stream has finished, so B must not be
 * produced by a b200 kernel launched immediately before on the same stream (torch kernels and the
 * library's load-time quantisers never release their dependents early, so they are safe).  The same
 * holds for the weight operand of b200_fp8_gemm_block128, b200_gemm_16bit and the W4A16 GEMV.     */
int b200_fp8_gemm_rowcol(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                         int sa_stride, const float* sb, int sb_stride, const void* bias,
                         void* out, int64_t ldc, int M, int N, int K, int out_dtype,
                         void* stream);

/* M <= max_m (default 128, the decode regime) is served by a dedicated weight-streaming kernel
 * (swap-AB tcgen05 tiles, thread-block-cluster split-K reduced through distributed shared memory);
 * this hook lets tests and sweeps force either path: max_m = 0 disables it, splits > 0 fixes the
 * cluster size, negative values restore the defaults (B200_SMALLM_MAX_M / B200_SMALLM_SPLITS).
 * Applies to b200_fp8_gemm_rowcol and b200_fp8_gemm_block128.                                   */
int b200_fp8_gemm_smallm_config(int max_m, int splits);

/* ---- FP8 block-scaled GEMM (128x128 weight blocks, 1x128 activation groups) -------- */
/* Replaces fp8_blockwise_scaled_mm (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423) and
 * w8a8_block_fp8_matmul (python/sglang/srt/layers/quantization/fp8_kernel.py:774-891):
 *   out[m,n] = sum_kb (A[m,kb] . B[n,kb]^T) * as[m*as_row_stride + kb*as_col_stride]
 *                                           * bs[(n/128)*bs_row_stride + kb]        (+ bias[n])
 * A [M,K] e4m3 (lda), B [N,K] e4m3 (ldb), scales fp32, out bf16/f16.                         */
int b200_fp8_gemm_block128(const void* a, int64_t lda, const void* b, int64_t ldb, const float* as,
                           int64_t as_row_stride, int64_t as_col_stride, const float* bs,
                           int64_t bs_row_stride, const void* bias, void* out, int64_t ldc, int M,
                           int N, int K, int out_dtype, void* stream);

/* ---- dense 16-bit GEMM (unquantised layers, lm_head) ------------------------------------------ */
/* out[M,N] = x[M,K] . w[N,K]^T (+ bias[N]); bf16 / f16 operands and output, fp32 accumulation on
 * tcgen05 (same persistent TMA kernel as the FP8 path, kind::f16).  Replaces F.linear in
 * UnquantizedLinearMethod.apply (python/sglang/srt/layers/linear.py:170-176) and the lm_head matmul of
 * LogitsProcessor._get_logits (layers/logits_processor.py:425-455).  x / w 16-byte aligned, row
 * strides (in elements) multiples of 8.                                                         */
int b200_gemm_16bit(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias,
                    void* out, int64_t ldc, int M, int N, int K, int dtype, void* stream);

/* ---- greedy sampling ----------------------------------------------------------------------- */
/* out[m] = index of the FIRST maximum of row m of x [M, V] (bf16 / f16, row stride ldx elements):
 * torch.argmax(logits, -1) of Sampler.forward (python/sglang/srt/layers/sampler.py:60-66), with a row
 * spread over a cluster of CTAs instead of torch's single CTA per row.                           */
int b200_argmax_rows(const void* x, int64_t ldx, int64_t* out, int M, int V, int dtype, void* stream);

/* ---- INT4 weight-only (W4A16): AWQ / GPTQ ------------------------------------- */
/* Bit-exact AWQ dequant with the reference layouts (replaces sgl_kernel.awq_dequantize,
 * sgl-kernel/csrc/gemm/awq_kernel.cu:126-219): qweight [K, N/8] int32, qzeros [K/g, N/8] int32,
 * scales [K/g, N] f16/bf16 -> out [K, N]; out[k, 8c+j] = (nib(qweight[k,c], ORDER[j]) -
 * nib(qzeros[k/g,c], ORDER[j])) * scales[k/g, 8c+j], ORDER = [0,4,1,5,2,6,3,7].                 */
int b200_awq_dequant(const void* qweight, const void* scales, const void* qzeros, void* out, int K,
                     int N, int group_size, int dtype, void* stream);
/* One-time repacks used by process_weights_after_loading (awq.py:192, vllm GPTQLinearMethod):
 *   packed zeros [G, N/8] -> 16-bit zeros [G, N] holding MAGIC + z (+1 for GPTQ v1), MAGIC = 1024
 *   (f16) / 128 (bf16); awq_order selects the AWQ interleaved or GPTQ sequential nibble order.
 *   GPTQ qweight [K/8, N] (K-packed) -> [K, N/8] in the AWQ nibble order; perm (nullable, int32 [K])
 *   gathers source rows (desc_act checkpoints: perm = argsort(g_idx)).                          */
int b200_w4a16_unpack_zeros(const void* qzeros, void* zeros_out, int num_groups, int N,
                            int awq_order, int add_one, int dtype, void* stream);
int b200_w4a16_repack_gptq(const void* qweight, const int32_t* perm, void* out, int K, int N,
                           void* stream);
/* Fused dequant + GEMM on tcgen05 (replaces awq_dequantize + torch.matmul, awq.py:203-215, and
 * vllm ops.gptq_gemm): out[M,N] = x[M,K] @ W[K,N] (+ bias), W[k,n] = round16((w - z) * s) with
 * qweight [K, N/8] in AWQ order, zeros/scales [K/g, N] as produced above; x/out/bias f16 or bf16. */
int b200_w4a16_gemm(const void* x, int64_t ldx, const void* qweight, const void* zeros,
                    const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                    int K, int group_size, int dtype, void* stream);

/* Decode-regime GEMV (1 <= M <= 16, N %% 32 == 0, group %% 16 == 0): the packed weights are streamed
 * once from HBM (0.5 byte / element) and consumed by mma.sync as exact integers,
 *   out[m,n] = sum_g s[g,n] * (sum_{k in g} x[m,k] w[k,n] - z[g,n] sum_{k in g} x[m,k]),
 * split-K over a thread-block cluster with a fixed-order reduction.  Same operands as
 * b200_w4a16_gemm, which dispatches here for M <= 16; b200_w4a16_gemv_config(max_m, splits) lets
 * tests and sweeps force either kernel (max_m = 0: never, negative: defaults).                   */
int b200_w4a16_gemv(const void* x, int64_t ldx, const void* qweight, const void* zeros,
                    const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                    int K, int group_size, int dtype, void* stream);
int b200_w4a16_gemv_config(int max_m, int splits);

/* Decode-regime GEMV, second generation (csrc/w4a16_gemv2.cu): the same arithmetic as
 * b200_w4a16_gemv on a one-time STREAM repack of the packed weight — tiles of 16 output columns x
 * 64 k, 512 contiguous bytes, ordered [n / 16][k / 64], each lane's 16 bytes being its mma.sync A
 * fragments — so that weights go HBM -> registers -> tensor core with no shared-memory staging.
 *   b200_w4a16_gemv_v2_repack:    qweight [K, N/8] (AWQ nibble order, the layout b200_w4a16_gemm
 *                                 takes) -> stream layout, K*N/2 bytes; K %% 64 == 0, N %% 16 == 0.
 *   b200_w4a16_gemv_v2:           1 <= M <= 16; zeros / scales / bias / x / out as b200_w4a16_gemm.
 *   b200_w4a16_gemv_v2_supported: 1 when the shape is accepted (group 32 or a multiple of 64).
 *   b200_w4a16_gemv_v2_config:    splits > 0 fixes the split-K cluster size (tests / sweeps).
 * Replaces awq_dequantize + torch.matmul (awq.py:203-215) / vllm ops.gptq_gemm for decode batches. */
int b200_w4a16_gemv_v2_repack(const void* qweight, void* out, int K, int N, void* stream);
int b200_w4a16_gemv_v2(const void* x, int64_t ldx, const void* qweight_stream, const void* zeros,
                       const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                       int K, int group_size, int dtype, void* stream);
int b200_w4a16_gemv_v2_supported(int M, int N, int K, int group_size);
int b200_w4a16_gemv_v2_config(int splits);

/* ---- steps either side of the GEMMs / attention ------------------ */
/* (residual +) RMSNorm, optionally fused with the FP8 activation quantiser of the linear layer that
 * consumes it.  Replaces RMSNorm.forward (python/sglang/srt/layers/layernorm.py:61-108; oracle =
 * forward_native :88-108; GPU = sgl_kernel rmsnorm / fused_add_rmsnorm) followed by
 * sgl_per_token_quant_fp8 / per_token_group_quant_fp8 (fp8_utils.py:454-648):
 *   s = x (+ residual) in fp32; residual <- T(s) (in place, if given);
 *   y = T((s * rsqrt(mean(s^2) + eps)) * weight)
 *   out (nullable) <- y;  quant_mode 1: scales[m] = absmax(y)/448, q = sat(y * (1/scale));
 *   quant_mode 2: per 128 columns, scales[m*row_stride + g*col_stride] = max(absmax, quant_eps)/448,
 *   q = sat(y / scale).  Same bytes as b200_fp8_quant_* applied to `out`.  K %% 8 == 0, K <= 16384. */
int b200_rmsnorm(const void* x, int64_t ldx, void* residual, int64_t ldr, const void* weight,
                 void* out, int64_t ldo, void* q, int64_t ldq, float* scales,
                 int64_t scale_row_stride, int64_t scale_col_stride, int M, int K, float eps,
                 int quant_mode, float quant_eps, int dtype, void* stream);
/* SiluAndMul (python/sglang/srt/layers/activation.py:42-51; oracle forward_native :43-45):
 *   y = T(T(silu(x[:, :d])) * x[:, d:]), x [M, 2d]; out / q / scales / quant_mode as above.     */
int b200_silu_and_mul(const void* x, int64_t ldx, void* out, int64_t ldo, void* q, int64_t ldq,
                      float* scales, int64_t scale_row_stride, int64_t scale_col_stride, int M,
                      int d, int quant_mode, float quant_eps, int dtype, void* stream);
/* Rotary embedding of q and k IN PLACE, optionally fused with the KV-pool write of (rotated k, v).
 * Replaces RotaryEmbedding.forward (python/sglang/srt/layers/rotary_embedding.py:119-176; oracle
 * forward_native + _apply_rotary_emb :30-57; GPU sgl-kernel apply_rope_with_cos_sin_cache_inplace)
 * and MHATokenToKVPool.set_kv_buffer (memory_pool.py:378-412).
 *   positions int64 [T]; q [T, Hq*D], k [T, Hkv*D], v [T, Hkv*D] token strides in elements;
 *   cos_sin_cache [max_pos, rotary_dim] (cos half then sin half), fp32 (B200_F32) or the
 *   activation dtype; is_neox: pairs (j, j + rotary_dim/2), else adjacent pairs;
 *   native_rounding != 0 reproduces torch's 16-bit arithmetic of forward_native bit for bit,
 *   0 computes in fp32 with one rounding (the reference's CUDA kernel);
 *   k_cache/v_cache (nullable, together): rows loc[t] receive the rotated k and v of token t. */
int b200_rope_kv_write(const int64_t* positions, void* q, int64_t q_stride, void* k,
                       int64_t k_stride, const void* v, int64_t v_stride,
                       const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                       int native_rounding, void* k_cache, void* v_cache, int64_t cache_row_stride,
                       const int64_t* loc, int T, int Hq, int Hkv, int D, int dtype, void* stream);

/* Same with an e4m3 KV pool (rows of bytes, cache_row_stride in bytes): the rotated k and v are
 * quantised exactly as b200_kv_write_fp8 does (torch's float8_e4m3fn cast after the optional
 * division by k_scale / v_scale, memory_pool.py:390-396; FlashInfer's set_kv_buffer call,
 * flashinfer_backend.py:427), so one launch replaces RoPE + set_kv_buffer on an FP8 pool. */
int b200_rope_kv_write_fp8(const int64_t* positions, void* q, int64_t q_stride, void* k,
                           int64_t k_stride, const void* v, int64_t v_stride,
                           const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                           int native_rounding, void* k_cache, void* v_cache,
                           int64_t cache_row_stride, const int64_t* loc, float k_scale,
                           float v_scale, int has_scale, int T, int Hq, int Hkv, int D, int dtype,
                           void* stream);

/* ---- tensor-parallel all-reduce over NVLink peer memory ------------------------------------- */
/* Replaces the reference's custom all-reduce for decode-sized messages: host class CustomAllreduce
 * (python/sglang/srt/distributed/device_communicators/custom_all_reduce.py:146-510; buffers are
 * cudaMalloc'ed, exported with cudaIpcGetMemHandle and all-gathered, :303-326), kernels
 * sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:127-225, called from GroupCoordinator.all_reduce
 * (parallel_state.py:398-461).  One process per GPU; every rank allocates ONE symmetric buffer of
 * b200_comm_buffer_bytes(world, max_bytes) with b200_comm_alloc (zero-filled; `handle_out`
 * receives b200_comm_handle_bytes() bytes to send to the peers through torch.distributed), opens
 * the peers' handles with b200_comm_open and builds the communicator from the `world` pointers
 * (own buffer at index `rank`).  Tests may pass `world` buffers of ONE device and drive each rank
 * from its own stream.  All ranks must issue the same sequence of collective calls. */
int b200_comm_handle_bytes(void);
size_t b200_comm_buffer_bytes(int world, size_t max_bytes);
int b200_comm_alloc(size_t bytes, void** ptr, void* handle_out);
int b200_comm_open(const void* handle, void** ptr);
int b200_comm_close(void* ptr);
int b200_comm_free(void* ptr);
int b200_comm_create(int rank, int world, void* const* peer_bufs, size_t buf_bytes,
                     void** comm_out);
int b200_comm_destroy(void* comm);
/* Largest message (bytes) one call can reduce (= one receive slot). */
size_t b200_comm_max_bytes(const void* comm);
/* out = sum over ranks of in (fp32 accumulation in rank order => bit-identical on every rank, one
 * rounding); one-shot: every rank pushes its vector into all peers' slots, flags, local reduce.
 * in may equal out.  numel * sizeof(dtype) must be a multiple of 16 (custom_all_reduce.py:418).  */
int b200_comm_all_reduce(void* comm, const void* in, void* out, int64_t numel, int dtype,
                         void* stream);

/* Row-parallel linear = GEMM -> all-reduce (RowParallelLinear.forward, layers/linear.py:1277-1299)
 * as ONE kernel in the decode regime (M <= 128): the swap-AB small-M kernel's epilogue pushes its
 * scaled bf16/f16 tile to every TP peer, waits for the peers' copies of the same tile and stores
 * their rank-ordered fp32 sum, so the transfer of one tile overlaps the weight streaming of the
 * others.  Result = round(sum_r float(round(out_r))) exactly.  Larger M: big-tile GEMM followed by
 * b200_comm_all_reduce in place (out must then be contiguous, ldc == N).  Arguments as the
 * un-fused entry points; bias must be passed on rank 0 only (linear.py:1290).                  */
int b200_fp8_gemm_rowcol_ar(void* comm, const void* a, int64_t lda, const void* b, int64_t ldb,
                            const float* sa, int sa_stride, const float* sb, int sb_stride,
                            const void* bias, void* out, int64_t ldc, int M, int N, int K,
                            int out_dtype, void* stream);
/* Largest M the *_ar entry points fuse (default 32, measured crossover; max 128; < 0 = default). */
int b200_comm_fuse_config(int max_m);

/* NVLS: all-reduce with the additions done inside the NVSwitch (multimem.ld_reduce over a multicast
 * mapping; csrc/comm_nvls.cu).  SURVEY section 8f row 3; reference analogue = the one-shot kernel of
 * custom_all_reduce.cuh:203-225 (adds on the SM) and NCCL's NVLS algorithms.  Life cycle (host side:
 * distributed/device_communicators/b200_all_reduce.py):
 *   b200_nvls_supported()                      1 when device + driver offer multicast objects
 *   b200_nvls_buffer_bytes(world, bytes)       size rounded to the multicast / allocation granularity
 *   b200_nvls_create(world, size, &buf, &fd)   rank 0: multicast object + POSIX fd for the peers
 *   b200_nvls_import(fd, world, size, &buf)    peers: import it (fd passed by SCM_RIGHTS)
 *   b200_nvls_add_device(buf)                  every rank; host barrier; then
 *   b200_nvls_bind(buf, &uc, &mc)              every rank: own physical memory bound + both views mapped
 *   b200_comm_attach_nvls(comm, buf)           b200_comm_all_reduce now uses the switch for messages of
 *                                              B200_COMM_NVLS_MIN (128 KiB) .. size / 2 bytes
 *   b200_nvls_free(buf)                                                                           */
int b200_nvls_supported(void);
size_t b200_nvls_buffer_bytes(int world, size_t bytes);
int b200_nvls_create(int world, size_t size, void** buf_out, int* fd_out);
int b200_nvls_import(int fd, int world, size_t size, void** buf_out);
int b200_nvls_add_device(void* buf);
int b200_nvls_bind(void* buf, void** uc_ptr, void** mc_ptr);
int b200_nvls_free(void* buf);
int b200_comm_attach_nvls(void* comm, void* nvls_buf);
int b200_fp8_gemm_block128_ar(void* comm, const void* a, int64_t lda, const void* b, int64_t ldb,
                              const float* as, int64_t as_row_stride, int64_t as_col_stride,
                              const float* bs, int64_t bs_row_stride, const void* bias, void* out,
                              int64_t ldc, int M, int N, int K, int out_dtype, void* stream);

/* ---- scheduler-side index kernels (page_size >= 1) --------------------------------------------- */
/* int64/int32 index arithmetic, bit-exact with the reference's Triton kernels.  All tensors device.
 * write_req_to_token_pool_triton (managers/schedule_batch.py:1789-1823, called :1272-1283):
 *   req_to_token[req_pool_indices[b], pre_lens[b] : seq_lens[b]] =
 *       out_cache_loc[sum(extend_lens[:b]) : sum(extend_lens[:b]) + seq_lens[b] - pre_lens[b]]    */
int b200_write_req_to_token(int32_t* req_to_token, int64_t req_to_token_stride,
                            const int64_t* req_pool_indices, const int64_t* pre_lens,
                            const int64_t* seq_lens, const int64_t* extend_lens,
                            const int64_t* out_cache_loc, int bs, void* stream);
/* get_last_loc (managers/schedule_batch.py:1826-1893):
 *   result[i] = prefix_lens[i] > 0 ? req_to_token[req_pool_indices[i], prefix_lens[i] - 1] : -1   */
int b200_get_last_loc(const int32_t* req_to_token, int64_t req_to_token_stride,
                      const int64_t* req_pool_indices, const int64_t* prefix_lens, int64_t* result,
                      int n, void* stream);
/* alloc_extend_kernel / alloc_decode_kernel of PagedTokenToKVPoolAllocator
 * (mem_cache/paged_allocator.py:29-155): token slot ids of the new tokens of every request — first
 * the rest of the partially filled last page (last_loc + 1 ...), then whole pages taken in order
 * from free_pages, then the head of one more page.  *ret_values = (new pages << 32) | new tokens
 * (extend) or the number of new pages (decode), which the host uses to pop free_pages.           */
int b200_alloc_extend(const int64_t* pre_lens, const int64_t* seq_lens, const int64_t* last_loc,
                      const int64_t* free_pages, int64_t* out_indices, int64_t* ret_values, int bs,
                      int page_size, void* stream);
int b200_alloc_decode(const int64_t* seq_lens, const int64_t* last_loc, const int64_t* free_pages,
                      int64_t* out_indices, int64_t* ret_values, int bs, int page_size,
                      void* stream);
/* compute_position_kernel (model_executor/forward_batch_info.py:639-692): positions of the new
 * tokens (prefix_len + i) and extend_start_loc = exclusive cumsum of extend_seq_lens.
 * extend_prefix_lens may be NULL (no prefix).                                                     */
int b200_compute_position(int64_t* positions, int32_t* extend_start_loc,
                          const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens, int bs,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_HOTPATH_H_ */
