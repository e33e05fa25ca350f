// Registration shim for the reference's own CPU attention kernels (TEST INFRASTRUCTURE ONLY).
// The kernels themselves are compiled from the reference sources where they lie
// (/root/reference/sgl-kernel/csrc/cpu/{decode,extend}.cpp) by oracle/build_ref.py; this file only
// forward-declares the entry points and exposes them as torch ops, the way the reference's
// torch_extension_cpu.cpp does for its full op set.
#include <ATen/ATen.h>
#include <torch/library.h>

void decode_attention_cpu(at::Tensor& query, at::Tensor& k_cache, at::Tensor& v_cache,
                          at::Tensor& output, at::Tensor& key, at::Tensor& value, at::Tensor& loc,
                          at::Tensor& attn_logits, at::Tensor& req_to_token,
                          at::Tensor& req_pool_indices, at::Tensor& seq_lens, double sm_scale,
                          double logit_cap);

void extend_attention_cpu(at::Tensor& q_extend, at::Tensor& k_extend, at::Tensor& v_extend,
                          at::Tensor& o_extend, at::Tensor& k_buffer, at::Tensor& v_buffer,
                          at::Tensor& req_to_token, at::Tensor& req_pool_indices,
                          at::Tensor& seq_lens, at::Tensor& extend_seq_lens,
                          at::Tensor& extend_start_loc, int64_t max_len_extend, double sm_scale,
                          double logit_cap);

// W8A16 block-FP8 CPU linear (sgl-kernel/csrc/cpu/gemm_fp8.cpp:480-545) and its weight pre-pack
// (gemm.cpp:401-464), registered as torch_extension_cpu.cpp:229-250 does.
at::Tensor convert_weight_packed(at::Tensor& weight);
at::Tensor fp8_scaled_mm_cpu(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales2,
                             std::vector<int64_t> block_size, const std::optional<at::Tensor>& bias,
                             at::ScalarType out_dtype, bool is_vnni);

TORCH_LIBRARY_FRAGMENT(B200_REF_NS, m) {
  m.def("convert_weight_packed(Tensor weight) -> Tensor");
  m.impl("convert_weight_packed", c10::kCPU, &convert_weight_packed);
  m.def(
      "fp8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, int[] block_size, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("fp8_scaled_mm_cpu", c10::kCPU, &fp8_scaled_mm_cpu);
  m.def(
      "decode_attention_cpu(Tensor query, Tensor k_cache, Tensor v_cache, Tensor(a!) output, "
      "Tensor key, Tensor value, Tensor loc, Tensor attn_logits, Tensor req_to_token, "
      "Tensor req_pool_indices, Tensor seq_lens, float sm_scale, float logit_cap) -> ()");
  m.impl("decode_attention_cpu", c10::kCPU, &decode_attention_cpu);
  m.def(
      "extend_attention_cpu(Tensor q_extend, Tensor k_extend, Tensor v_extend, Tensor(a!) o_extend, "
      "Tensor k_buffer, Tensor v_buffer, Tensor req_to_token, Tensor req_pool_indices, "
      "Tensor seq_lens, Tensor extend_seq_lens, Tensor extend_start_loc, int max_len_extend, "
      "float sm_scale, float logit_cap) -> ()");
  m.impl("extend_attention_cpu", c10::kCPU, &extend_attention_cpu);
}
