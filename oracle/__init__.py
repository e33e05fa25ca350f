"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's algorithms for the hot path.  Only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import this package, and only as the checker or the timed CPU baseline — never as (part of) the
product path.  Nothing under `sglang_security_b200/` imports it.

Pinning status (details in DESIGN.md §3):
  attention decode/extend : pinned — checked against outputs of the reference's own
      torch_native_backend.py run in the authoring container (tests/golden/attn_*.pt, generated
      by oracle/gen_golden.py), against the reference sgl-kernel CPU kernels (oracle/_ref) and
      against the reference's own Triton decode / extend kernels run in Triton's CPU interpreter
      (tests/golden/triton_attn.pt, oracle/gen_golden_triton_attn.py).
  pools / kv-indices      : pinned by the same fixtures (bit-exact) and the reference test's
      python-loop definition (test/srt/test_create_kvindices.py:40-49); FP8 (e4m3 / e5m2) pool
      writes by the reference's own set_kv_buffer run on CPU (tests/golden/kv_write_fp8.pt).
  FP8 quant / GEMM, AWQ   : restatements of the torch references inside the reference's tests
      (file:line in each function); pinned by fixtures generated from those test files.
  scheduler index kernels : pinned — get_last_loc / compute_position by the reference's torch twins,
      alloc_extend / alloc_decode by the reference's own Triton kernels run in Triton's CPU
      interpreter (tests/golden/sched_index.pt, sched_alloc.pt).
  GPTQ                    : arithmetic lives in vllm 0.8.4 (absent, CUDA only): packing order pinned
      by vllm's gptq_pack (tests/golden/gptq_pack_vllm.pt); dequant arithmetic checked on the GPU box
      against the image's vllm wheel (tests/test_gptq_vs_vllm_gpu.py) - unpinned in the CPU suite.
"""
