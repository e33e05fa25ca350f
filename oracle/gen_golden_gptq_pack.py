"""Generates tests/golden/gptq_pack_vllm.pt: the GPTQ qweight packing ([K, N] 4-bit integers ->
[K/8, N] int32, nibble k % 8) produced by the DEPENDENCY the reference delegates GPTQ to — vllm's
own `gptq_pack` (vllm/model_executor/layers/quantization/utils/quant_utils.py).  The reference pins
vllm 0.8.4 (python/sglang/srt/layers/quantization/gptq.py:17-31 imports it); the wheel present in
this image is a later vllm (version recorded in the fixture) — the checkpoint format did not change.
This pins the PACKING ORDER of oracle/int4.py::unpack_gptq_qweight only; the dequant arithmetic
(zero + 1, exllama CUDA kernels) has no CPU implementation in vllm; it is checked on the GPU box (tests/test_gptq_vs_vllm_gpu.py).
TEST INFRASTRUCTURE ONLY."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    import vllm
    from vllm.model_executor.layers.quantization.utils.quant_utils import gptq_pack

    g = torch.Generator().manual_seed(0)
    cases = []
    for k, n in [(8, 8), (128, 64), (512, 384)]:
        q = torch.randint(0, 16, (k, n), generator=g, dtype=torch.int32)
        packed = gptq_pack(q, 4, k, n)
        cases.append(dict(q=q, qweight=packed.to(torch.int32)))
    out = os.path.join(ROOT, "tests", "golden", "gptq_pack_vllm.pt")
    torch.save(dict(vllm_version=vllm.__version__, cases=cases), out)
    print(out, vllm.__version__, [tuple(c["qweight"].shape) for c in cases])


if __name__ == "__main__":
    sys.exit(main())
