"""GPTQ arithmetic against the dependency the reference delegates it to: vllm's own `ops.gptq_gemm`
(python/sglang/srt/layers/quantization/gptq.py:17-31 imports vllm's GPTQLinearMethod).  The reference
pins vllm 0.8.4; the GPU image carries a later vllm wheel, used here ONLY as the checker (skipped when
the wheel or its compiled ops are unavailable).  AutoGPTQ v1 checkpoint format: stored zero = z - 1."""
import numpy as np
import pytest
import torch

from oracle import int4 as O

pytestmark = pytest.mark.gpu


def _vllm_gptq_gemm(x, qweight, qzeros, scales, g_idx):
    try:
        import vllm._custom_ops as ops

        return ops.gptq_gemm(x, qweight, qzeros, scales, g_idx, False, False, 4)
    except Exception as e:  # noqa: BLE001 - no wheel / no compiled ops / different signature
        pytest.skip(f"vllm gptq_gemm unavailable: {type(e).__name__}: {str(e)[:120]}")


@pytest.mark.parametrize("m", [4, 45])
def test_gptq_linear_method_matches_vllm_gptq_gemm(m):
    from sglang_security_b200.layers.linear import RowParallelLinear
    from sglang_security_b200.layers.quantization.gptq import GPTQConfig

    k, n, group, dtype = 1024, 768, 128, torch.float16
    rs = np.random.RandomState(7)
    q = torch.from_numpy(rs.randint(0, 16, size=(k, n)).astype(np.int64))
    z = torch.from_numpy(rs.randint(0, 15, size=(k // group, n)).astype(np.int64))   # stored zero, z + 1 <= 15
    qweight = sum((q.view(k // 8, 8, n)[:, j] << (4 * j)) for j in range(8))
    qzeros = sum((z.view(k // group, n // 8, 8)[..., j] << (4 * j)) for j in range(8))
    to_i32 = lambda t: torch.where(t >= 2 ** 31, t - 2 ** 32, t).to(torch.int32)
    qweight, qzeros = to_i32(qweight), to_i32(qzeros)
    g = torch.Generator().manual_seed(3)
    scales = (torch.rand((k // group, n), generator=g) * 0.02 + 0.001).to(dtype)
    g_idx = (torch.arange(k, dtype=torch.int32) // group)
    x = torch.randn((m, k), generator=g).to(dtype)
    ref = _vllm_gptq_gemm(x.cuda(), qweight.cuda(), qzeros.cuda(), scales.cuda(), g_idx.cuda()).cpu()
    # the oracle restatement against the dependency's kernel
    orc = O.w4a16_linear_oracle(x, O.gptq_dequantize_oracle(qweight, scales, qzeros, g_idx, group))
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert ((orc.float() - ref.float()).abs() <= tol).all()
    # the B200 linear method against the dependency's kernel
    layer = RowParallelLinear(k, n, bias=False, params_dtype=dtype,
                              quant_config=GPTQConfig(4, group, False), tp_rank=0, tp_size=1).cuda()
    layer.qweight.data.copy_(qweight)
    layer.qzeros.data.copy_(qzeros)
    layer.scales.data.copy_(scales)
    layer.g_idx.data.copy_(g_idx)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x.cuda())
    assert ((out.cpu().float() - ref.float()).abs() <= tol).all()
