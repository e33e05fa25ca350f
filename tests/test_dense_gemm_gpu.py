"""GPU parity of the dense 16-bit tcgen05 GEMM (b200_gemm_16bit) behind UnquantizedLinearMethod and
the decode step's lm_head, against a plain fp32 torch reference of the same op (TF32 off)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n,k", [(1, 256, 128), (64, 1000, 4096), (130, 136, 72), (64, 16032, 4096),
                                   (300, 4096, 4096), (2500, 6144, 1024), (4100, 512, 264)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("with_bias", [False, True])
def test_gemm_16bit_matches_fp32_reference(m, n, k, dtype, with_bias):
    from sglang_security_b200._lib import load
    from sglang_security_b200.layers.dense_kernel import gemm_16bit

    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn((m, k), device="cuda", generator=g).to(dtype)
    w = (torch.randn((n, k), device="cuda", generator=g) * 0.05).to(dtype)
    b = torch.randn((n,), device="cuda", generator=g).to(dtype) if with_bias else None
    out = gemm_16bit(x, w, b)
    assert b"gemm_tn kind=" in load().b200_last_variant()
    ref = x.float() @ w.float().t()
    if b is not None:
        ref = ref + b.float()
    # one rounding of an fp32-accumulated sum to 16 bit: 2^-8 (bf16) / 2^-11 (f16) relative
    rtol = 2 ** -7 if dtype == torch.bfloat16 else 2 ** -10
    err = (out.float() - ref).abs()
    tol = rtol * ref.abs() + 1e-3 * ref.abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.abs().max().item())
    assert torch.equal(out, gemm_16bit(x, w, b))


def test_unquantized_linear_method_runs_the_in_tree_kernel():
    from sglang_security_b200._lib import load
    from sglang_security_b200.layers.linear import ColumnParallelLinear

    layer = ColumnParallelLinear(512, 384, bias=True, params_dtype=torch.bfloat16, tp_rank=0,
                                 tp_size=1).cuda()
    layer.weight.data.normal_(0, 0.05)
    layer.bias.data.normal_()
    x = torch.randn((3, 7, 512), device="cuda", dtype=torch.bfloat16)
    out, _ = layer(x)
    assert out.shape == (3, 7, 384)
    assert b"gemm_tn kind=1" in load().b200_last_variant()
    ref = torch.nn.functional.linear(x.float(), layer.weight.float(), layer.bias.float())
    assert (out.float() - ref).abs().max() <= 2 ** -6 * ref.abs().max()
