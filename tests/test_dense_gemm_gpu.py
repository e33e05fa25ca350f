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


@pytest.mark.parametrize("m,v", [(1, 128256), (64, 128256), (3, 1000), (700, 32000), (5, 151936), (2, 8)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_argmax_rows_equals_torch_argmax(m, v, dtype):
    """Greedy sampling kernel: first index of the row maximum, ties included (bf16 logits tie often)."""
    from sglang_security_b200.layers.dense_kernel import argmax_rows

    g = torch.Generator(device="cuda").manual_seed(m * 7 + v)
    x = torch.randn((m, v), device="cuda", generator=g).to(dtype)
    x[0, : min(v, 5)] = x[0].max()            # a tie at the very start of a row
    if v > 300:
        x[-1, 299] = x[-1, v - 1] = x[-1].max() + 1  # a tie across CTAs of the cluster
    assert torch.equal(argmax_rows(x), torch.argmax(x, dim=-1))
    xs = torch.randn((m, v + 8), device="cuda", generator=g).to(dtype)[:, 3:v + 3]   # unaligned rows
    assert torch.equal(argmax_rows(xs), torch.argmax(xs, dim=-1))
