"""GPU parity of the FP8 linear path (quantisers, tcgen05 GEMMs, LinearMethod.apply) against the
CPU oracle (oracle/fp8.py restatements of the reference's in-test torch references)."""
import pytest
import torch

from oracle import fp8 as O

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


def _rand_fp8(shape, g):
    # reference test distribution: uniform over the whole e4m3 range (test_fp8_gemm.py:18-22)
    x = (torch.rand(shape, generator=g) - 0.5) * 2 * 448.0
    return x.clamp(-448.0, 448.0).to(FP8)


@pytest.mark.parametrize("m,k", [(1, 512), (7, 4096), (128, 2048), (300, 14336), (5, 1032)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_per_token_quant_bit_exact(m, k, dtype):
    from sglang_security_b200.layers.quantization.fp8_kernel import sglang_per_token_quant_fp8

    g = torch.Generator().manual_seed(m * 7 + k)
    x = (torch.randn((m, k), generator=g) * 3).to(dtype)
    if m > 2:
        x[1].zero_()  # all-zero row: scale 0, zeros out
    q_ref, s_ref = O.per_token_quant_fp8_oracle(x)
    q, s = sglang_per_token_quant_fp8(x.cuda())
    assert torch.equal(s.cpu(), s_ref)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref.view(torch.uint8))


@pytest.mark.parametrize("m,k", [(1, 128), (33, 4096), (257, 7168)])
@pytest.mark.parametrize("column_major", [False, True])
def test_group_quant_bit_exact(m, k, column_major):
    from sglang_security_b200.layers.quantization.fp8_kernel import per_token_group_quant_fp8

    g = torch.Generator().manual_seed(m + k)
    x = (torch.randn((m, k), generator=g) * 2).to(torch.bfloat16)
    x[0, :128] = 0  # eps path
    q_ref, s_ref = O.per_token_group_quant_fp8_oracle(x, 128)
    q, s = per_token_group_quant_fp8(x.cuda(), 128, column_major_scales=column_major,
                                     scale_tma_aligned=column_major)
    assert s.shape == s_ref.shape
    if column_major:
        assert s.stride(0) == 1
    assert torch.equal(s.cpu(), s_ref)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref.view(torch.uint8))


def test_static_and_rowwise_division_quant():
    from sglang_security_b200.layers.quantization.fp8_kernel import (per_token_group_quant_fp8,
                                                                    static_quant_fp8)

    g = torch.Generator().manual_seed(3)
    x = torch.randn((19, 1024), generator=g).to(torch.bfloat16)
    scale = torch.tensor([0.0123])
    q, s = static_quant_fp8(x.cuda(), scale.cuda(), repeat_scale=True)
    assert torch.equal(q.cpu().view(torch.uint8),
                       O.static_quant_fp8_oracle(x, scale).view(torch.uint8))
    assert s.shape == (19, 1) and torch.equal(s.cpu(), scale.expand(19, 1))
    # group_size == K (load-time per-channel weight quantisation, fp8.py:340-343)
    q_ref, s_ref = O.per_token_group_quant_fp8_oracle(x, 1024)
    q, s = per_token_group_quant_fp8(x.cuda(), 1024)
    assert torch.equal(s.cpu(), s_ref)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref.view(torch.uint8))


@pytest.mark.parametrize("m", [1, 17, 128, 300, 1024])
@pytest.mark.parametrize("n,k", [(16, 512), (128, 512), (512, 1040), (4096, 4096), (6144, 4096),
                                 (1000, 2048)])
@pytest.mark.parametrize("with_bias", [False, True])
def test_fp8_scaled_mm_reference_distribution(m, n, k, with_bias):
    """Shapes/tolerances of sgl-kernel/tests/test_fp8_gemm.py:16-49 (rtol 0.02, atol 1)."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    g = torch.Generator().manual_seed(m * 31 + n * 7 + k)
    a = _rand_fp8((m, k), g)
    b = _rand_fp8((n, k), g)
    sa = torch.randn((m,), generator=g) * 0.001
    sb = torch.randn((n,), generator=g) * 0.001
    out_dtype = torch.bfloat16
    bias = torch.randn((n,), generator=g).to(out_dtype) if with_bias else None
    if m * n * k > 2 ** 33:
        # too slow on the host: the same oracle function on the GPU in exact fp32 (TF32 off)
        torch.backends.cuda.matmul.allow_tf32 = False
        ref = O.scaled_mm_oracle(a.cuda(), b.cuda().t(), sa.cuda(), sb.cuda(), out_dtype,
                                 None if bias is None else bias.cuda()).cpu()
    else:
        ref = O.scaled_mm_oracle(a, b.t(), sa, sb, out_dtype, bias)
    out = fp8_scaled_mm(a.cuda(), b.cuda().t(), sa.cuda(), sb.cuda(), out_dtype,
                        None if bias is None else bias.cuda())
    torch.testing.assert_close(out.cpu(), ref, rtol=0.02, atol=1)
    # tighter: fp32 accumulation of exact fp8 products, one bf16 rounding (+ bias rounding)
    err = (out.cpu().float() - ref.float()).abs().max().item()
    assert err <= 2e-2 * max(1.0, ref.float().abs().max().item()), err


def test_fp8_scaled_mm_large_tile_path_and_f16():
    """M large enough for the 128x256 tile configuration; f16 output; scalar scales."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    g = torch.Generator().manual_seed(5)
    m, n, k = 2500, 4096, 1024
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    sa = torch.rand((m,), generator=g) * 1e-3
    sb = torch.rand((n,), generator=g) * 1e-3
    ref = O.scaled_mm_oracle(a, b.t(), sa, sb, torch.float16, None)
    out = fp8_scaled_mm(a.cuda(), b.cuda().t(), sa.cuda(), sb.cuda(), torch.float16)
    torch.testing.assert_close(out.cpu(), ref, rtol=0.02, atol=1)
    ref1 = O.scaled_mm_oracle(a, b.t(), sa[:1].expand(m), sb[:1].expand(n), torch.float16, None)
    out1 = fp8_scaled_mm(a.cuda(), b.cuda().t(), sa[:1].cuda(), sb[:1].cuda(), torch.float16)
    torch.testing.assert_close(out1.cpu(), ref1, rtol=0.02, atol=1)


@pytest.mark.parametrize("m,n,k", [(1, 128, 128), (16, 512, 7168), (83, 7168, 1024),
                                   (300, 1536, 2048), (1024, 2112, 1152)])
def test_fp8_block_scaled_mm(m, n, k):
    """python/sglang/test/test_block_fp8.py:328-396: rel-mean-err < 1e-3 vs the native reference."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_blockwise_scaled_mm

    g = torch.Generator().manual_seed(m + n + k)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    a_s = torch.rand((m, (k + 127) // 128), generator=g) * 1e-2
    b_s = torch.rand(((n + 127) // 128, (k + 127) // 128), generator=g) * 1e-2
    ref = O.block_scaled_mm_oracle(a, b, a_s, b_s, (128, 128), torch.bfloat16)
    out = fp8_blockwise_scaled_mm(a.cuda(), b.cuda(), a_s.cuda(), b_s.cuda(), torch.bfloat16)
    rel = (out.cpu().float() - ref.float()).abs().mean() / ref.float().abs().mean()
    assert rel.item() < 1e-3, rel.item()
    # column-major activation scales (fp8_kernel.py:218-238) give the same result
    a_s_col = a_s.t().contiguous().t()
    out2 = fp8_blockwise_scaled_mm(a.cuda(), b.cuda(), a_s_col.cuda().t().contiguous().t(),
                                   b_s.cuda(), torch.bfloat16)
    assert torch.equal(out, out2)


def _make_linear(cls, quant_config, n_parts, k, dtype, bias):
    from sglang_security_b200.layers.linear import MergedColumnParallelLinear

    torch.manual_seed(0)
    layer = MergedColumnParallelLinear(k, n_parts, bias=bias, params_dtype=dtype,
                                       quant_config=quant_config, tp_rank=0, tp_size=1)
    return layer.cuda()


@pytest.mark.parametrize("m", [1, 64, 333])
def test_fp8_linear_method_16bit_checkpoint(m):
    """Fp8LinearMethod / W8A8Fp8LinearMethod on a bf16 checkpoint: per-channel weight quant at
    load, dynamic per-token activation quant, fused scaled GEMM (+bias)."""
    from sglang_security_b200.layers.quantization.fp8 import Fp8Config
    from sglang_security_b200.layers.quantization.w8a8_fp8 import W8A8Fp8Config

    g = torch.Generator().manual_seed(m)
    k, parts = 1024, [512, 256]
    w = (torch.randn((sum(parts), k), generator=g) * 0.02).to(torch.bfloat16)
    bias = torch.randn((sum(parts),), generator=g).to(torch.bfloat16)
    x = torch.randn((m, k), generator=g).to(torch.bfloat16)
    wq_ref, ws_ref = O.per_token_group_quant_fp8_oracle(w, k)
    ref = O.fp8_linear_oracle(x, wq_ref, ws_ref, bias)
    for cfg in (Fp8Config(False, "dynamic"), W8A8Fp8Config(False)):
        layer = _make_linear(None, cfg, parts, k, torch.bfloat16, True)
        layer.weight.data.copy_(w)
        layer.bias.data.copy_(bias)
        layer.quant_method.process_weights_after_loading(layer)
        assert layer.weight.dtype == FP8 and layer.weight.shape == (k, sum(parts))
        assert torch.equal(layer.weight.t().cpu().view(torch.uint8), wq_ref.view(torch.uint8))
        assert torch.equal(layer.weight_scale.cpu().reshape(-1), ws_ref.reshape(-1))
        out, _ = layer(x.cuda())
        st = (out.cpu().float() - ref.float()).abs()
        assert st.max().item() <= 2e-2 * max(1.0, ref.float().abs().max().item()), st.max()


def test_fp8_linear_method_serialized_checkpoint_static_and_block():
    from sglang_security_b200.layers.quantization.fp8 import Fp8Config

    g = torch.Generator().manual_seed(9)
    k, parts, m = 2048, [256, 128, 128], 37
    n = sum(parts)
    w = _rand_fp8((n, k), g)
    x = torch.randn((m, k), generator=g).to(torch.bfloat16)
    # (a) per-shard tensor scales + static input scale
    layer = _make_linear(None, Fp8Config(True, "static"), parts, k, torch.bfloat16, False)
    shard_scales = torch.tensor([1e-3, 2e-3, 5e-4])
    in_scales = torch.tensor([0.02, 0.03, 0.025])
    layer.weight.data.copy_(w.cuda().view(torch.uint8).view(FP8))
    layer.weight_scale.data.copy_(shard_scales)
    layer.input_scale.data.copy_(in_scales)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x.cuda())
    ws = torch.repeat_interleave(shard_scales, torch.tensor(parts))
    ref = O.fp8_linear_oracle(x, w, ws, None, input_scale=in_scales.max())
    assert (out.cpu().float() - ref.float()).abs().max().item() <= 2e-2 * ref.float().abs().max().item()
    # (b) 128x128 block scales, dynamic activations
    layer = _make_linear(None, Fp8Config(True, "dynamic", weight_block_size=[128, 128]), parts, k,
                         torch.bfloat16, False)
    b_s = torch.rand((n // 128, k // 128), generator=g) * 1e-3
    layer.weight.data.copy_(w.cuda().view(torch.uint8).view(FP8))
    layer.weight_scale_inv.data.copy_(b_s)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x.cuda())
    ref = O.block_fp8_linear_oracle(x, w, b_s)
    rel = (out.cpu().float() - ref.float()).abs().mean() / ref.float().abs().mean()
    assert rel.item() < 1e-3, rel.item()


@pytest.fixture
def smallm_cfg():
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_gemm_smallm_config

    yield fp8_gemm_smallm_config
    fp8_gemm_smallm_config(-1, -1)


@pytest.mark.parametrize("m", [1, 3, 16, 17, 33, 64, 100, 128])
@pytest.mark.parametrize("n,k,splits", [(4096, 4096, 0), (6144, 4096, 3), (1000, 2048, 8),
                                        (256, 14336, 5), (4096, 1152, 8), (128, 128, 2)])
def test_fp8_smallm_cluster_splitk_matches_oracle_and_big_tile_path(m, n, k, splits, smallm_cfg):
    """Decode-regime kernel (swap-AB, cluster split-K over DSMEM) against the oracle and against
    the big-tile kernel on identical inputs; forced cluster sizes incl. ragged K splits (empty
    ranks), N % 128 != 0, M % 16 != 0; two runs must be bit-identical (fixed reduction order)."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    g = torch.Generator().manual_seed(m * 131 + n + k + splits)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    sa = torch.randn((m,), generator=g) * 0.001
    sb = torch.randn((n,), generator=g) * 0.001
    bias = torch.randn((n,), generator=g).to(torch.bfloat16)
    ref = O.scaled_mm_oracle(a, b.t(), sa, sb, torch.bfloat16, bias)
    args = (a.cuda(), b.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, bias.cuda())
    smallm_cfg(128, splits)
    out = fp8_scaled_mm(*args)
    out_again = fp8_scaled_mm(*args)
    smallm_cfg(0, -1)
    out_big = fp8_scaled_mm(*args)
    assert torch.equal(out, out_again)
    torch.testing.assert_close(out.cpu(), ref, rtol=0.02, atol=1)
    scale = max(1.0, ref.float().abs().max().item())
    assert (out.cpu().float() - ref.float()).abs().max().item() <= 2e-2 * scale
    # same fp8 products, fp32 accumulation in a different order: at most one bf16 ulp apart
    assert (out.float() - out_big.float()).abs().max().item() <= 2 ** -7 * scale


@pytest.mark.parametrize("m", [33, 48, 64, 65])
def test_fp8_scaled_mm_wide_weight_default_dispatch(m):
    """A weight with more 128-row tiles than SMs (here 150) takes the swap-AB kernel up to M = 64 and
    the big tiles above; both sides of the switch against the oracle with the library defaults."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    n, k = 150 * 128, 384
    g = torch.Generator().manual_seed(m + n)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    sa = torch.randn((m,), generator=g) * 0.001
    sb = torch.randn((n,), generator=g) * 0.001
    ref = O.scaled_mm_oracle(a, b.t(), sa, sb, torch.bfloat16, None)
    out = fp8_scaled_mm(a.cuda(), b.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, None)
    torch.testing.assert_close(out.cpu(), ref, rtol=0.02, atol=1)
    scale = max(1.0, ref.float().abs().max().item())
    assert (out.cpu().float() - ref.float()).abs().max().item() <= 2e-2 * scale


@pytest.mark.parametrize("m", [1, 16, 31, 64, 83, 128])
@pytest.mark.parametrize("n,k,splits", [(512, 7168, 0), (7168, 1024, 0), (1536, 2048, 8),
                                        (2112, 1152, 4), (128, 128, 1)])
def test_fp8_block_scaled_smallm(m, n, k, splits, smallm_cfg):
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_blockwise_scaled_mm

    g = torch.Generator().manual_seed(m + n + k + splits)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    a_s = torch.rand((m, (k + 127) // 128), generator=g) * 1e-2
    b_s = torch.rand(((n + 127) // 128, (k + 127) // 128), generator=g) * 1e-2
    ref = O.block_scaled_mm_oracle(a, b, a_s, b_s, (128, 128), torch.bfloat16)
    args = (a.cuda(), b.cuda(), a_s.cuda(), b_s.cuda(), torch.bfloat16)
    smallm_cfg(128, splits)
    out = fp8_blockwise_scaled_mm(*args)
    assert torch.equal(out, fp8_blockwise_scaled_mm(*args))
    smallm_cfg(0, -1)
    out_big = fp8_blockwise_scaled_mm(*args)
    rel = (out.cpu().float() - ref.float()).abs().mean() / ref.float().abs().mean()
    assert rel.item() < 1e-3, rel.item()
    rel2 = (out.float() - out_big.float()).abs().mean() / out_big.float().abs().mean()
    assert rel2.item() < 1e-3, rel2.item()
    a_s_col = a_s.cuda().t().contiguous().t()  # column-major activation scales
    smallm_cfg(128, splits)
    assert torch.equal(out, fp8_blockwise_scaled_mm(a.cuda(), b.cuda(), a_s_col, b_s.cuda(),
                                                    torch.bfloat16))
