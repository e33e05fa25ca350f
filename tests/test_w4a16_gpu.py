"""GPU parity of the INT4 weight-only path: bit-exact unpack / dequant / repack and the fused
tcgen05 W4A16 GEMM behind AWQLinearMethod / GPTQLinearMethod."""
import numpy as np
import pytest
import torch

from oracle import int4 as O

pytestmark = pytest.mark.gpu
I32_MAX = torch.iinfo(torch.int32).max


def _rand_i32(shape, g):
    # random bit patterns incl. the sign bit (the reference test only draws non-negative words)
    return torch.from_numpy(np.random.RandomState(int(g)).randint(-2**31, 2**31, size=shape,
                                                                  dtype=np.int64).astype(np.int32))


@pytest.mark.parametrize("rows,cols", [(128, 16), (256, 64), (1024, 128), (3584, 448), (512, 576)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("group", [None, 128])
def test_awq_dequant_bit_exact(rows, cols, dtype, group):
    """Shapes / distributions of sgl-kernel/tests/test_awq_dequant.py:67-112 (its tolerance is
    rtol 1e-3 / atol 1e-5; here the result must be bit-identical to the torch reference)."""
    from sglang_security_b200.layers.quantization.w4a16_kernel import awq_dequantize

    g = torch.Generator().manual_seed(rows + cols)
    group_size = group or rows
    qweight = _rand_i32((rows, cols), rows * 3 + cols)
    qzeros = _rand_i32((rows // group_size, cols), rows * 5 + cols)
    scales = torch.rand((rows // group_size, cols * 8), generator=g).to(dtype)
    ref = O.awq_dequantize_oracle(qweight, scales, qzeros, group_size)
    out = awq_dequantize(qweight.cuda(), scales.cuda(), qzeros.cuda())
    assert out.dtype == dtype and out.shape == (rows, cols * 8)
    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


def test_gptq_repack_and_zero_unpack_bit_exact():
    from sglang_security_b200.layers.quantization.w4a16_kernel import repack_gptq, unpack_zeros

    k, n, group = 512, 384, 128
    qweight = _rand_i32((k // 8, n), 1)
    qzeros = _rand_i32((k // group, n // 8), 2)
    ints = O.unpack_gptq_qweight(qweight)                       # [K, N] integers
    packed = repack_gptq(qweight.cuda()).cpu()                  # [K, N/8] AWQ order
    assert torch.equal(O.unpack_awq_int4(packed), ints)
    perm = torch.from_numpy(np.random.RandomState(3).permutation(k))
    packed_p = repack_gptq(qweight.cuda(), perm.cuda()).cpu()
    assert torch.equal(O.unpack_awq_int4(packed_p), ints[perm])
    for dtype, magic in ((torch.float16, 1024), (torch.bfloat16, 128)):
        z = unpack_zeros(qzeros.cuda(), dtype, awq_order=False, add_one=True).cpu()
        assert torch.equal(z.float(), (O.unpack_gptq_qzeros(qzeros) + magic).float())
        z = unpack_zeros(qzeros.cuda(), dtype, awq_order=True, add_one=False).cpu()
        assert torch.equal(z.float(), (O.unpack_awq_int4(qzeros) + magic).float())


def _awq_layer(k, n, group, dtype, seed, realistic=False):
    g = torch.Generator().manual_seed(seed)
    if realistic:
        # asymmetric group quantisation of N(0, 0.02) weights (SURVEY §8d cfg 4 "realistic variant")
        w = torch.randn((k, n), generator=g) * 0.02
        wg = w.view(k // group, group, n)
        lo, hi = wg.amin(1), wg.amax(1)
        scale = ((hi - lo).clamp(min=1e-5) / 15).to(dtype)
        zero = (-lo / scale.float()).round().clamp(0, 15).to(torch.int32)
        q = ((wg / scale.float().unsqueeze(1)).round() + zero.unsqueeze(1)).clamp(0, 15).to(torch.int32).view(k, n)
        pack = lambda t: sum(((t.view(t.shape[0], -1, 8)[..., j]) & 0xF) << (4 * O.AWQ_ORDER[j])
                             for j in range(8)).to(torch.int32)
        return pack(q), scale, pack(zero)
    return (_rand_i32((k, n // 8), seed), (torch.rand((k // group, n), generator=g) * 0.02).to(dtype),
            _rand_i32((k // group, n // 8), seed + 1))


@pytest.mark.parametrize("m", [1, 16, 24, 64, 100, 130, 256, 1024])
@pytest.mark.parametrize("k,n,group", [(512, 256, 128), (5120, 3584, 128), (2560, 5120, 128),
                                       (256, 136, 32), (1024, 1000, 64)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_awq_linear_method(m, k, n, group, dtype):
    from sglang_security_b200.layers.linear import ColumnParallelLinear
    from sglang_security_b200.layers.quantization.awq import AWQConfig

    big = m * k * n > 2 ** 32   # too slow on the host: same oracle functions on the GPU in fp32
    qweight, scales, qzeros = _awq_layer(k, n, group, dtype, seed=m + k + n, realistic=(n % 128 == 0))
    g = torch.Generator().manual_seed(7)
    x = torch.randn((m, k), generator=g).to(dtype)
    bias = torch.randn((n,), generator=g).to(dtype)
    layer = ColumnParallelLinear(k, n, bias=True, params_dtype=dtype,
                                 quant_config=AWQConfig(4, group, True), tp_rank=0, tp_size=1).cuda()
    layer.qweight.data.copy_(qweight)
    layer.qzeros.data.copy_(qzeros)
    layer.scales.data.copy_(scales)
    layer.bias.data.copy_(bias)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x.cuda())
    if big:
        torch.backends.cuda.matmul.allow_tf32 = False
        ref = O.w4a16_linear_oracle(x.cuda(), O.awq_dequantize_oracle(
            qweight.cuda(), scales.cuda(), qzeros.cuda(), group), bias.cuda()).cpu()
    else:
        ref = O.w4a16_linear_oracle(x, O.awq_dequantize_oracle(qweight, scales, qzeros, group), bias)
    # reference tolerance for x @ dequant in 16 bit: rtol 1e-2 (SURVEY §8d cfg 4)
    err = (out.cpu().float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())


@pytest.mark.parametrize("m", [130, 256])
def test_awq_deep_k_swap256(m):
    """K >= 8192 with few column tiles takes the 256-token swap-AB / split-K instance."""
    from sglang_security_b200.layers.quantization.w4a16_kernel import unpack_zeros, w4a16_gemm

    k, n, group, dtype = 8192, 384, 128, torch.bfloat16
    qweight, scales, qzeros = _awq_layer(k, n, group, dtype, seed=3, realistic=True)
    x = torch.randn((m, k), generator=torch.Generator().manual_seed(5)).to(dtype)
    z = unpack_zeros(qzeros.cuda(), dtype, True, False)
    out = w4a16_gemm(x.cuda(), qweight.cuda(), z, scales.cuda(), group)
    ref = O.w4a16_linear_oracle(x, O.awq_dequantize_oracle(qweight, scales, qzeros, group), None)
    err = (out.cpu().float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())


@pytest.mark.parametrize("desc_act", [False, True])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_gptq_linear_method(desc_act, dtype):
    from sglang_security_b200.layers.linear import RowParallelLinear
    from sglang_security_b200.layers.quantization.gptq import GPTQConfig

    k, n, group, m = 1024, 768, 128, 45
    g = torch.Generator().manual_seed(11)
    qweight = _rand_i32((k // 8, n), 21)
    qzeros = _rand_i32((k // group, n // 8), 22)
    scales = (torch.rand((k // group, n), generator=g) * 0.02).to(dtype)
    g_idx = torch.arange(k, dtype=torch.int32) // group
    if desc_act:
        g_idx = g_idx[torch.from_numpy(np.random.RandomState(5).permutation(k))]
    x = torch.randn((m, k), generator=g).to(dtype)
    layer = RowParallelLinear(k, n, bias=False, params_dtype=dtype,
                              quant_config=GPTQConfig(4, group, desc_act), tp_rank=0, tp_size=1).cuda()
    layer.qweight.data.copy_(qweight)
    layer.qzeros.data.copy_(qzeros)
    layer.scales.data.copy_(scales)
    layer.g_idx.data.copy_(g_idx)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x.cuda())
    ref = O.w4a16_linear_oracle(x, O.gptq_dequantize_oracle(qweight, scales, qzeros, g_idx, group))
    err = (out.cpu().float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())


@pytest.fixture
def gemv_cfg():
    from sglang_security_b200.layers.quantization.w4a16_kernel import w4a16_gemv_config

    yield w4a16_gemv_config
    w4a16_gemv_config(-1, -1)


def _prep(qweight, scales, qzeros, dtype):
    from sglang_security_b200.layers.quantization.w4a16_kernel import unpack_zeros

    return (qweight.cuda().contiguous(), unpack_zeros(qzeros.cuda(), dtype, True, False),
            scales.cuda().contiguous())


@pytest.mark.parametrize("m", [1, 2, 7, 8, 9, 16])
@pytest.mark.parametrize("k,n,group,splits", [(512, 256, 128, 0), (5120, 3584, 128, 0),
                                              (2560, 5120, 128, 3), (1024, 288, 64, 8),
                                              (256, 32, 32, 2), (13824, 5120, 128, 0),
                                              (640, 2048, 128, 8)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_w4a16_gemv_matches_oracle_and_tcgen05_kernel(m, k, n, group, splits, dtype, gemv_cfg):
    """Decode-regime GEMV (integer MMA + per-group zero/scale correction, cluster split-K incl.
    empty ranks and partial 256-column tiles) against x @ dequant (oracle) and against the tcgen05
    dequant-GEMM on identical operands; two runs bit-identical (fixed reduction order)."""
    from sglang_security_b200.layers.quantization.w4a16_kernel import w4a16_gemm

    qweight, scales, qzeros = _awq_layer(k, n, group, dtype, seed=m * 3 + k + n, realistic=(n % 128 == 0))
    g = torch.Generator().manual_seed(m + n)
    x = torch.randn((m, k), generator=g).to(dtype)
    bias = torch.randn((n,), generator=g).to(dtype)
    ref = O.w4a16_linear_oracle(x, O.awq_dequantize_oracle(qweight, scales, qzeros, group), bias)
    qw, zb, sc = _prep(qweight, scales, qzeros, dtype)
    gemv_cfg(16, splits)
    out = w4a16_gemm(x.cuda(), qw, zb, sc, group, bias.cuda())
    assert torch.equal(out, w4a16_gemm(x.cuda(), qw, zb, sc, group, bias.cuda()))
    gemv_cfg(0, -1)
    out_big = w4a16_gemm(x.cuda(), qw, zb, sc, group, bias.cuda())
    err = (out.cpu().float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())
    # the GEMV skips the per-weight 16-bit rounding of the dequantised value: it is at least as
    # close to the exact fp64 result as the dequant-GEMM
    w_exact = (O.unpack_awq_int4(qweight).double() -
               O.unpack_awq_int4(qzeros).double().repeat_interleave(group, 0)) * \
        scales.double().repeat_interleave(group, 0)
    exact = x.double() @ w_exact + bias.double()
    e_gemv = (out.cpu().double() - exact).abs().mean().item()
    e_big = (out_big.cpu().double() - exact).abs().mean().item()
    assert e_gemv <= 1.05 * e_big + 1e-6, (e_gemv, e_big)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_w4a16_gemv_integer_exact(dtype, gemv_cfg):
    """INT4 unpacking through the GEMV is exact: unit scales and small-integer activations make
    every partial sum an integer below 2^24 (fp32-exact) and the result fit the 16-bit output."""
    from sglang_security_b200.layers.quantization.w4a16_kernel import w4a16_gemm

    k, n, group, m = 256, 512, 64, 5
    qweight = _rand_i32((k, n // 8), 31)
    qzeros = _rand_i32((k // group, n // 8), 32)
    scales = torch.ones((k // group, n), dtype=dtype)
    g = torch.Generator().manual_seed(1)
    x = torch.randint(-1, 2, (m, k), generator=g).to(dtype)   # -1, 0, 1
    w = O.unpack_awq_int4(qweight).long() - O.unpack_awq_int4(qzeros).long().repeat_interleave(group, 0)
    exact = x.long() @ w                                      # |.| <= 256 * 15
    qw, zb, sc = _prep(qweight, scales, qzeros, dtype)
    gemv_cfg(16, 2)
    out = w4a16_gemm(x.cuda(), qw, zb, sc, group)
    assert torch.equal(out.cpu(), exact.to(dtype))


# ---- decode GEMV, second generation (csrc/w4a16_gemv2.cu): stream layout + mma.sync ---------------
@pytest.mark.parametrize("k,n", [(64, 16), (256, 272), (1024, 512), (5120, 3584)])
def test_gemv_stream_repack_bit_exact(k, n):
    from sglang_security_b200.layers.quantization.w4a16_kernel import gemv_stream_repack

    qweight = _rand_i32((k, n // 8), k + n)
    got = gemv_stream_repack(qweight.cuda()).cpu()
    assert torch.equal(got, O.gemv_stream_layout_oracle(qweight))


def test_gemv_stream_repack_unsupported_shapes():
    from sglang_security_b200.layers.quantization.w4a16_kernel import gemv_stream_repack, gemv_v2_supported

    assert gemv_stream_repack(_rand_i32((96, 4), 1).cuda()) is None      # K % 64
    assert gemv_stream_repack(_rand_i32((64, 3), 1).cuda()) is None      # N % 16
    assert not gemv_v2_supported(17, 256, 512, 128)
    assert not gemv_v2_supported(4, 256, 512, 96)
    assert gemv_v2_supported(16, 27648, 5120, 128) and gemv_v2_supported(1, 5120, 13824, 128)


@pytest.fixture
def gemv2_cfg():
    from sglang_security_b200.layers.quantization.w4a16_kernel import w4a16_gemv_v2_config

    yield w4a16_gemv_v2_config
    w4a16_gemv_v2_config(0)


@pytest.mark.parametrize("m", [1, 2, 7, 8, 9, 16])
@pytest.mark.parametrize("k,n,group,splits", [(512, 256, 128, 0), (5120, 3584, 128, 0),
                                              (2560, 5120, 128, 0), (1024, 288, 64, 8),
                                              (256, 32, 32, 2), (13824, 5120, 128, 0),
                                              (640, 2048, 128, 4), (5120, 27648, 128, 0),
                                              (1024, 272, 1024, 1)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_w4a16_gemv_v2_matches_oracle(m, k, n, group, splits, dtype, gemv2_cfg):
    """Stream-layout GEMV against x @ dequant (oracle) at the reference tolerance, at least as
    close to the exact fp64 result as the dequant-GEMM, and run-to-run bit-identical (fixed
    reduction order); covers empty cluster ranks, partial 128-column groups, groups of 32 / 64 /
    128 / K, both token-tile counts and the cfg-4 TP = 2 shapes."""
    from sglang_security_b200._lib import load
    from sglang_security_b200.layers.quantization.w4a16_kernel import (gemv_stream_repack, w4a16_gemm,
                                                                        w4a16_gemv_v2)

    qweight, scales, qzeros = _awq_layer(k, n, group, dtype, seed=m * 3 + k + n, realistic=(n % 128 == 0))
    g = torch.Generator().manual_seed(m + n)
    x = torch.randn((m, k), generator=g).to(dtype)
    bias = torch.randn((n,), generator=g).to(dtype)
    big = m * k * n > 2 ** 30
    dev = "cuda" if big else "cpu"
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = O.w4a16_linear_oracle(x.to(dev), O.awq_dequantize_oracle(
        qweight.to(dev), scales.to(dev), qzeros.to(dev), group), bias.to(dev)).cpu()
    qw, zb, sc = _prep(qweight, scales, qzeros, dtype)
    stream = gemv_stream_repack(qw)
    gemv2_cfg(splits)
    out = w4a16_gemv_v2(x.cuda(), stream, zb, sc, group, bias.cuda())
    assert b"w4a16_gemv2" in load().b200_last_variant()
    assert torch.equal(out, w4a16_gemv_v2(x.cuda(), stream, zb, sc, group, bias.cuda()))
    err = (out.cpu().float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())
    if not big:
        out_big = w4a16_gemm(x.cuda(), qw, zb, sc, group, bias.cuda())
        w_exact = (O.unpack_awq_int4(qweight).double() -
                   O.unpack_awq_int4(qzeros).double().repeat_interleave(group, 0)) * \
            scales.double().repeat_interleave(group, 0)
        exact = x.double() @ w_exact + bias.double()
        e_gemv = (out.cpu().double() - exact).abs().mean().item()
        e_big = (out_big.cpu().double() - exact).abs().mean().item()
        assert e_gemv <= 1.05 * e_big + 1e-6, (e_gemv, e_big)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m", [5, 13])
def test_w4a16_gemv_v2_integer_exact(dtype, m, gemv2_cfg):
    """INT4 unpacking through the stream layout is exact: unit scales and activations in
    {-1, 0, 1} make every partial sum an integer below 2^24 and the result fit the 16-bit output
    (f16: this also proves the 1/16 pre-scaling of the shift-free nibble pairs is exact)."""
    from sglang_security_b200.layers.quantization.w4a16_kernel import gemv_stream_repack, w4a16_gemv_v2

    k, n, group = 256, 512, 64
    qweight = _rand_i32((k, n // 8), 31)
    qzeros = _rand_i32((k // group, n // 8), 32)
    scales = torch.ones((k // group, n), dtype=dtype)
    g = torch.Generator().manual_seed(1)
    x = torch.randint(-1, 2, (m, k), generator=g).to(dtype)
    w = O.unpack_awq_int4(qweight).long() - O.unpack_awq_int4(qzeros).long().repeat_interleave(group, 0)
    exact = x.long() @ w
    qw, zb, sc = _prep(qweight, scales, qzeros, dtype)
    gemv2_cfg(2)
    out = w4a16_gemv_v2(x.cuda(), gemv_stream_repack(qw), zb, sc, group)
    assert torch.equal(out.cpu(), exact.to(dtype))


def test_awq_method_decode_batches_take_the_stream_gemv():
    from sglang_security_b200._lib import load
    from sglang_security_b200.layers.linear import ColumnParallelLinear
    from sglang_security_b200.layers.quantization.awq import AWQConfig

    k, n, group, dtype = 512, 256, 128, torch.float16
    qweight, scales, qzeros = _awq_layer(k, n, group, dtype, seed=3, realistic=True)
    layer = ColumnParallelLinear(k, n, bias=False, params_dtype=dtype,
                                 quant_config=AWQConfig(4, group, True), tp_rank=0, tp_size=1).cuda()
    layer.qweight.data.copy_(qweight)
    layer.qzeros.data.copy_(qzeros)
    layer.scales.data.copy_(scales)
    layer.quant_method.process_weights_after_loading(layer)
    assert layer.b200_qweight_stream is not None
    x = torch.randn((4, k), dtype=dtype, device="cuda")
    layer(x)
    assert b"w4a16_gemv2" in load().b200_last_variant()
    layer(torch.randn((40, k), dtype=dtype, device="cuda"))
    assert b"w4a16_gemv2" not in load().b200_last_variant()
