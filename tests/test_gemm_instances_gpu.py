"""GPU parity of the LARGE-shape GEMM instances — the ones whose roofline fractions are quoted —
against the oracle functions run on the GPU in fp32 (TF32 off; the checker may use library matmuls,
the product may not).  Every case asserts WHICH kernel instance the dispatcher picked
(b200_last_variant) so a shape cannot silently fall back to a smaller tile.

  FP8 per-token x per-channel: 128x256 tiles, CTA pairs with TMA-multicast weights (even number of
      M blocks, bf16 and f16), the non-multicast 128x256 path (odd), 128x128 tiles with cluster
      split-K; all four Llama-3.1-8B shapes of BASELINE cfg 3 at M in {512, 2048, 8192}
  FP8 block-scaled: the 128x256-tile kernel (needs >= 1.5 tiles per SM)
  W4A16: the 256-row (two-accumulator) instance and the 128-row persistent instance on the
      Qwen2.5-32B TP=2 shapes of cfg 4
"""
import pytest
import torch

from oracle import fp8 as OF
from oracle import int4 as OI

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


@pytest.fixture(autouse=True)
def _exact_fp32_matmul():
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.set_float32_matmul_precision("highest")
    yield
    torch.backends.cuda.matmul.allow_tf32 = old


def _rand_fp8(shape, g):
    x = (torch.rand(shape, generator=g, device="cuda") - 0.5) * 2 * 448.0
    return x.clamp(-448.0, 448.0).to(FP8)


def _variant():
    from sglang_security_b200 import _lib

    return _lib.last_variant()


LLAMA8B = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]


@pytest.mark.parametrize("n,k", LLAMA8B)
@pytest.mark.parametrize("m", [512, 2048, 8192])
def test_fp8_rowcol_cfg3_shapes(m, n, k):
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    sa = torch.randn((m,), generator=g, device="cuda") * 0.001
    sb = torch.randn((n,), generator=g, device="cuda") * 0.001
    bias = torch.randn((n,), generator=g, device="cuda").to(torch.bfloat16)
    out = fp8_scaled_mm(a, b.t(), sa, sb, torch.bfloat16, bias)
    v = _variant()
    ref = OF.scaled_mm_oracle(a, b.t(), sa, sb, torch.bfloat16, bias)
    torch.testing.assert_close(out, ref, rtol=0.02, atol=1)
    scale = max(1.0, ref.float().abs().max().item())
    assert (out.float() - ref.float()).abs().max().item() <= 2e-2 * scale
    tiles256 = ((m + 127) // 128) * ((n + 255) // 256)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert v.startswith("gemm_tn kind=0"), v
    assert ("bn=256" in v) == (tiles256 >= sms), (v, tiles256)
    if tiles256 >= sms:
        assert "mc=1" in v, v          # even number of M blocks -> multicast pairs


@pytest.mark.parametrize("m,out_dtype,mc", [(2048, torch.bfloat16, 1), (2048, torch.float16, 1),
                                            (2176, torch.bfloat16, 0), (2176, torch.float16, 0),
                                            (4096 + 128, torch.bfloat16, 0)])
def test_fp8_rowcol_bn256_multicast_and_plain(m, out_dtype, mc):
    """fp8_gemm.cu tile dispatch: even ceil(M/128) -> CTA pairs sharing the weight tile by TMA
    multicast; odd -> the plain 128x256 persistent kernel.  Ragged N (not a multiple of 256)."""
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    n, k = 6144 + 80, 1152
    g = torch.Generator(device="cuda").manual_seed(m)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    sa = torch.rand((m,), generator=g, device="cuda") * 1e-3
    sb = torch.rand((n,), generator=g, device="cuda") * 1e-3
    out = fp8_scaled_mm(a, b.t(), sa, sb, out_dtype)
    v = _variant()
    assert "bn=256" in v and f"mc={mc}" in v and ("f16" if out_dtype == torch.float16 else "bf16") in v, v
    ref = OF.scaled_mm_oracle(a, b.t(), sa, sb, out_dtype, None)
    torch.testing.assert_close(out, ref, rtol=0.02, atol=1)
    scale = max(1.0, ref.float().abs().max().item())
    assert (out.float() - ref.float()).abs().max().item() <= 2e-2 * scale


def _block_oracle_gpu(a, b, a_s, b_s, out_dtype):
    """native_w8a8_block_fp8_matmul (test_block_fp8.py:276-325) with the n-tile loop vectorised:
    per K block, C += (A_kb @ B_kb^T) * (As[:, kb] x Bs[n // 128, kb]) in fp32."""
    m, k = a.shape
    n = b.shape[0]
    c = torch.zeros((m, n), dtype=torch.float32, device=a.device)
    for kb in range((k + 127) // 128):
        ks = slice(kb * 128, min(k, (kb + 1) * 128))
        s = a_s[:, kb:kb + 1].float() * b_s[:, kb].float().repeat_interleave(128)[:n].unsqueeze(0)
        c += (a[:, ks].float() @ b[:, ks].float().t()) * s
    return c.to(out_dtype)


@pytest.mark.parametrize("m,n,k", [(4096, 6144, 4096), (8192, 4096, 14336), (3000, 28672, 4096),
                                   (2048, 2112, 1152)])
def test_fp8_block_bn256_instance(m, n, k):
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_blockwise_scaled_mm

    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    a_s = torch.rand((m, (k + 127) // 128), generator=g, device="cuda") * 1e-2
    b_s = torch.rand(((n + 127) // 128, (k + 127) // 128), generator=g, device="cuda") * 1e-2
    out = fp8_blockwise_scaled_mm(a, b, a_s, b_s, torch.bfloat16)
    v = _variant()
    tiles256 = ((m + 127) // 128) * ((n + 255) // 256)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert v.startswith("fp8_block_gemm bn=256" if 2 * tiles256 >= 3 * sms else "fp8_block_gemm bn=128"), v
    if (m, n, k) != (2048, 2112, 1152):
        assert "bn=256" in v, v
    ref = _block_oracle_gpu(a, b, a_s, b_s, torch.bfloat16)
    rel = (out.float() - ref.float()).abs().mean() / ref.float().abs().mean()
    assert rel.item() < 1e-3, rel.item()
    # element-wise: fp32 accumulation of exact products in a different order + one bf16 rounding
    scale = ref.float().abs().max().item()
    assert (out.float() - ref.float()).abs().max().item() <= 1e-2 * scale


def test_block_oracle_gpu_form_equals_cpu_oracle():
    """The vectorised GPU form used above is the same arithmetic as oracle/fp8.py (which is pinned
    to the reference's function by tests/test_oracle_pinned.py)."""
    g = torch.Generator(device="cuda").manual_seed(1)
    m, n, k = 70, 384, 640
    a, b = _rand_fp8((m, k), g), _rand_fp8((n, k), g)
    a_s = torch.rand((m, k // 128), generator=g, device="cuda") * 1e-2
    b_s = torch.rand((n // 128, k // 128), generator=g, device="cuda") * 1e-2
    x = _block_oracle_gpu(a, b, a_s, b_s, torch.bfloat16).cpu()
    y = OF.block_scaled_mm_oracle(a.cpu(), b.cpu(), a_s.cpu(), b_s.cpu(), (128, 128), torch.bfloat16)
    rel = (x.float() - y.float()).abs().mean() / y.float().abs().mean()
    assert rel.item() < 1e-5, rel.item()


def _awq_realistic(k, n, group, dtype, g):
    """Asymmetric group quantisation of N(0, 0.02) weights (SURVEY §8d cfg 4 realistic variant)."""
    w = torch.randn((k, n), generator=g, device="cuda") * 0.02
    wg = w.view(k // group, group, n)
    lo, hi = wg.amin(1), wg.amax(1)
    scale = ((hi - lo).clamp(min=1e-5) / 15).to(dtype)
    zero = (-lo / scale.float()).round().clamp(0, 15).to(torch.int32)
    q = ((wg / scale.float().unsqueeze(1)).round() + zero.unsqueeze(1)).clamp(0, 15).to(torch.int32).view(k, n)
    pack = lambda t: sum(((t.view(t.shape[0], -1, 8)[..., j]) & 0xF) << (4 * OI.AWQ_ORDER[j])
                         for j in range(8)).to(torch.int32)
    return pack(q), scale, pack(zero)


@pytest.mark.parametrize("m,n,k,bm", [(1024, 3584, 5120, 256), (4096, 3584, 5120, 256),
                                      (1024, 27648, 5120, 256), (4096, 27648, 5120, 256),
                                      (512, 3584, 5120, 128), (2048, 5120, 13824, 256)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_w4a16_large_m_instances(m, n, k, bm, dtype):
    """Qwen2.5-32B TP=2 shapes (cfg 4) at prefill-chunk M: the 256-row two-accumulator instance
    (w4a16.cu w4_pick_bm) and the 128-row persistent one, through AWQLinearMethod.apply."""
    from sglang_security_b200.layers.linear import ColumnParallelLinear
    from sglang_security_b200.layers.quantization.awq import AWQConfig

    group = 128
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    qweight, scales, qzeros = _awq_realistic(k, n, group, dtype, g)
    x = torch.randn((m, k), generator=g, device="cuda").to(dtype)
    bias = torch.randn((n,), generator=g, device="cuda").to(dtype)
    layer = ColumnParallelLinear(k, n, bias=True, params_dtype=dtype,
                                 quant_config=AWQConfig(4, group, True), tp_rank=0, tp_size=1).cuda()
    layer.qweight.data.copy_(qweight)
    layer.qzeros.data.copy_(qzeros)
    layer.scales.data.copy_(scales)
    layer.bias.data.copy_(bias)
    layer.quant_method.process_weights_after_loading(layer)
    out, _ = layer(x)
    v = _variant()
    assert v.startswith(f"w4a16_gemm bm={bm} swap=0"), v
    w = OI.awq_dequantize_oracle(qweight, scales, qzeros, group)
    ref = OI.w4a16_linear_oracle(x, w, bias)
    err = (out.float() - ref.float()).abs()
    tol = 1e-2 * ref.float().abs() + 1e-2 * ref.float().abs().max()
    assert (err <= tol).all(), (err.max().item(), ref.float().abs().max().item())
    # regression: the packed ring used to be released to its TMA refill without a cross-proxy fence
    # (stale weights whenever the MMA pipe was the bottleneck, different on every run)
    out2, _ = layer(x)
    assert torch.equal(out, out2), "w4a16 GEMM is not deterministic run to run"
