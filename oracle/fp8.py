"""CPU restatements of the reference's FP8 quantisers and scaled GEMMs (torch, fp32 math).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Each function names the reference code it
follows; the formulas are the torch-native references the reference's own tests compare its CUDA
kernels against.
"""
from __future__ import annotations

import torch

FP8 = torch.float8_e4m3fn
FP8_MAX = torch.finfo(FP8).max  # 448


def per_token_quant_fp8_oracle(x: torch.Tensor):
    """Dynamic per-token quantisation: scale[m] = absmax(x[m]) / 448, q = clamp(x * (1/scale)).
    Follows sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:9-76 and its torch twin
    sgl-kernel/tests/test_per_token_quant_fp8.py:14-22 (reciprocal, then multiply)."""
    xf = x.float()
    scale = xf.abs().amax(dim=-1, keepdim=True) / FP8_MAX
    inv = torch.where(scale > 0, scale.reciprocal(), torch.zeros_like(scale))
    q = (xf * inv).clamp(-FP8_MAX, FP8_MAX).to(FP8)
    return q, scale


def static_quant_fp8_oracle(x: torch.Tensor, scale: torch.Tensor):
    """Static per-tensor quantisation with a given scale (fp8_kernel.py:336-376,
    native_static_quant_fp8 python/sglang/test/test_block_fp8.py:97-115): q = clamp(x * (1/s))."""
    inv = scale.float().reciprocal()
    return (x.float() * inv).clamp(-FP8_MAX, FP8_MAX).to(FP8)


def per_token_group_quant_fp8_oracle(x: torch.Tensor, group_size: int = 128, eps: float = 1e-10):
    """Per-(token, group) quantisation, true division by the scale
    (native_per_token_group_quant_fp8, python/sglang/test/test_block_fp8.py:23-48).  The eps clamp is
    applied to the fp32 absmax as the reference's kernels do (fp8_kernel.py:133,
    per_token_group_quant_8bit.cu:40); the in-test twin clamps in the input dtype, which only
    differs for all-zero groups (tests/test_oracle_pinned.py spells that out)."""
    assert x.shape[-1] % group_size == 0
    g = x.float().reshape(-1, group_size)
    amax = g.abs().amax(dim=-1, keepdim=True).clamp(min=eps)
    s = amax / FP8_MAX
    q = (g / s).clamp(-FP8_MAX, FP8_MAX).to(FP8).reshape(x.shape)
    return q, s.reshape(x.shape[:-1] + (x.shape[-1] // group_size,))


def scaled_mm_oracle(a, b_t, scale_a, scale_b, out_dtype, bias=None):
    """out = ((A @ B) * sa[:,None] * sb[None,:]).to(out_dtype) (+ bias)   with A [M,K] fp8,
    B given as the [K,N] (column-major) view the reference passes
    (torch_scaled_mm, sgl-kernel/tests/test_fp8_gemm.py:6-14)."""
    o = a.float() @ b_t.float()
    o = o * scale_a.float().reshape(-1, 1)
    o = o * scale_b.float().reshape(1, -1)
    o = o.to(out_dtype)
    if bias is not None:
        o = o + bias.reshape(1, -1)
    return o


def block_scaled_mm_oracle(a, b, a_s, b_s, block_size=(128, 128), out_dtype=torch.bfloat16):
    """C = sum_kb (A[:,kb] @ B[nb,kb]^T) * As[:,kb] * Bs[nb,kb]: A [M,K] fp8 with [M,K/bk] scales,
    B [N,K] fp8 with [N/bn, K/bk] scales (native_w8a8_block_fp8_matmul,
    python/sglang/test/test_block_fp8.py:276-325)."""
    bn, bk = block_size
    af, bf = a.float(), b.float()
    m, k = af.shape
    n = bf.shape[0]
    c = torch.zeros((m, n), dtype=torch.float32, device=af.device)
    for kb in range((k + bk - 1) // bk):
        ks = slice(kb * bk, min((kb + 1) * bk, k))
        for nb in range((n + bn - 1) // bn):
            ns = slice(nb * bn, min((nb + 1) * bn, n))
            c[:, ns] += (af[:, ks] @ bf[ns, ks].t()) * (a_s[:, kb:kb + 1].float() * b_s[nb, kb].float())
    return c.to(out_dtype)


def fp8_linear_oracle(x, weight_fp8_nk, weight_scale, bias=None, input_scale=None):
    """apply_fp8_linear on the per-channel path (fp8_utils.py:454-648): dynamic per-token (or
    static) activation quant, scaled GEMM, output in x.dtype."""
    x2 = x.reshape(-1, x.shape[-1])
    if input_scale is not None:
        q = static_quant_fp8_oracle(x2, input_scale)
        sx = input_scale.float().reshape(1).expand(x2.shape[0])
    else:
        q, sx = per_token_quant_fp8_oracle(x2)
    out = scaled_mm_oracle(q, weight_fp8_nk.t(), sx.reshape(-1), weight_scale.reshape(-1), x.dtype,
                           bias)
    return out.reshape(*x.shape[:-1], weight_fp8_nk.shape[0])


def block_fp8_linear_oracle(x, weight_fp8_nk, weight_scale_inv, block_size=(128, 128), bias=None):
    """Block-quant linear (fp8_utils.py:152-304): per-token-group(128) activation quant +
    block-scaled GEMM (+ bias), output in x.dtype."""
    x2 = x.reshape(-1, x.shape[-1])
    q, sx = per_token_group_quant_fp8_oracle(x2, block_size[1])
    out = block_scaled_mm_oracle(q, weight_fp8_nk, sx, weight_scale_inv, block_size, x.dtype)
    if bias is not None:
        out = out + bias
    return out.to(x.dtype).reshape(*x.shape[:-1], weight_fp8_nk.shape[0])
