"""INT4 weight-only kernels bound through the C-ABI (AWQ dequant, one-time repacks, fused GEMM)."""
from __future__ import annotations

from typing import Optional

import torch

from ... import _lib


def _cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise _lib.B200Error(f"{what}: tensor must live on a CUDA device (no CPU path)")


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """Same call as sgl_kernel.awq_dequantize (awq.py:210): -> [K, N] in scales.dtype."""
    _cuda(qweight, "awq_dequantize")
    k, nc = qweight.shape
    n = nc * 8
    group = k // scales.shape[0]
    out = torch.empty((k, n), device=qweight.device, dtype=scales.dtype)
    rc = _lib.load().b200_awq_dequant(qweight.contiguous().data_ptr(), scales.contiguous().data_ptr(),
                                      qzeros.contiguous().data_ptr(), out.data_ptr(), k, n, group,
                                      _lib.dtype_code(scales.dtype), _lib.current_stream())
    _lib.check(rc, "b200_awq_dequant")
    return out


def unpack_zeros(qzeros: torch.Tensor, dtype: torch.dtype, awq_order: bool, add_one: bool) -> torch.Tensor:
    _cuda(qzeros, "unpack_zeros")
    g, nc = qzeros.shape
    out = torch.empty((g, nc * 8), device=qzeros.device, dtype=dtype)
    rc = _lib.load().b200_w4a16_unpack_zeros(qzeros.contiguous().data_ptr(), out.data_ptr(), g,
                                             nc * 8, 1 if awq_order else 0, 1 if add_one else 0,
                                             _lib.dtype_code(dtype), _lib.current_stream())
    _lib.check(rc, "b200_w4a16_unpack_zeros")
    return out


def repack_gptq(qweight: torch.Tensor, perm: Optional[torch.Tensor] = None) -> torch.Tensor:
    """GPTQ [K/8, N] -> B200 [K, N/8] (AWQ nibble order); perm gathers source rows."""
    _cuda(qweight, "repack_gptq")
    k8, n = qweight.shape
    k = k8 * 8
    out = torch.empty((k, n // 8), device=qweight.device, dtype=torch.int32)
    if perm is not None:
        perm = perm.to(torch.int32).contiguous()
    rc = _lib.load().b200_w4a16_repack_gptq(qweight.contiguous().data_ptr(), _lib.ptr(perm),
                                            out.data_ptr(), k, n, _lib.current_stream())
    _lib.check(rc, "b200_w4a16_repack_gptq")
    return out


def w4a16_gemm(x: torch.Tensor, qweight: torch.Tensor, zeros: torch.Tensor, scales: torch.Tensor,
               group_size: int, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [M, K] f16/bf16 @ dequant(qweight [K, N/8], zeros/scales [K/g, N]) (+ bias) -> [M, N]."""
    _cuda(x, "w4a16_gemm")
    if x.dim() != 2 or x.stride(-1) != 1:
        raise _lib.B200Error("w4a16_gemm: x must be 2-D with contiguous rows")
    m, k = x.shape
    n = qweight.shape[1] * 8
    if qweight.shape[0] != k or scales.dtype != x.dtype or zeros.dtype != x.dtype:
        raise _lib.B200Error("w4a16_gemm: shape / dtype mismatch between x and the packed weight")
    out = torch.empty((m, n), device=x.device, dtype=x.dtype)
    if bias is not None and bias.dtype != x.dtype:
        bias = bias.to(x.dtype)
    rc = _lib.load().b200_w4a16_gemm(x.data_ptr(), x.stride(0), qweight.data_ptr(), zeros.data_ptr(),
                                     scales.data_ptr(), _lib.ptr(bias), out.data_ptr(),
                                     out.stride(0), m, n, k, group_size, _lib.dtype_code(x.dtype),
                                     _lib.current_stream())
    _lib.check(rc, "b200_w4a16_gemm")
    return out


def w4a16_gemv_config(max_m: int = -1, splits: int = -1) -> None:
    """Test/sweep hook (b200_w4a16_gemv_config): max_m = 0 keeps b200_w4a16_gemm on the tcgen05
    kernel for every M, splits > 0 fixes the GEMV's split-K cluster size, negative = defaults."""
    _lib.check(_lib.load().b200_w4a16_gemv_config(int(max_m), int(splits)), "b200_w4a16_gemv_config")
