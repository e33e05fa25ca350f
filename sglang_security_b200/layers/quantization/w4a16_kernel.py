"""INT4 weight-only kernels bound through the C-ABI (AWQ dequant, one-time repacks, fused GEMM)."""
from __future__ import annotations

import os
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


def gemv_stream_repack(qweight: torch.Tensor) -> Optional[torch.Tensor]:
    """One-time repack of the B200 [K, N/8] packed weight into the decode GEMV's stream layout
    (csrc/w4a16_gemv2.cu): same bytes, ordered so that a warp's 512-byte load is its mma.sync A
    fragments.  Returns None when the shape has no stream layout (K % 64 or N % 16 != 0)."""
    _cuda(qweight, "gemv_stream_repack")
    k, nc = qweight.shape
    n = nc * 8
    if k % 64 != 0 or n % 16 != 0:
        return None
    out = torch.empty(k * nc, device=qweight.device, dtype=torch.int32)
    rc = _lib.load().b200_w4a16_gemv_v2_repack(qweight.contiguous().data_ptr(), out.data_ptr(), k, n,
                                               _lib.current_stream())
    _lib.check(rc, "b200_w4a16_gemv_v2_repack")
    return out


def gemv_v2_supported(m: int, n: int, k: int, group_size: int) -> bool:
    return bool(_lib.load().b200_w4a16_gemv_v2_supported(int(m), int(n), int(k), int(group_size)))


def w4a16_gemv_v2(x: torch.Tensor, qweight_stream: torch.Tensor, zeros: torch.Tensor,
                  scales: torch.Tensor, group_size: int,
                  bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [M <= 16, K] @ dequant(stream-layout weight) (+ bias) -> [M, N]; N from the scales."""
    _cuda(x, "w4a16_gemv_v2")
    if x.dim() != 2 or x.stride(-1) != 1:
        raise _lib.B200Error("w4a16_gemv_v2: x must be 2-D with contiguous rows")
    m, k = x.shape
    n = scales.shape[1]
    if qweight_stream.numel() * 8 != k * n or scales.dtype != x.dtype or zeros.dtype != x.dtype:
        raise _lib.B200Error("w4a16_gemv_v2: shape / dtype mismatch between x and the packed weight")
    out = torch.empty((m, n), device=x.device, dtype=x.dtype)
    if bias is not None and bias.dtype != x.dtype:
        bias = bias.to(x.dtype)
    rc = _lib.load().b200_w4a16_gemv_v2(x.data_ptr(), x.stride(0), qweight_stream.data_ptr(),
                                        zeros.data_ptr(), scales.data_ptr(), _lib.ptr(bias),
                                        out.data_ptr(), out.stride(0), m, n, k, group_size,
                                        _lib.dtype_code(x.dtype), _lib.current_stream())
    _lib.check(rc, "b200_w4a16_gemv_v2")
    return out


def w4a16_gemv_v2_config(splits: int = 0) -> None:
    _lib.check(_lib.load().b200_w4a16_gemv_v2_config(int(splits)), "b200_w4a16_gemv_v2_config")


# largest token count handed to the stream GEMV (its two-tile instance for 9..16 tokens is at par or
# ahead of the tcgen05 swap-AB kernel on every cfg-4 shape, profiles/r02_bench_w4a16_gemv2.json)
GEMV_V2_MAX_M = int(os.environ.get("B200_W4A16_GEMV2_MAX_M", "16"))


def stream_copy_enabled() -> bool:
    return os.environ.get("B200_W4A16_STREAM_COPY", "1") != "0"


def w4a16_linear(x2: torch.Tensor, layer, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """Dispatch of AWQ/GPTQ apply(): decode batches (M <= 16) stream the repacked weight through
    the GEMV, everything else runs the fused dequant + tcgen05 GEMM on the [K, N/8] layout."""
    stream = getattr(layer, "b200_qweight_stream", None)
    m = x2.shape[0]
    if stream is not None and 1 <= m <= GEMV_V2_MAX_M and gemv_v2_supported(
            m, layer.scales.shape[1], x2.shape[1], layer.w4a16_group_size):
        return w4a16_gemv_v2(x2, stream, layer.b200_zeros, layer.scales, layer.w4a16_group_size, bias)
    packed = layer.b200_qweight if hasattr(layer, "b200_qweight") else layer.qweight  # GPTQ / AWQ
    return w4a16_gemm(x2, packed, layer.b200_zeros, layer.scales, layer.w4a16_group_size, bias)


def w4a16_gemv_config(max_m: int = -1, splits: int = -1) -> None:
    """Test/sweep hook (b200_w4a16_gemv_config): max_m = 0 keeps b200_w4a16_gemm on the tcgen05
    kernel for every M, splits > 0 fixes the GEMV's split-K cluster size, negative = defaults."""
    _lib.check(_lib.load().b200_w4a16_gemv_config(int(max_m), int(splits)), "b200_w4a16_gemv_config")
