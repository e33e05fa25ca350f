"""Dense 16-bit GEMM bound through the C-ABI (csrc/dense_gemm.cu): the unquantised layers and the
lm_head run on the in-tree tcgen05 kernel instead of a library call."""
from __future__ import annotations

from typing import Optional

import torch

from .. import _lib


def dense_supported(x: torch.Tensor, weight: torch.Tensor) -> bool:
    """CUDA, bf16/f16, K-contiguous operands that satisfy the TMA alignment rules."""
    return (x.is_cuda and weight.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and
            weight.dtype == x.dtype and weight.dim() == 2 and weight.stride(1) == 1 and
            weight.stride(0) % 8 == 0 and weight.data_ptr() % 16 == 0 and x.shape[-1] % 8 == 0)


def gemm_16bit(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [..., K] @ weight[N, K]^T (+ bias) -> [..., N] in x.dtype (same contract as F.linear)."""
    if not x.is_cuda:
        raise _lib.B200Error("gemm_16bit: tensors must live on a CUDA device (no CPU path)")
    k = x.shape[-1]
    n = weight.shape[0]
    x2 = x.reshape(-1, k)
    if x2.stride(-1) != 1 or x2.stride(0) % 8 != 0 or x2.data_ptr() % 16 != 0:
        x2 = x2.contiguous()
    out = torch.empty((x2.shape[0], n), device=x.device, dtype=x.dtype)
    if bias is not None and bias.dtype != x.dtype:
        bias = bias.to(x.dtype)
    rc = _lib.load().b200_gemm_16bit(x2.data_ptr(), x2.stride(0), weight.data_ptr(), weight.stride(0),
                                     _lib.ptr(bias), out.data_ptr(), out.stride(0), x2.shape[0], n, k,
                                     _lib.dtype_code(x.dtype), _lib.current_stream())
    _lib.check(rc, "b200_gemm_16bit")
    return out.reshape(x.shape[:-1] + (n,))


def argmax_rows(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """torch.argmax(logits, -1) for CUDA bf16/f16 [M, V] logits (first index of the maximum) -> int64 [M]."""
    if not logits.is_cuda or logits.dim() != 2 or logits.stride(-1) != 1:
        raise _lib.B200Error("argmax_rows: logits must be a 2-D CUDA tensor with contiguous rows")
    m, v = logits.shape
    if out is None:
        out = torch.empty((m,), dtype=torch.int64, device=logits.device)
    rc = _lib.load().b200_argmax_rows(logits.data_ptr(), logits.stride(0), out.data_ptr(), m, v,
                                      _lib.dtype_code(logits.dtype), _lib.current_stream())
    _lib.check(rc, "b200_argmax_rows")
    return out
