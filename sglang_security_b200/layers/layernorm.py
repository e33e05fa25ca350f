"""RMSNorm with the reference's interface (python/sglang/srt/layers/layernorm.py:42-108): same
constructor, `forward(x, residual=None)` returning `x` or `(x, residual)` with the residual updated
in place like the reference's fused_add_rmsnorm path (:69-71).  One sm_100a kernel through the
C-ABI (b200_rmsnorm); no torch/eager fallback.

`forward_quant_fp8` is the fused variant for a following FP8 linear layer: it returns the quantised
activation and its scales (what Fp8LinearMethod.apply would compute from the 16-bit output with
sgl_per_token_quant_fp8 / per_token_group_quant_fp8, fp8_utils.py:454-648) without materialising
the 16-bit tensor; the bytes are identical to the unfused chain."""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch
import torch.nn as nn

from .. import _lib

FP8 = torch.float8_e4m3fn


def _rows(x: torch.Tensor) -> torch.Tensor:
    if not x.is_cuda:
        raise _lib.B200Error("RMSNorm: tensor must live on a CUDA device (no CPU path)")
    x2 = x.reshape(-1, x.shape[-1])
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    return x2


def rmsnorm_call(x2, residual2, weight, eps, out2, q2, scales, quant_mode, quant_eps=1e-10):
    """Raw wrapper over b200_rmsnorm on 2-D row tensors."""
    s_row, s_col = (scales.stride(0), scales.stride(1)) if (scales is not None and scales.dim() == 2) \
        else (1, 0)
    rc = _lib.load().b200_rmsnorm(
        x2.data_ptr(), x2.stride(0), _lib.ptr(residual2),
        residual2.stride(0) if residual2 is not None else 0, weight.data_ptr(),
        _lib.ptr(out2), out2.stride(0) if out2 is not None else 0,
        _lib.ptr(q2), q2.stride(0) if q2 is not None else 0, _lib.ptr(scales), s_row, s_col,
        x2.shape[0], x2.shape[1], float(eps), int(quant_mode), float(quant_eps),
        _lib.dtype_code(x2.dtype), _lib.current_stream())
    _lib.check(rc, "b200_rmsnorm")


class RMSNorm(nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.variance_epsilon = eps

    def _weight_like(self, x):
        """Weight in the activation dtype (the reference multiplies the fp32-promoted weight,
        layernorm.py:104; checkpoints store it in the model dtype, so the cast is exact).  The
        cast copy is cached until the parameter changes."""
        w = self.weight.data
        if w.dtype == x.dtype:
            return w
        key = (w.data_ptr(), w._version, x.dtype)
        if getattr(self, "_wcast_key", None) != key:
            self._wcast, self._wcast_key = w.to(x.dtype), key
        return self._wcast

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None
                ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        x2 = _rows(x)
        out = torch.empty_like(x2)
        if residual is not None:
            r2 = _rows(residual)
            rmsnorm_call(x2, r2, self._weight_like(x2), self.variance_epsilon, out, None, None, 0)
            if r2.data_ptr() != residual.data_ptr():
                residual.copy_(r2.view(residual.shape))
            return out.view(x.shape), residual
        rmsnorm_call(x2, None, self._weight_like(x2), self.variance_epsilon, out, None, None, 0)
        return out.view(x.shape)

    def forward_quant_fp8(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None,
                          group_size: Optional[int] = None, column_major_scales: bool = False):
        """-> (x_q e4m3 [M,K], x_s fp32 [M,1] or [M,K/128]) (+ residual updated in place)."""
        x2 = _rows(x)
        m, k = x2.shape
        r2 = None
        if residual is not None:
            r2 = _rows(residual)
            assert r2.data_ptr() == residual.data_ptr(), "residual must have contiguous rows"
        q = torch.empty((m, k), device=x.device, dtype=FP8)
        if group_size is None:
            s = torch.empty((m, 1), device=x.device, dtype=torch.float32)
            rmsnorm_call(x2, r2, self._weight_like(x2), self.variance_epsilon, None, q, s, 1)
        else:
            if group_size != 128:
                raise _lib.B200Error("RMSNorm.forward_quant_fp8: group_size must be 128")
            if column_major_scales:
                s = torch.empty((k // 128, m), device=x.device, dtype=torch.float32).t()
            else:
                s = torch.empty((m, k // 128), device=x.device, dtype=torch.float32)
            rmsnorm_call(x2, r2, self._weight_like(x2), self.variance_epsilon, None, q, s, 2)
        return q, s
