"""SiluAndMul with the reference's interface (python/sglang/srt/layers/activation.py:42-51), one
sm_100a kernel through the C-ABI (b200_silu_and_mul).  `forward_quant_fp8` additionally fuses the
FP8 activation quantiser of the following down_proj (fp8_utils.py:454-648)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import _lib

FP8 = torch.float8_e4m3fn


def _call(x2, d, out, q, scales, mode, eps=1e-10):
    s_row, s_col = (scales.stride(0), scales.stride(1)) if (scales is not None and scales.dim() == 2) \
        else (1, 0)
    rc = _lib.load().b200_silu_and_mul(
        x2.data_ptr(), x2.stride(0), _lib.ptr(out), out.stride(0) if out is not None else 0,
        _lib.ptr(q), q.stride(0) if q is not None else 0, _lib.ptr(scales), s_row, s_col,
        x2.shape[0], d, int(mode), float(eps), _lib.dtype_code(x2.dtype), _lib.current_stream())
    _lib.check(rc, "b200_silu_and_mul")


def _rows(x):
    if not x.is_cuda:
        raise _lib.B200Error("SiluAndMul: tensor must live on a CUDA device (no CPU path)")
    x2 = x.reshape(-1, x.shape[-1])
    return x2 if x2.stride(-1) == 1 else x2.contiguous()


class SiluAndMul(nn.Module):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        d = x.shape[-1] // 2
        x2 = _rows(x)
        out = torch.empty((x2.shape[0], d), dtype=x.dtype, device=x.device)
        _call(x2, d, out, None, None, 0)
        return out.view(x.shape[:-1] + (d,))

    def forward_quant_fp8(self, x: torch.Tensor, group_size: Optional[int] = None,
                          column_major_scales: bool = False):
        d = x.shape[-1] // 2
        x2 = _rows(x)
        m = x2.shape[0]
        q = torch.empty((m, d), device=x.device, dtype=FP8)
        if group_size is None:
            s = torch.empty((m, 1), device=x.device, dtype=torch.float32)
            _call(x2, d, None, q, s, 1)
        else:
            if group_size != 128:
                raise _lib.B200Error("SiluAndMul.forward_quant_fp8: group_size must be 128")
            s = (torch.empty((d // 128, m), device=x.device, dtype=torch.float32).t()
                 if column_major_scales else
                 torch.empty((m, d // 128), device=x.device, dtype=torch.float32))
            _call(x2, d, None, q, s, 2)
        return q, s
