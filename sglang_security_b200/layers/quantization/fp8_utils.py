"""FP8 linear glue: activation quantisation + scaled GEMM, mirroring
python/sglang/srt/layers/quantization/fp8_utils.py (apply_fp8_linear :454-693,
apply_w8a8_block_fp8_linear :152-304) and quantization/utils.py:68-88 (convert_to_channelwise).
Only the sm_100a tcgen05 kernels are dispatched — no cuBLASLt / Triton / DeepGEMM alternatives."""
from __future__ import annotations

from typing import List, Optional

import torch

from .fp8_kernel import (fp8_blockwise_scaled_mm, fp8_scaled_mm, per_token_group_quant_fp8,
                         sglang_per_token_quant_fp8, static_quant_fp8)


class QuantizedActivation:
    """An activation that was quantised by the kernel that produced it (fused RMSNorm->quant,
    SiLU*mul->quant): `q` e4m3 [M, K], `scale` fp32 [M, 1] (per token) or [M, K/128] (per group),
    `dtype` = the 16-bit dtype the linear output must have.  Passing it to LinearBase.forward /
    LinearMethod.apply skips the activation quantiser the reference runs inside apply_fp8_linear
    (fp8_utils.py:454-648); the bytes are the ones that quantiser would have produced."""

    __slots__ = ("q", "scale", "dtype")

    def __init__(self, q: torch.Tensor, scale: torch.Tensor, dtype: torch.dtype):
        self.q, self.scale, self.dtype = q, scale, dtype

    @property
    def shape(self):
        return self.q.shape


def convert_to_channelwise(weight_scale: torch.Tensor, logical_widths: List[int]) -> torch.Tensor:
    """Per-partition (fused qkv / gate_up shard) scales -> one scale per output channel [N, 1]."""
    total = sum(logical_widths)
    if weight_scale.dim() == 0:
        return weight_scale.reshape(1, 1).expand(total, 1).to(torch.float32).contiguous()
    widths = torch.tensor(logical_widths, device=weight_scale.device)
    return torch.repeat_interleave(weight_scale.to(torch.float32), widths).reshape(total, 1)


def apply_fp8_linear(input: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
                     input_scale: Optional[torch.Tensor] = None,
                     input_scale_ub: Optional[torch.Tensor] = None,
                     bias: Optional[torch.Tensor] = None, cutlass_fp8_supported: bool = True,
                     use_per_token_if_dynamic: bool = False, comm=None) -> torch.Tensor:
    """weight is the [K, N] column-major view of the e4m3 [N, K] weight; weight_scale has one
    entry per output channel (or is a scalar).  `comm` (a B200AllReduce) makes the GEMM reduce its
    output across the tensor-parallel group (row-parallel layers)."""
    if weight_scale.numel() not in (1, weight.shape[1]):
        raise ValueError("B200 fp8 linear supports per-channel (or per-tensor) weight scales only")
    if isinstance(input, QuantizedActivation):
        if input_scale is not None or input.scale.shape[-1] != 1:
            raise ValueError("pre-quantised activation must be dynamic per-token for this layer")
        return fp8_scaled_mm(input.q, weight, input.scale, weight_scale, out_dtype=input.dtype,
                             bias=bias, comm=comm)
    input_2d = input.reshape(-1, input.shape[-1])
    if input_2d.stride(-1) != 1:
        input_2d = input_2d.contiguous()
    output_shape = [*input.shape[:-1], weight.shape[1]]
    if input_scale is not None:
        qinput, x_scale = static_quant_fp8(input_2d, input_scale, repeat_scale=False)
    else:
        qinput, x_scale = sglang_per_token_quant_fp8(input_2d)
    output = fp8_scaled_mm(qinput, weight, x_scale, weight_scale, out_dtype=input.dtype, bias=bias,
                           comm=comm)
    return output.view(*output_shape)


def apply_w8a8_block_fp8_linear(input: torch.Tensor, weight: torch.Tensor, block_size: List[int],
                                weight_scale: torch.Tensor,
                                input_scale: Optional[torch.Tensor] = None,
                                bias: Optional[torch.Tensor] = None, comm=None) -> torch.Tensor:
    """Block-quantised linear: weight [N, K] e4m3 with [ceil(N/128), ceil(K/128)] fp32 scales,
    activations quantised per (token, 128-group) on the fly."""
    assert input_scale is None
    if list(block_size) != [128, 128]:
        raise ValueError(f"B200 block-fp8 linear supports 128x128 weight blocks, got {block_size}")
    if isinstance(input, QuantizedActivation):
        return fp8_blockwise_scaled_mm(input.q, weight, input.scale, weight_scale,
                                       out_dtype=input.dtype, bias=bias, comm=comm)
    input_2d = input.reshape(-1, input.shape[-1])
    if input_2d.stride(-1) != 1:
        input_2d = input_2d.contiguous()
    output_shape = [*input.shape[:-1], weight.shape[0]]
    q_input, x_scale = per_token_group_quant_fp8(input_2d, block_size[1], column_major_scales=False)
    output = fp8_blockwise_scaled_mm(q_input, weight, x_scale, weight_scale, out_dtype=input.dtype,
                                     bias=bias, comm=comm)
    return output.view(*output_shape)


def dispatch_w8a8_block_fp8_linear():
    return apply_w8a8_block_fp8_linear
