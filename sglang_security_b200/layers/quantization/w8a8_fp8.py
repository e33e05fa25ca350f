"""W8A8Fp8Config / W8A8Fp8LinearMethod: static per-channel weight scales x dynamic per-token
activation scales (python/sglang/srt/layers/quantization/w8a8_fp8.py:28-183) — BASELINE's
"per-channel" FP8 configuration."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .base_config import LinearMethodBase, QuantizationConfig, set_weight_attrs
from .fp8_kernel import per_token_group_quant_fp8
from .fp8_utils import apply_fp8_linear


class W8A8Fp8Config(QuantizationConfig):
    def __init__(self, is_checkpoint_fp8_serialized: bool = False):
        self.is_checkpoint_fp8_serialized = is_checkpoint_fp8_serialized

    def get_name(self) -> str:
        return "w8a8_fp8"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "W8A8Fp8Config":
        quant_method = cls.get_from_keys(config, ["quant_method"])
        return cls(is_checkpoint_fp8_serialized="compressed-tensors" in quant_method)

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = ""):
        from ..linear import LinearBase

        return W8A8Fp8LinearMethod(self) if isinstance(layer, LinearBase) else None


class W8A8Fp8LinearMethod(LinearMethodBase):
    def __init__(self, quantization_config: W8A8Fp8Config):
        self.cutlass_fp8_supported = True
        self.quantization_config = quantization_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        serialized = self.quantization_config.is_checkpoint_fp8_serialized
        weight_loader = extra_weight_attrs.get("weight_loader")
        self.logical_widths = output_partition_sizes
        n = sum(output_partition_sizes)
        weight = Parameter(torch.empty(n, input_size_per_partition,
                                       dtype=torch.float8_e4m3fn if serialized else params_dtype),
                           requires_grad=False)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0, "weight_loader": weight_loader})
        layer.register_parameter("weight", weight)
        if serialized:
            scale = Parameter(torch.empty((n, 1), dtype=torch.float32), requires_grad=False)
            set_weight_attrs(scale, {"output_dim": 0, "weight_loader": weight_loader})
            layer.register_parameter("weight_scale", scale)
        else:
            layer.weight_scale = None

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        if self.quantization_config.is_checkpoint_fp8_serialized:
            layer.weight = Parameter(layer.weight.data.t(), requires_grad=False)
            layer.weight_scale = Parameter(layer.weight_scale.data.detach(), requires_grad=False)
        else:
            qweight, weight_scale = per_token_group_quant_fp8(layer.weight.data,
                                                              layer.weight.shape[-1])
            layer.weight = Parameter(qweight.t(), requires_grad=False)
            layer.weight_scale = Parameter(weight_scale.t().contiguous(), requires_grad=False)
            layer.input_scale = None

    supports_fused_all_reduce = True  # see Fp8LinearMethod

    def apply(self, layer: torch.nn.Module, x: torch.Tensor,
              bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        return apply_fp8_linear(x, layer.weight, layer.weight_scale, bias=bias,
                                cutlass_fp8_supported=True,
                                comm=getattr(layer, "fused_all_reduce", None))
