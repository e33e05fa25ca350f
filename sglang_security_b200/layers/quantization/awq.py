"""AWQConfig / AWQLinearMethod with the reference's parameter contract
(python/sglang/srt/layers/quantization/awq.py:27-215): qweight [K, N/8] int32, qzeros [K/g, N/8]
int32, scales [K/g, N].  apply() is ONE fused dequant+GEMM kernel instead of the reference's
full-weight dequant followed by a dense matmul on every call."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .base_config import LinearMethodBase, QuantizationConfig, set_weight_attrs
from .w4a16_kernel import gemv_stream_repack, stream_copy_enabled, unpack_zeros, w4a16_linear


def is_layer_skipped_awq(prefix: str, modules_to_not_convert: List[str]) -> bool:
    return any(module_name in prefix for module_name in modules_to_not_convert)


class AWQConfig(QuantizationConfig):
    def __init__(self, weight_bits: int, group_size: int, zero_point: bool,
                 modules_to_not_convert: Optional[List[str]] = None):
        self.weight_bits = weight_bits
        self.group_size = group_size
        self.zero_point = zero_point
        self.modules_to_not_convert = modules_to_not_convert or []
        if weight_bits != 4:
            raise ValueError("Currently, only 4-bit weight quantization is supported for AWQ, "
                             f"but got {weight_bits} bits.")
        self.pack_factor = 32 // weight_bits

    def __repr__(self) -> str:
        return (f"AWQConfig(weight_bits={self.weight_bits}, group_size={self.group_size}, "
                f"zero_point={self.zero_point}, modules_to_not_convert={self.modules_to_not_convert})")

    def get_name(self) -> str:
        return "awq"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        # the reference declares fp16 only (awq.py:67-68) although its dequant kernel also handles
        # bf16 (awq_kernel.cu:160-180); both are supported and tested here
        return [torch.half, torch.bfloat16]

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quant_config.json", "quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "AWQConfig":
        return cls(cls.get_from_keys(config, ["w_bit", "bits"]),
                   cls.get_from_keys(config, ["q_group_size", "group_size"]),
                   cls.get_from_keys(config, ["zero_point"]),
                   cls.get_from_keys_or(config, ["modules_to_not_convert"], None))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = ""):
        from ..linear import LinearBase, UnquantizedLinearMethod

        if isinstance(layer, LinearBase):
            if is_layer_skipped_awq(prefix, self.modules_to_not_convert):
                return UnquantizedLinearMethod()
            return AWQLinearMethod(self)
        return None


class AWQLinearMethod(LinearMethodBase):
    def __init__(self, quant_config: AWQConfig):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        cfg = self.quant_config
        group = cfg.group_size if cfg.group_size != -1 else input_size_per_partition
        if input_size_per_partition % group != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        n = sum(output_partition_sizes)
        if n % cfg.pack_factor != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        weight_loader = extra_weight_attrs.get("weight_loader")
        packed = {"input_dim": 0, "output_dim": 1, "packed_dim": 1,
                  "packed_factor": cfg.pack_factor, "weight_loader": weight_loader}
        qweight = Parameter(torch.empty(input_size_per_partition, n // cfg.pack_factor,
                                        dtype=torch.int32), requires_grad=False)
        qzeros = Parameter(torch.empty(input_size_per_partition // group, n // cfg.pack_factor,
                                       dtype=torch.int32), requires_grad=False)
        scales = Parameter(torch.empty(input_size_per_partition // group, n, dtype=params_dtype),
                           requires_grad=False)
        set_weight_attrs(qweight, packed)
        set_weight_attrs(qzeros, packed)
        set_weight_attrs(scales, {"input_dim": 0, "output_dim": 1, "weight_loader": weight_loader})
        layer.register_parameter("qweight", qweight)
        layer.register_parameter("qzeros", qzeros)
        layer.register_parameter("scales", scales)
        layer.w4a16_group_size = group

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        layer.qweight = Parameter(layer.qweight.data.contiguous(), requires_grad=False)
        layer.qzeros = Parameter(layer.qzeros.data, requires_grad=False)
        layer.scales = Parameter(layer.scales.data.contiguous(), requires_grad=False)
        # one-time repack: packed zeros -> biased 16-bit zeros the fused kernel subtracts directly
        layer.b200_zeros = unpack_zeros(layer.qzeros.data, layer.scales.dtype, awq_order=True,
                                        add_one=False)
        # decode batches stream a second, MMA-fragment-ordered copy of the same nibbles
        # (+0.5 byte per weight; B200_W4A16_STREAM_COPY=0 keeps only the [K, N/8] layout)
        layer.b200_qweight_stream = (gemv_stream_repack(layer.qweight.data)
                                     if stream_copy_enabled() else None)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor,
              bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        out_shape = x.shape[:-1] + (layer.qweight.shape[-1] * self.quant_config.pack_factor,)
        x2 = x.reshape(-1, x.shape[-1])
        if x2.stride(-1) != 1:
            x2 = x2.contiguous()
        out = w4a16_linear(x2, layer, bias)
        return out.reshape(out_shape)
