"""Fp8Config / Fp8LinearMethod with the reference's parameter contract
(python/sglang/srt/layers/quantization/fp8.py:97-439) and B200 kernels behind apply()."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .base_config import LinearMethodBase, QuantizationConfig, set_weight_attrs
from .fp8_kernel import per_token_group_quant_fp8
from .fp8_utils import apply_fp8_linear, apply_w8a8_block_fp8_linear, convert_to_channelwise

ACTIVATION_SCHEMES = ["static", "dynamic"]


class Fp8Config(QuantizationConfig):
    def __init__(self, is_checkpoint_fp8_serialized: bool = False,
                 activation_scheme: str = "dynamic", ignored_layers: Optional[List[str]] = None,
                 weight_block_size: Optional[List[int]] = None):
        if activation_scheme not in ACTIVATION_SCHEMES:
            raise ValueError(f"Unsupported activation scheme {activation_scheme}")
        self.is_checkpoint_fp8_serialized = is_checkpoint_fp8_serialized
        self.activation_scheme = activation_scheme
        self.ignored_layers = ignored_layers or []
        if weight_block_size is not None:
            if not is_checkpoint_fp8_serialized:
                raise ValueError("block-wise fp8 needs an fp8-serialized checkpoint")
            if len(weight_block_size) != 2:
                raise ValueError("weight_block_size must be [block_n, block_k]")
            if activation_scheme != "dynamic":
                raise ValueError("block-wise fp8 only supports dynamic activation quantisation")
        self.weight_block_size = weight_block_size

    def get_name(self) -> str:
        return "fp8"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.bfloat16, torch.half]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "Fp8Config":
        quant_method = cls.get_from_keys(config, ["quant_method"])
        return cls(is_checkpoint_fp8_serialized="fp8" in quant_method,
                   activation_scheme=cls.get_from_keys(config, ["activation_scheme"]),
                   ignored_layers=cls.get_from_keys_or(config, ["ignored_layers"], None),
                   weight_block_size=cls.get_from_keys_or(config, ["weight_block_size"], None))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = ""):
        from ..linear import LinearBase

        if isinstance(layer, LinearBase):
            if any(prefix.startswith(p) or p in prefix for p in self.ignored_layers if p):
                from ..linear import UnquantizedLinearMethod

                return UnquantizedLinearMethod()
            return Fp8LinearMethod(self)
        return None


class Fp8LinearMethod(LinearMethodBase):
    """FP8 linear: fp8-serialized checkpoints (per-tensor-per-shard or 128x128 block scales,
    dynamic or static activation scale) and bf16/f16 checkpoints quantised per channel at load."""

    def __init__(self, quant_config: Fp8Config):
        self.quant_config = quant_config
        self.cutlass_fp8_supported = True  # per-channel scaled GEMM is always available here
        self.block_quant = quant_config.weight_block_size is not None

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        n = sum(output_partition_sizes)
        k = input_size_per_partition
        tp_size = extra_weight_attrs.pop("tp_size", 1)
        weight_loader = extra_weight_attrs.get("weight_loader")
        if self.block_quant:
            block_n, block_k = self.quant_config.weight_block_size
            if tp_size > 1 and input_size // k == tp_size and k % block_k != 0:
                raise ValueError(f"Weight input_size_per_partition = {k} is not divisible by "
                                 f"weight quantization block_k = {block_k}.")
            if (tp_size > 1 and output_size // n == tp_size) or len(output_partition_sizes) > 1:
                for part in output_partition_sizes:
                    if part % block_n != 0:
                        raise ValueError(f"Weight output_partition_size = {part} is not divisible "
                                         f"by weight quantization block_n = {block_n}.")
        layer.logical_widths = output_partition_sizes
        layer.input_size_per_partition = k
        layer.output_size_per_partition = n
        layer.orig_dtype = params_dtype
        serialized = self.quant_config.is_checkpoint_fp8_serialized
        weight = Parameter(torch.empty(n, k, dtype=torch.float8_e4m3fn if serialized else params_dtype),
                           requires_grad=False)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0, "weight_loader": weight_loader})
        layer.register_parameter("weight", weight)
        if not serialized:
            return
        fmin = torch.finfo(torch.float32).min
        if self.block_quant:
            scale = Parameter(torch.full(((n + block_n - 1) // block_n, (k + block_k - 1) // block_k),
                                         fmin, dtype=torch.float32), requires_grad=False)
            # block_n: fused-shard loaders convert row windows to scale-grid rows (linear.py:731-741)
            set_weight_attrs(scale, {"input_dim": 1, "output_dim": 0, "block_n": block_n,
                                     "weight_loader": weight_loader})
            layer.register_parameter("weight_scale_inv", scale)
        else:
            scale = Parameter(torch.full((len(output_partition_sizes),), fmin, dtype=torch.float32),
                              requires_grad=False)
            set_weight_attrs(scale, {"weight_loader": weight_loader})
            layer.register_parameter("weight_scale", scale)
        if self.quant_config.activation_scheme == "static":
            scale = Parameter(torch.full((len(output_partition_sizes),), fmin, dtype=torch.float32),
                              requires_grad=False)
            set_weight_attrs(scale, {"weight_loader": weight_loader})
            layer.register_parameter("input_scale", scale)
        else:
            layer.register_parameter("input_scale", None)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        if self.block_quant:
            layer.weight = Parameter(layer.weight.data, requires_grad=False)
            layer.weight_scale_inv = Parameter(layer.weight_scale_inv.data, requires_grad=False)
            return
        if not self.quant_config.is_checkpoint_fp8_serialized:
            # 16-bit checkpoint: per-channel quantisation (one group per weight row)
            qweight, weight_scale = per_token_group_quant_fp8(layer.weight.data,
                                                              layer.weight.shape[-1])
            layer.weight = Parameter(qweight.t(), requires_grad=False)
            layer.weight_scale = Parameter(weight_scale.t().contiguous(), requires_grad=False)
            layer.input_scale = None
            return
        # fp8 checkpoint with one scale per fused shard -> one scale per output channel
        weight_scale = convert_to_channelwise(layer.weight_scale.data, layer.logical_widths)
        layer.weight = Parameter(layer.weight.data.t(), requires_grad=False)
        layer.weight_scale = Parameter(weight_scale, requires_grad=False)
        if self.quant_config.activation_scheme == "static":
            layer.input_scale = Parameter(layer.input_scale.data.max(), requires_grad=False)

    # RowParallelLinear.forward may hand the TP communicator over in `layer.fused_all_reduce`: the
    # GEMM then returns the all-reduced output (one kernel in the decode regime).
    supports_fused_all_reduce = True

    def apply(self, layer: torch.nn.Module, x: torch.Tensor,
              bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        comm = getattr(layer, "fused_all_reduce", None)
        if self.block_quant:
            return apply_w8a8_block_fp8_linear(input=x, weight=layer.weight,
                                               block_size=self.quant_config.weight_block_size,
                                               weight_scale=layer.weight_scale_inv,
                                               input_scale=None, bias=bias, comm=comm)
        return apply_fp8_linear(input=x, weight=layer.weight, weight_scale=layer.weight_scale,
                                input_scale=layer.input_scale, bias=bias,
                                cutlass_fp8_supported=True, use_per_token_if_dynamic=False,
                                comm=comm)
