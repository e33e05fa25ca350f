"""GPTQConfig / GPTQLinearMethod.  The reference only parses the config
(python/sglang/srt/layers/quantization/gptq.py:57-160) and takes the linear method from vllm 0.8.4
(gptq.py:17-31, quantization/__init__.py:216-230) — absent here, so the parameter contract follows
the public AutoGPTQ v1 checkpoint layout that vllm's GPTQLinearMethod registers:
  qweight [K/8, N] int32 (K-packed), qzeros [K/g, N/8] int32 (zero - 1), scales [K/g, N], g_idx [K].
process_weights_after_loading repacks once into the B200 layout shared with AWQ; desc_act
checkpoints are handled by sorting rows by group (perm = argsort(g_idx)) and gathering x columns."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .base_config import LinearMethodBase, QuantizationConfig, set_weight_attrs
from .w4a16_kernel import (gemv_stream_repack, repack_gptq, stream_copy_enabled, unpack_zeros,
                           w4a16_linear)


class GPTQConfig(QuantizationConfig):
    def __init__(self, weight_bits: int, group_size: int, desc_act: bool, lm_head_quantized: bool = False,
                 dynamic: Optional[Dict[str, Dict[str, Any]]] = None):
        self.dynamic = dynamic or {}
        self.weight_bits = weight_bits
        self.group_size = group_size
        self.desc_act = desc_act
        self.lm_head_quantized = lm_head_quantized
        if weight_bits != 4:
            raise ValueError("The B200 GPTQ path supports 4-bit weights only "
                             f"(reference allows 2/3/4/8 through vllm), got {weight_bits} bits.")
        self.pack_factor = 32 // weight_bits

    def __repr__(self) -> str:
        return (f"GPTQConfig(weight_bits={self.weight_bits}, group_size={self.group_size}, "
                f"desc_act={self.desc_act}), lm_head_quantized={self.lm_head_quantized}), "
                f"dynamic={self.dynamic}")

    def get_name(self) -> str:
        return "gptq"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.half, torch.bfloat16]

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "GPTQConfig":
        dynamic = cls.get_from_keys_or(config, ["dynamic"], default={}) or {}
        return cls(cls.get_from_keys(config, ["bits"]), cls.get_from_keys(config, ["group_size"]),
                   cls.get_from_keys(config, ["desc_act"]),
                   cls.get_from_keys_or(config, ["lm_head"], default=False), dynamic)

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = ""):
        from ..linear import LinearBase

        return GPTQLinearMethod(self) if isinstance(layer, LinearBase) else None


class GPTQLinearMethod(LinearMethodBase):
    def __init__(self, quant_config: GPTQConfig):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        cfg = self.quant_config
        group = cfg.group_size if cfg.group_size != -1 else input_size
        if input_size_per_partition % group != 0 or input_size_per_partition % cfg.pack_factor != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        n = sum(output_partition_sizes)
        if n % cfg.pack_factor != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        if cfg.desc_act and input_size_per_partition != input_size:
            # act-order rows of one K shard belong to arbitrary GLOBAL groups: the reference's
            # dependency (vllm GPTQLinearMethod.create_weights) then loads scales / zeros
            # unsharded and indexes them through g_idx.  The B200 kernel needs group-contiguous
            # rows per rank, which such a shard cannot provide -> fail loudly, not silently wrong.
            raise NotImplementedError("GPTQ desc_act=True on a row-parallel layer with tp_size > 1 "
                                      "is not supported by the B200 W4A16 kernel")
        weight_loader = extra_weight_attrs.get("weight_loader")
        k = input_size_per_partition
        qweight = Parameter(torch.empty(k // cfg.pack_factor, n, dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, {"input_dim": 0, "output_dim": 1, "packed_dim": 0,
                                   "packed_factor": cfg.pack_factor, "weight_loader": weight_loader})
        g_idx = Parameter(torch.arange(k, dtype=torch.int32) // group, requires_grad=False)
        set_weight_attrs(g_idx, {"input_dim": 0, "weight_loader": weight_loader})
        qzeros = Parameter(torch.empty(k // group, n // cfg.pack_factor, dtype=torch.int32),
                           requires_grad=False)
        set_weight_attrs(qzeros, {"input_dim": 0, "output_dim": 1, "packed_dim": 1,
                                  "packed_factor": cfg.pack_factor, "weight_loader": weight_loader})
        scales = Parameter(torch.empty(k // group, n, dtype=params_dtype), requires_grad=False)
        set_weight_attrs(scales, {"input_dim": 0, "output_dim": 1, "weight_loader": weight_loader})
        layer.register_parameter("qweight", qweight)
        layer.register_parameter("g_idx", g_idx)
        layer.register_parameter("qzeros", qzeros)
        layer.register_parameter("scales", scales)
        layer.w4a16_group_size = group

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        k = layer.qweight.shape[0] * self.quant_config.pack_factor
        group = layer.w4a16_group_size
        g_idx = layer.g_idx.data.to(torch.int64)
        # desc_act=False: checkpoints still store g_idx (AutoGPTQ writes arange(K)//group with
        # GLOBAL group ids, so a K shard of rank > 0 never equals the local arange) -> ignore it.
        trivial = (not self.quant_config.desc_act) or bool(
            torch.equal(g_idx, torch.arange(k, device=g_idx.device) // group))
        perm = None
        if not trivial:
            # desc_act: make every group's rows contiguous, remember how to gather x columns
            if not bool((torch.bincount(g_idx, minlength=k // group) == group).all()):
                raise ValueError("GPTQ g_idx does not describe equally sized groups")
            perm = torch.argsort(g_idx, stable=True)
        layer.b200_qweight = repack_gptq(layer.qweight.data, perm)
        layer.b200_zeros = unpack_zeros(layer.qzeros.data, layer.scales.dtype, awq_order=False,
                                        add_one=True)
        layer.b200_perm = perm
        layer.scales = Parameter(layer.scales.data.contiguous(), requires_grad=False)
        # decode batches stream a second, MMA-fragment-ordered copy of the same nibbles
        # (+0.5 byte per weight; B200_W4A16_STREAM_COPY=0 keeps only the [K, N/8] layout)
        layer.b200_qweight_stream = (gemv_stream_repack(layer.b200_qweight)
                                     if stream_copy_enabled() else None)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor,
              bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        out_shape = x.shape[:-1] + (layer.b200_qweight.shape[-1] * self.quant_config.pack_factor,)
        x2 = x.reshape(-1, x.shape[-1])
        if layer.b200_perm is not None:
            x2 = x2.index_select(-1, layer.b200_perm)
        elif x2.stride(-1) != 1:
            x2 = x2.contiguous()
        out = w4a16_linear(x2, layer, bias)
        return out.reshape(out_shape)
