"""KV-cache quantisation method: gives an attention layer loadable `k_scale` / `v_scale` parameters
and resolves them after loading — mirror of the reference's BaseKVCacheMethod
(python/sglang/srt/layers/quantization/kv_cache.py:17-87; Fp8KVCacheMethod, fp8.py:1096-1102).
The resolved host floats `k_scale_float` / `v_scale_float` are what b200_kv_write_fp8 /
b200_rope_kv_write_fp8 divide by and b200_attn_decode_fp8kv multiplies with."""
from __future__ import annotations

import torch

from .base_config import QuantizationConfig, QuantizeMethodBase


class BaseKVCacheMethod(QuantizeMethodBase):
    def __init__(self, quant_config: QuantizationConfig):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module):
        # -1.0 = "not in the checkpoint"; a loaded k_scale / v_scale overwrites it
        layer.k_scale = torch.nn.Parameter(torch.tensor(-1.0, dtype=torch.float32), requires_grad=False)
        layer.v_scale = torch.nn.Parameter(torch.tensor(-1.0, dtype=torch.float32), requires_grad=False)

    def apply(self, layer: torch.nn.Module) -> torch.Tensor:
        raise RuntimeError(f"{self.__class__.__name__}.apply should not be called.")

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        ks, vs = layer.k_scale, layer.v_scale
        if ks.numel() != 1 or vs.numel() != 1:
            raise ValueError("Only support per-tensor scaling factor for fp8 KV cache")
        k, v = float(ks), float(vs)
        if k > 0.0 and v > 0.0:        # separate scales in the checkpoint
            pass
        elif k < 0.0 and v < 0.0:      # none loaded
            k = v = 1.0
        else:                          # a single (deprecated) kv_scale was remapped to k_scale
            assert k > 0.0, "a lone v_scale without k_scale is not a supported checkpoint format"
            k = v = max(k, v)
        with torch.no_grad():
            ks.fill_(k)
            vs.fill_(v)
        layer.k_scale_float = k        # host copies: read per step without a device sync
        layer.v_scale_float = v


class Fp8KVCacheMethod(BaseKVCacheMethod):
    """KV-cache scaling factors of FP8 checkpoints."""
