"""RadixAttention layer stand-in: the per-layer attributes an AttentionBackend reads and the
k/v reshape + dispatch of the reference layer (python/sglang/srt/layers/radix_attention.py:38-109)."""
from __future__ import annotations

from enum import Enum

import torch.nn as nn


class AttentionType(Enum):
    DECODER = "decoder"
    ENCODER_ONLY = "encoder_only"


class RadixAttention(nn.Module):
    def __init__(self, num_heads: int, head_dim: int, scaling: float, num_kv_heads: int,
                 layer_id: int, logit_cap: float = 0.0, v_head_dim: int = -1,
                 sliding_window_size: int = -1, is_cross_attention: bool = False,
                 attn_type: AttentionType = AttentionType.DECODER, quant_config=None,
                 prefix: str = ""):
        super().__init__()
        self.tp_q_head_num = num_heads
        self.tp_k_head_num = num_kv_heads
        self.tp_v_head_num = num_kv_heads
        self.head_dim = head_dim
        self.qk_head_dim = head_dim
        self.v_head_dim = head_dim if v_head_dim == -1 else v_head_dim
        self.scaling = scaling
        self.layer_id = layer_id
        self.logit_cap = logit_cap
        self.sliding_window_size = sliding_window_size or -1
        self.is_cross_attention = is_cross_attention
        self.attn_type = attn_type
        self.k_scale = None
        self.v_scale = None
        self.k_scale_float = None   # host copies (radix_attention.py:74-75): no .item() per step
        self.v_scale_float = None
        # a KV-cache quant method (layers/quantization/kv_cache.py) turns k_scale / v_scale into
        # loadable parameters (radix_attention.py:76-80)
        method = quant_config.get_quant_method(self, prefix=prefix) if quant_config is not None else None
        self.quant_method = method
        if method is not None:
            method.create_weights(self)

    def forward(self, q, k, v, forward_batch, save_kv_cache: bool = True, **kwargs):
        if k is not None:
            assert v is not None
            k = k.view(-1, self.tp_k_head_num, self.qk_head_dim)
            v = v.view(-1, self.tp_v_head_num, self.v_head_dim)
        return forward_batch.attn_backend.forward(q, k, v, self, forward_batch, save_kv_cache,
                                                  **kwargs)
