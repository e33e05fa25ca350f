"""One decoder layer of the Llama / Qwen2 family over the B200 hot-path ops.

Call sequence = the reference's LlamaDecoderLayer.forward / LlamaAttention.forward / LlamaMLP.forward
(python/sglang/srt/models/llama.py:94-98, 180-191, 245-267; qwen2.py has the same structure with a
qkv bias):

    hidden, residual = input_layernorm(hidden, residual)
    qkv = qkv_proj(hidden); q, k, v = split(qkv); q, k = rotary_emb(positions, q, k)
    attn = RadixAttention(q, k, v, forward_batch); hidden = o_proj(attn)
    hidden, residual = post_attention_layernorm(hidden, residual)
    hidden = down_proj(SiluAndMul(gate_up_proj(hidden)))

With `fuse=True` (default) the steps either side of the GEMMs / attention use the fused kernels
(SURVEY §8f rows 1-2): RMSNorm->FP8 quant, RoPE->KV-pool write, SiLU*mul->FP8 quant — 11 launches
per layer in decode instead of 16.  With `fuse=False` every op is launched separately exactly in
the reference's order (used by the tests to prove both give the same result).
The scheduler / model runner stay the reference's: this class only needs `forward_batch`.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from ..layers.activation import SiluAndMul
from ..layers.layernorm import RMSNorm
from ..layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
from ..layers.quantization.fp8_utils import QuantizedActivation
from ..layers.radix_attention import RadixAttention
from ..layers.rotary_embedding import RotaryEmbedding


def _fp8_mode(quant_config):
    """-> None (16-bit / int4 linear), "token" (per-token activations) or "group" (block-128)."""
    if quant_config is None:
        return None
    name = type(quant_config).__name__
    if name == "W8A8Fp8Config":
        return "token"
    if name == "Fp8Config":
        if getattr(quant_config, "weight_block_size", None):
            return "group"
        return "token" if quant_config.activation_scheme == "dynamic" else None
    return None


class B200DecoderLayer(nn.Module):
    def __init__(self, hidden_size: int, num_heads: int, num_kv_heads: int, head_dim: int,
                 intermediate_size: int, layer_id: int, rotary_emb: RotaryEmbedding,
                 quant_config=None, params_dtype: torch.dtype = torch.bfloat16,
                 rms_norm_eps: float = 1e-6, qkv_bias: bool = False, fuse: bool = True,
                 tp_rank: Optional[int] = None, tp_size: Optional[int] = None):
        super().__init__()
        kw = dict(params_dtype=params_dtype, quant_config=quant_config, tp_rank=tp_rank,
                  tp_size=tp_size)
        self.qkv_proj = QKVParallelLinear(hidden_size, head_dim, num_heads, num_kv_heads,
                                          bias=qkv_bias, **kw)
        self.o_proj = RowParallelLinear(num_heads * head_dim, hidden_size, bias=False, **kw)
        self.gate_up_proj = MergedColumnParallelLinear(hidden_size, [intermediate_size] * 2,
                                                       bias=False, **kw)
        self.down_proj = RowParallelLinear(intermediate_size, hidden_size, bias=False, **kw)
        self.q_size = self.qkv_proj.num_heads * head_dim
        self.kv_size = self.qkv_proj.num_kv_heads * head_dim
        self.attn = RadixAttention(self.qkv_proj.num_heads, head_dim, head_dim ** -0.5,
                                   self.qkv_proj.num_kv_heads, layer_id)
        self.rotary_emb = rotary_emb
        self.input_layernorm = RMSNorm(hidden_size, rms_norm_eps)
        self.post_attention_layernorm = RMSNorm(hidden_size, rms_norm_eps)
        self.act_fn = SiluAndMul()
        self.fuse = fuse
        self.fp8_mode = _fp8_mode(quant_config)
        self.params_dtype = params_dtype

    # -- helpers ---------------------------------------------------------------------------------
    def _norm(self, norm: RMSNorm, x, residual):
        """-> (activation for the next linear, residual)."""
        if residual is None:
            residual = x
            if self.fuse and self.fp8_mode:
                q, s = norm.forward_quant_fp8(x, None, 128 if self.fp8_mode == "group" else None)
                return QuantizedActivation(q, s, x.dtype), residual
            return norm(x), residual
        if self.fuse and self.fp8_mode:
            q, s = norm.forward_quant_fp8(x, residual, 128 if self.fp8_mode == "group" else None)
            return QuantizedActivation(q, s, x.dtype), residual
        x, residual = norm(x, residual)
        return x, residual

    def forward(self, positions: torch.Tensor, hidden_states: torch.Tensor, forward_batch,
                residual: Optional[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
        x, residual = self._norm(self.input_layernorm, hidden_states, residual)
        qkv, _ = self.qkv_proj(x)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        if self.fuse:
            self.rotary_emb.forward_with_kv_write(positions, q, k, v, self.attn, forward_batch)
            # decode with a per-token FP8 o_proj: the attention kernel's tail quantises its own
            # output row (same bytes as the stand-alone quantiser), one launch less per layer
            if (self.fp8_mode == "token" and forward_batch.forward_mode.is_decode() and
                    getattr(forward_batch.attn_backend, "fused_tail", False)):
                a = self.attn(q, k, v, forward_batch, save_kv_cache=False, fp8_out=True)
            else:
                a = self.attn(q, k, v, forward_batch, save_kv_cache=False)
        else:
            q, k = self.rotary_emb(positions, q, k)
            a = self.attn(q, k, v, forward_batch)
        o, _ = self.o_proj(a)
        x, residual = self._norm(self.post_attention_layernorm, o, residual)
        gate_up, _ = self.gate_up_proj(x)
        if self.fuse and self.fp8_mode:
            qa, sa = self.act_fn.forward_quant_fp8(gate_up, 128 if self.fp8_mode == "group" else None)
            y = QuantizedActivation(qa, sa, gate_up.dtype)
        else:
            y = self.act_fn(gate_up)
        out, _ = self.down_proj(y)
        return out, residual
