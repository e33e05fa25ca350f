"""Whole-model decode forward over the B200 hot-path ops, for benches and end-to-end tests.

What `bench_one_batch.py` times in the reference (python/sglang/bench_one_batch.py:234-263 `decode`:
`model_runner.forward(batch)` + `logits_processor` + greedy sample) for the Llama / Qwen2 family
(models/llama.py:270-330 LlamaModel.forward, :420-440 LlamaForCausalLM.forward):

    h = embed_tokens(input_ids)   (vocab-parallel + all-reduce, vocab_parallel_embedding.py:480-497)
    for layer in layers: h, residual = layer(positions, h, forward_batch, residual)
    h, _ = norm(h, residual)
    logits = lm_head(h)  (vocab-parallel: each rank a vocab/tp slice, all-gathered,
                          layers/logits_processor.py:470-490)
    next_token_ids = argmax(logits)

This is NOT a model zoo or a model runner: it only strings `B200DecoderLayer`s together on
random-init weights of a named geometry so that BASELINE.json's headline metric (decode tokens/s of
the whole model, TP all-reduce included) can be measured through the hot path.  The scheduler's
per-step work (slot allocation, req_to_token writes) is done by the caller.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Optional

import torch
import torch.nn.functional as F

from ..distributed.parallel_state import (tensor_model_parallel_all_gather,
                                          tensor_model_parallel_all_reduce)
from ..layers.attention.b200_backend import B200AttnBackend
from ..layers.dense_kernel import argmax_rows, gemm_16bit
from ..layers.layernorm import RMSNorm
from ..layers.rotary_embedding import RotaryEmbedding
from ..mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
from .decoder_layer import B200DecoderLayer
from .forward_batch_info import ForwardBatch, ForwardMode

_LM_HEAD_LIBRARY = os.environ.get("B200_LM_HEAD_LIBRARY", "0") == "1"

FP8 = torch.float8_e4m3fn


@dataclass
class ModelGeometry:
    name: str
    hidden: int
    heads: int
    kv_heads: int
    head_dim: int
    inter: int
    layers: int
    vocab: int
    rope_theta: float
    qkv_bias: bool
    rms_eps: float = 1e-5


# public HF configs of the models BASELINE.json names
GEOMETRIES = {
    "llama3.1-8b": ModelGeometry("llama3.1-8b", 4096, 32, 8, 128, 14336, 32, 128256, 500000.0, False),
    "llama3.1-70b": ModelGeometry("llama3.1-70b", 8192, 64, 8, 128, 28672, 80, 128256, 500000.0, False),
    "qwen2.5-32b": ModelGeometry("qwen2.5-32b", 5120, 40, 8, 128, 27648, 64, 152064, 1000000.0, True,
                                 1e-6),
    "qwen2.5-0.5b": ModelGeometry("qwen2.5-0.5b", 896, 14, 2, 64, 4864, 24, 151936, 1000000.0, True,
                                  1e-6),
}


def quant_config(kind: Optional[str]):
    from ..layers.quantization.awq import AWQConfig
    from ..layers.quantization.fp8 import Fp8Config
    from ..layers.quantization.w8a8_fp8 import W8A8Fp8Config

    if kind in (None, "bf16"):
        return None
    if kind == "fp8":            # per-channel weights, dynamic per-token activations (W8A8)
        return W8A8Fp8Config(True)
    if kind == "fp8_block":      # 128x128 weight blocks, 1x128 activation groups
        return Fp8Config(True, "dynamic", weight_block_size=[128, 128])
    if kind == "awq":
        return AWQConfig(4, 128, True)
    raise ValueError(kind)


def fill_random(layer: B200DecoderLayer, seed: int, device) -> None:
    """Random-init weights in their storage types (the `--load-format dummy` convention of the
    reference, model_loader/weight_utils.py:638-663, with N(0, 0.02)-like magnitudes), then
    process_weights_after_loading."""
    g = torch.Generator(device=device).manual_seed(seed)
    for name in ("qkv_proj", "o_proj", "gate_up_proj", "down_proj"):
        lin = getattr(layer, name)
        for pname, p in lin.named_parameters(recurse=False):
            if p.dtype == FP8:
                p.data.copy_((torch.randn(p.shape, device=device, generator=g) * 0.5).to(FP8))
            elif p.dtype == torch.int32:
                p.data.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, p.shape, device=device,
                                           generator=g, dtype=torch.int64).to(torch.int32))
            elif "scale" in pname:
                p.data.copy_(torch.rand(p.shape, device=device, generator=g) * 0.02 + 0.005)
            else:
                p.data.normal_(0, 0.02, generator=g)
        lin.quant_method.process_weights_after_loading(lin)
    for n in ("input_layernorm", "post_attention_layernorm"):
        norm = getattr(layer, n)
        norm.weight.data = torch.ones_like(norm.weight.data).to(torch.bfloat16)


class B200DecodeStep:
    """Pools + backend + decoder layers of one TP rank, and the decode forward over them."""

    def __init__(self, geometry: ModelGeometry, quant: Optional[str], device, tp_rank: int = 0,
                 tp_size: int = 1, max_bs: int = 64, max_ctx: int = 4096, pool_tokens: int = 0,
                 kv_dtype: torch.dtype = torch.bfloat16, layers: int = 0, fuse: bool = True,
                 num_kv_splits: int = 0, seed: int = 1234, fill_kv: bool = True):
        g = geometry
        self.g, self.device, self.tp_rank, self.tp_size = g, device, tp_rank, tp_size
        self.layers_n = layers or g.layers
        self.hq, self.hkv = g.heads // tp_size, max(1, g.kv_heads // tp_size)
        self.kv_dtype = kv_dtype
        self.max_bs, self.max_ctx = max_bs, max_ctx
        pool_tokens = pool_tokens or max_bs * max_ctx
        self.rope = RotaryEmbedding(g.head_dim, g.head_dim, max_ctx + 8, g.rope_theta, True,
                                    torch.bfloat16).to(device)
        self.req_to_token_pool = ReqToTokenPool(max_bs + 1, max_ctx + 8, device)
        self.token_to_kv_pool = MHATokenToKVPool(pool_tokens + max_bs, 1, kv_dtype, self.hkv,
                                                 g.head_dim, self.layers_n, device)
        if fill_kv:
            for l in range(self.layers_n):
                for buf in (self.token_to_kv_pool.k_buffer[l], self.token_to_kv_pool.v_buffer[l]):
                    if kv_dtype == FP8:      # random e4m3 bytes without the NaN patterns
                        buf.random_(0, 256)
                        buf[(buf & 0x7f) == 0x7f] = 0x38
                    else:
                        buf.normal_()
        cfg = SimpleNamespace(num_attention_heads=g.heads, context_len=max_ctx + 4,
                              get_num_kv_heads=lambda t: max(1, g.kv_heads // t))
        self.attn_backend = B200AttnBackend(SimpleNamespace(
            device=device, req_to_token_pool=self.req_to_token_pool,
            token_to_kv_pool=self.token_to_kv_pool, model_config=cfg, tp_size=tp_size,
            server_args=SimpleNamespace(b200_attention_num_kv_splits=num_kv_splits)))
        qc = quant_config(quant)
        self.layers = []
        for l in range(self.layers_n):
            L = B200DecoderLayer(g.hidden, g.heads, g.kv_heads, g.head_dim, g.inter, l, self.rope,
                                 quant_config=qc, params_dtype=torch.bfloat16,
                                 rms_norm_eps=g.rms_eps, qkv_bias=g.qkv_bias, fuse=fuse,
                                 tp_rank=tp_rank, tp_size=tp_size).to(device)
            fill_random(L, seed + 100 + l * 16 + tp_rank, device)
            self.layers.append(L)
        genr = torch.Generator(device=device).manual_seed(seed + 7 + tp_rank)
        self.vocab_per_rank = g.vocab // tp_size
        self.embed = (torch.randn(self.vocab_per_rank, g.hidden, device=device, generator=genr)
                      * 0.02).to(torch.bfloat16)
        self.lm_head = (torch.randn(g.vocab // tp_size, g.hidden, device=device, generator=genr)
                        * 0.02).to(torch.bfloat16)
        self.final_norm = RMSNorm(g.hidden, g.rms_eps).to(device)
        self.final_norm.weight.data = self.final_norm.weight.data.to(torch.bfloat16)

    # ------------------------------------------------------------------ FP8 KV-cache scales
    def load_kv_cache_scales(self, quantization_param_path: str, model_type: Optional[str] = None) -> int:
        """`--quantization-param-path` (model_runner.py:560-569 -> models/llama.py:359-378): per-layer
        k/v scaling factors of an FP8 KV pool from the reference's JSON format.  -> layers set."""
        from ..model_loader.weight_utils import load_kv_cache_scales

        return load_kv_cache_scales([L.attn for L in self.layers], quantization_param_path,
                                    self.tp_rank, self.tp_size, model_type)

    # ------------------------------------------------------------------ byte accounting
    def linear_weight_bytes(self) -> int:
        n = 0
        for L in self.layers:
            for name in ("qkv_proj", "o_proj", "gate_up_proj", "down_proj"):
                for p in getattr(L, name).parameters():
                    n += p.numel() * p.element_size()
        return n

    def kv_bytes_per_token(self) -> int:
        """K + V bytes one cached token occupies on this rank, all layers."""
        return 2 * self.hkv * self.g.head_dim * self.token_to_kv_pool.k_buffer[0].element_size() \
            * self.layers_n

    def step_algorithmic_bytes(self, bs: int, ctx: int) -> int:
        """Bytes one decode step must move on this rank: every linear weight and the lm_head slice
        once, every cached K/V row of the batch once, the slot ids once per layer (activations are
        three orders of magnitude smaller and left out)."""
        return (self.linear_weight_bytes() + self.lm_head.numel() * 2
                + bs * ctx * self.kv_bytes_per_token() + bs * ctx * 4 * self.layers_n)

    # ------------------------------------------------------------------ forward
    def make_forward_batch(self, bs, req_pool_indices, seq_lens, out_cache_loc, seq_lens_sum):
        return ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs,
                            req_pool_indices=req_pool_indices, seq_lens=seq_lens,
                            out_cache_loc=out_cache_loc, seq_lens_sum=seq_lens_sum,
                            req_to_token_pool=self.req_to_token_pool,
                            token_to_kv_pool=self.token_to_kv_pool, attn_backend=self.attn_backend)

    def forward(self, input_ids, positions, forward_batch, out_token_ids) -> None:
        """One decode step; the greedy next-token ids land in `out_token_ids` (int64 [bs])."""
        if self.tp_size > 1:   # VocabParallelEmbedding.forward: masked local lookup, then all-reduce
            local = input_ids - self.tp_rank * self.vocab_per_rank
            outside = (local < 0) | (local >= self.vocab_per_rank)
            h = self.embed[local.masked_fill(outside, 0)]
            h.masked_fill_(outside.unsqueeze(-1), 0)
            h = tensor_model_parallel_all_reduce(h)
        else:
            h = self.embed[input_ids]
        residual = None
        for L in self.layers:
            h, residual = L(positions, h, forward_batch, residual)
        h, _ = self.final_norm(h, residual)
        # in-tree tcgen05 dense GEMM (B200_LM_HEAD_LIBRARY=1: the library matmul, for A/B runs only)
        logits = F.linear(h, self.lm_head) if _LM_HEAD_LIBRARY else gemm_16bit(h, self.lm_head)
        if self.tp_size > 1:
            logits = tensor_model_parallel_all_gather(logits)
        argmax_rows(logits.reshape(-1, logits.shape[-1]), out_token_ids)   # greedy sampling, in-tree kernel

    def forward_extend(self, input_ids, positions, forward_batch, last_index, out_token_ids) -> None:
        """One extend (prefill over a cached radix prefix) step: `input_ids` / `positions` hold the
        NEW tokens of all requests back to back, `forward_batch` is in EXTEND mode
        (extend_seq_lens / extend_prefix_lens / out_cache_loc filled), `last_index` [bs] = row of every
        request's last new token.  Only those rows go through the lm_head, as in
        LogitsProcessor.forward (layers/logits_processor.py:236-260)."""
        if self.tp_size > 1:
            local = input_ids - self.tp_rank * self.vocab_per_rank
            outside = (local < 0) | (local >= self.vocab_per_rank)
            h = self.embed[local.masked_fill(outside, 0)]
            h.masked_fill_(outside.unsqueeze(-1), 0)
            h = tensor_model_parallel_all_reduce(h)
        else:
            h = self.embed[input_ids]
        residual = None
        for L in self.layers:
            h, residual = L(positions, h, forward_batch, residual)
        h, _ = self.final_norm(h, residual)
        logits = gemm_16bit(h[last_index].contiguous(), self.lm_head)
        if self.tp_size > 1:
            logits = tensor_model_parallel_all_gather(logits)
        argmax_rows(logits.reshape(-1, logits.shape[-1]), out_token_ids)   # greedy sampling, in-tree kernel
