"""Test-only decoder stack (Llama/Qwen2 layer structure, models/llama.py:94-98,180-191,245-267)
used to check greedy token streams end to end.  Two executions of the SAME weights:

  mode "b200"   : GPU.  sglang_security_b200.model_executor.decoder_layer.B200DecoderLayer:
                  attention through RadixAttention -> B200AttnBackend, linear layers through the
                  configured quant method, RMSNorm / RoPE(+KV write) / SiLU*mul through the fused
                  C-ABI kernels (`fuse=False` launches every op separately in the reference order).
                  Embedding lookup and lm_head are plain torch ops on the device.
  mode "oracle" : CPU.  Attention through oracle/attention.py over the oracle pools, linear layers
                  through torch (or the oracle's quantised-linear restatements), RMSNorm / RoPE /
                  SiLU*mul through oracle/elementwise.py (the reference's forward_native).
The host-side bookkeeping (slot allocation, req_to_token writes, ForwardBatch fields) emulates
ScheduleBatch.prepare_for_extend / prepare_for_decode (schedule_batch.py:1117-1291,1464-1527).
"""
from __future__ import annotations

from dataclasses import dataclass
from types import SimpleNamespace
from typing import List, Optional

import torch
import torch.nn.functional as F


@dataclass
class StackConfig:
    hidden: int = 896
    heads: int = 14
    kv_heads: int = 2
    head_dim: int = 64
    inter: int = 4864
    layers: int = 24
    vocab: int = 32000
    rope_theta: float = 1e6
    eps: float = 1e-6
    qkv_bias: bool = True          # Qwen2 has qkv bias (models/qwen2.py:138)
    quant: Optional[str] = None    # None | "fp8" | "fp8_block"


def init_weights(cfg: StackConfig, seed: int = 0, std: float = 0.02, decisive_embed_std: float = 0.0):
    """N(0, std) weights.  With `decisive_embed_std` > 0 the head is TIED to the embedding through a
    fixed random permutation (lm_head[v] = embed[perm[v]]; Qwen2.5-0.5B ties its head, HF config
    tie_word_embeddings) and the embedding gets that std, so the residual stream carries a token
    signal as large as the 24 layers' contributions: the greedy token then wins by a margin far
    above the bf16 / fp8 numeric noise instead of being a near-tie among 32000 random logits
    (SURVEY §8d cfg 1: random-init logits are nearly tied, report the margin)."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: (torch.randn(*s, generator=g) * std).to(torch.bfloat16)
    qkv_out = (cfg.heads + 2 * cfg.kv_heads) * cfg.head_dim
    w = {"embed": r(cfg.vocab, cfg.hidden), "lm_head": r(cfg.vocab, cfg.hidden),
         "final_norm": torch.ones(cfg.hidden, dtype=torch.bfloat16), "layers": []}
    if decisive_embed_std > 0:
        w["embed"] = (torch.randn(cfg.vocab, cfg.hidden, generator=g) * decisive_embed_std).to(
            torch.bfloat16)
        w["lm_head"] = w["embed"][torch.randperm(cfg.vocab, generator=g)].contiguous()
    for _ in range(cfg.layers):
        w["layers"].append(dict(
            in_norm=(1 + 0.1 * torch.randn(cfg.hidden, generator=g)).to(torch.bfloat16),
            post_norm=(1 + 0.1 * torch.randn(cfg.hidden, generator=g)).to(torch.bfloat16),
            qkv=r(qkv_out, cfg.hidden), qkv_bias=r(qkv_out) if cfg.qkv_bias else None,
            o=r(cfg.hidden, cfg.heads * cfg.head_dim), gate_up=r(2 * cfg.inter, cfg.hidden),
            down=r(cfg.hidden, cfg.inter)))
    return w


def rmsnorm(x, w, eps):
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * w


def rope(positions, q, k, head_dim, theta):
    """neox-style rotary embedding on [T, H*D] tensors (layers/rotary_embedding.py)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=q.device).float() / head_dim))
    ang = positions.float().unsqueeze(-1) * inv
    cos, sin = ang.cos().unsqueeze(1), ang.sin().unsqueeze(1)

    def rot(x):
        t = x.view(x.shape[0], -1, head_dim).float()
        x1, x2 = t[..., : head_dim // 2], t[..., head_dim // 2:]
        return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], -1).to(x.dtype).view(x.shape)

    return rot(q), rot(k)


class Stack:
    def __init__(self, cfg: StackConfig, weights, mode: str, pool_size: int, max_reqs: int,
                 max_ctx: int, fuse: bool = True):
        self.cfg, self.mode = cfg, mode
        self.max_pos = max_ctx + 8
        self.dev = "cuda" if mode == "b200" else "cpu"
        self.w = weights
        self.next_slot = 1
        self.scaling = cfg.head_dim ** -0.5
        if mode == "b200":
            from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
            from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding
            from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
            from sglang_security_b200.model_executor.decoder_layer import B200DecoderLayer

            self.rtp = ReqToTokenPool(max_reqs, max_ctx, self.dev)
            self.kvp = MHATokenToKVPool(pool_size, 1, torch.bfloat16, cfg.kv_heads, cfg.head_dim,
                                        cfg.layers, self.dev)
            mc = SimpleNamespace(num_attention_heads=cfg.heads, context_len=max_ctx - 4,
                                 get_num_kv_heads=lambda tp: cfg.kv_heads)
            self.backend = B200AttnBackend(SimpleNamespace(
                device=self.dev, req_to_token_pool=self.rtp, token_to_kv_pool=self.kvp,
                model_config=mc, tp_size=1, server_args=SimpleNamespace()))
            qc = self._quant_config()
            # native_rounding: bit-exact with the oracle's (= forward_native's) 16-bit RoPE arithmetic
            rope_mod = RotaryEmbedding(cfg.head_dim, cfg.head_dim, self.max_pos, cfg.rope_theta, True,
                                       torch.bfloat16, native_rounding=True).to(self.dev)
            self.layers = []
            for l, lw in enumerate(weights["layers"]):
                L = B200DecoderLayer(cfg.hidden, cfg.heads, cfg.kv_heads, cfg.head_dim, cfg.inter, l,
                                     rope_mod, quant_config=qc, params_dtype=torch.bfloat16,
                                     rms_norm_eps=cfg.eps, qkv_bias=cfg.qkv_bias, fuse=fuse,
                                     tp_rank=0, tp_size=1).to(self.dev)
                for mod, key in ((L.qkv_proj, "qkv"), (L.o_proj, "o"), (L.gate_up_proj, "gate_up"),
                                 (L.down_proj, "down")):
                    self._load(mod, lw[key])
                if cfg.qkv_bias:
                    L.qkv_proj.bias.data.copy_(lw["qkv_bias"])
                L.input_layernorm.weight.data = lw["in_norm"].to(self.dev)
                L.post_attention_layernorm.weight.data = lw["post_norm"].to(self.dev)
                self.layers.append(L)
            self.embed = weights["embed"].to(self.dev)
            self.lm_head = weights["lm_head"].to(self.dev)
            self.final_norm = weights["final_norm"].to(self.dev)
        else:
            from oracle.pools import OracleMHATokenToKVPool, OracleReqToTokenPool

            self.rtp = OracleReqToTokenPool(max_reqs, max_ctx)
            self.kvp = OracleMHATokenToKVPool(pool_size, 1, torch.bfloat16, cfg.kv_heads,
                                              cfg.head_dim, cfg.layers)
            self._oracle_q = {}

    # ---------------------------------------------------------------- quantisation plumbing
    def _quant_config(self):
        if self.cfg.quant is None:
            return None
        from sglang_security_b200.layers.quantization.fp8 import Fp8Config

        if self.cfg.quant == "fp8":
            return Fp8Config(False, "dynamic")          # 16-bit checkpoint, per-channel at load
        raise ValueError(self.cfg.quant)

    def _load(self, mod, w):
        mod.weight.data.copy_(w)
        mod.quant_method.process_weights_after_loading(mod)

    def _oracle_linear(self, name, x, w, bias=None):
        if self.cfg.quant is None:
            return F.linear(x, w, bias)
        from oracle import fp8 as OF

        if name not in self._oracle_q:
            self._oracle_q[name] = OF.per_token_group_quant_fp8_oracle(w, w.shape[-1])
        wq, ws = self._oracle_q[name]
        return OF.fp8_linear_oracle(x, wq, ws, bias)

    # ---------------------------------------------------------------- host bookkeeping
    def alloc(self, n: int) -> torch.Tensor:
        s = torch.arange(self.next_slot, self.next_slot + n, dtype=torch.int64)
        self.next_slot += n
        return s

    # ---------------------------------------------------------------- one forward pass
    def forward(self, input_ids: torch.Tensor, positions: torch.Tensor, fb_fields: dict):
        """fb_fields: decode(bool), req_pool_indices, seq_lens, out_cache_loc, prefix_lens,
        extend_lens (CPU tensors).  Returns logits [bs, vocab] of each request's LAST token."""
        cfg, dev = self.cfg, self.dev
        decode = fb_fields["decode"]
        ext = fb_fields["extend_lens"]
        last_idx = torch.cumsum(ext, 0) - 1
        if self.mode == "b200":
            from sglang_security_b200.model_executor.forward_batch_info import (ForwardBatch,
                                                                                ForwardMode)

            fb = ForwardBatch(
                forward_mode=ForwardMode.DECODE if decode else ForwardMode.EXTEND,
                batch_size=len(ext), req_pool_indices=fb_fields["req_pool_indices"].to(dev),
                seq_lens=fb_fields["seq_lens"].to(dev),
                out_cache_loc=fb_fields["out_cache_loc"].to(dev),
                seq_lens_sum=int(fb_fields["seq_lens"].sum()),
                extend_seq_lens=ext.to(torch.int32).to(dev),
                extend_prefix_lens=fb_fields["prefix_lens"].to(torch.int32).to(dev),
                extend_seq_lens_cpu=ext.tolist(),
                extend_prefix_lens_cpu=fb_fields["prefix_lens"].tolist(),
                req_to_token_pool=self.rtp, token_to_kv_pool=self.kvp, attn_backend=self.backend)
            self.backend.init_forward_metadata(fb)
            h = self.embed[input_ids.to(dev)]
            pos = positions.to(dev)
            residual = None
            for L in self.layers:
                h, residual = L(pos, h, fb, residual)
            h = rmsnorm((h + residual)[last_idx.to(dev)], self.final_norm, cfg.eps)
            return F.linear(h, self.lm_head).float().cpu()
        # ---- oracle (CPU)
        from oracle import attention as OA
        from oracle import elementwise as OE

        w = self.w
        h = w["embed"][input_ids]
        if not hasattr(self, "_cos_sin"):
            self._cos_sin = OE.cos_sin_cache_oracle(cfg.head_dim, self.max_pos, cfg.rope_theta)
        layer_ns = SimpleNamespace(layer_id=0)
        residual = None
        for l, lw in enumerate(w["layers"]):
            # LlamaDecoderLayer.forward (models/llama.py:245-267) over the forward_native ops
            if residual is None:
                residual, x = h, OE.rmsnorm_oracle(h, lw["in_norm"], cfg.eps)
            else:
                x, residual = OE.rmsnorm_oracle(h, lw["in_norm"], cfg.eps, residual)
            qkv = self._oracle_linear(f"qkv{l}", x, lw["qkv"], lw["qkv_bias"])
            q, k, v = qkv.split([cfg.heads * cfg.head_dim, cfg.kv_heads * cfg.head_dim,
                                 cfg.kv_heads * cfg.head_dim], dim=-1)
            q, k = OE.rope_oracle(positions, q.contiguous(), k.contiguous(), self._cos_sin,
                                  cfg.head_dim, cfg.head_dim, True)
            layer_ns.layer_id = l
            self.kvp.set_kv_buffer(layer_ns, fb_fields["out_cache_loc"],
                                   k.reshape(-1, cfg.kv_heads, cfg.head_dim).contiguous(),
                                   v.reshape(-1, cfg.kv_heads, cfg.head_dim).contiguous())
            qh = q.reshape(-1, cfg.heads, cfg.head_dim)
            if decode:
                a = OA.decode_attention_oracle(qh, self.kvp.k_buffer[l], self.kvp.v_buffer[l],
                                               self.rtp.req_to_token, fb_fields["req_pool_indices"],
                                               fb_fields["seq_lens"], self.scaling)
            else:
                a = OA.extend_attention_oracle(qh, self.kvp.k_buffer[l], self.kvp.v_buffer[l],
                                               self.rtp.req_to_token, fb_fields["req_pool_indices"],
                                               fb_fields["seq_lens"], fb_fields["prefix_lens"], ext,
                                               self.scaling)
            o = self._oracle_linear(f"o{l}", a.reshape(a.shape[0], -1), lw["o"])
            x, residual = OE.rmsnorm_oracle(o, lw["post_norm"], cfg.eps, residual)
            gu = self._oracle_linear(f"gu{l}", x, lw["gate_up"])
            h = self._oracle_linear(f"dn{l}", OE.silu_and_mul_oracle(gu), lw["down"])
        h = rmsnorm((h + residual)[last_idx], w["final_norm"], cfg.eps)
        return F.linear(h, w["lm_head"]).float()


def run_shared_prefix_session(stack: Stack, prefix_ids, suffix_ids: List[torch.Tensor],
                              decode_steps: int, forced_tokens=None):
    """cfg 1: (A) extend one request with the shared prefix, (B) extend bs requests
    that reuse the SAME prefix slots plus their own suffix, (C) greedy decode.  If forced_tokens
    is given the decode inputs are taken from it (teacher forcing) instead of this run's argmax.
    Returns (tokens [steps+1, bs], logits list)."""
    bs, p = len(suffix_ids), len(prefix_ids)
    # (A) prefix request -> request slot 0
    slots = stack.alloc(p)
    stack.rtp.req_to_token[0, :p] = slots.to(torch.int32).to(stack.rtp.req_to_token.device)
    stack.forward(prefix_ids, torch.arange(p), dict(
        decode=False, req_pool_indices=torch.tensor([0]), seq_lens=torch.tensor([p]),
        out_cache_loc=slots, prefix_lens=torch.tensor([0]), extend_lens=torch.tensor([p])))
    # (B) bs requests sharing the prefix slots (prefix_indices copy, schedule_batch.py:1155-1158)
    req_idx = torch.arange(1, bs + 1)
    ext = torch.tensor([len(s) for s in suffix_ids])
    locs, pos = [], []
    for i, s in enumerate(suffix_ids):
        own = stack.alloc(len(s))
        row = torch.cat([slots, own]).to(torch.int32)
        stack.rtp.req_to_token[i + 1, :len(row)] = row.to(stack.rtp.req_to_token.device)
        locs.append(own)
        pos.append(torch.arange(p, p + len(s)))
    seq_lens = ext + p
    logits = stack.forward(torch.cat(suffix_ids), torch.cat(pos), dict(
        decode=False, req_pool_indices=req_idx, seq_lens=seq_lens, out_cache_loc=torch.cat(locs),
        prefix_lens=torch.full((bs,), p), extend_lens=ext))
    all_logits, tokens = [logits], [logits.argmax(-1)]
    # (C) greedy decode
    for step in range(decode_steps):
        inp = tokens[-1] if forced_tokens is None else forced_tokens[step]
        new = stack.alloc(bs)
        for i in range(bs):
            stack.rtp.req_to_token[i + 1, int(seq_lens[i])] = int(new[i])
        positions = seq_lens.clone()
        seq_lens = seq_lens + 1
        logits = stack.forward(inp, positions, dict(
            decode=True, req_pool_indices=req_idx, seq_lens=seq_lens, out_cache_loc=new,
            prefix_lens=seq_lens - 1, extend_lens=torch.ones(bs, dtype=torch.int64)))
        all_logits.append(logits)
        tokens.append(logits.argmax(-1))
    return torch.stack(tokens), all_logits
