"""CPU restatement of the reference attention semantics (torch_native backend).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows python/sglang/srt/layers/attention/torch_native_backend.py:
  decode  (:112-180, 226-267): per request gather K/V rows `req_to_token[req, :seq_len]` from the
          cache (the new token's K/V has already been written, :244-247), softmax(q.K^T*scale)@V.
  extend  (:27-110, 182-224): per request, the new tokens attend causally to all seq_len cached
          keys (prefix + themselves); only rows [prefix_len:] are produced.
The reference calls SDPA on bf16 tensors; here the math is spelled out in fp32 (softmax in fp32,
single rounding of the output to the input dtype), which is what SDPA's math does up to the
rounding of intermediate products.  `logit_cap` (tanh capping) is NOT part of torch_native; it is
restated from triton_ops/decode_attention.py:110-111 for the optional capped path.
"""
from __future__ import annotations

import torch


def _attend(q, k, v, scale, causal_offset=None, logit_cap=0.0):
    """q [Hq, Lq, D], k [Hkv, Lk, D], v [Hkv, Lk, Dv] fp32 -> [Hq, Lq, Dv] fp32.
    causal_offset = number of keys every query row may see beyond its own index (prefix_len)."""
    hq, hkv = q.shape[0], k.shape[0]
    if hq != hkv:  # GQA: query head h uses kv head h // group
        k = k.repeat_interleave(hq // hkv, dim=0)
        v = v.repeat_interleave(hq // hkv, dim=0)
    s = torch.matmul(q, k.transpose(1, 2)) * scale
    if logit_cap and logit_cap > 0:
        s = logit_cap * torch.tanh(s / logit_cap)
    if causal_offset is not None:
        lq, lk = q.shape[1], k.shape[1]
        qi = torch.arange(lq).unsqueeze(1) + causal_offset
        kj = torch.arange(lk).unsqueeze(0)
        s = s.masked_fill(kj > qi, float("-inf"))
    return torch.matmul(torch.softmax(s, dim=-1), v)


def decode_attention_oracle(q, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens, scale,
                            logit_cap=0.0, k_scale=1.0, v_scale=1.0):
    """q [bs, Hq, D] -> o [bs, Hq, Dv] in q.dtype.
    FP8 KV pool (k_cache/v_cache of dtype float8_e4m3fn): the stored values are de-quantised as
    value * k_scale / value * v_scale — the FlashInfer-backend semantics of the reference
    (flashinfer_backend.py:499-509; the Triton backend is the k_scale = v_scale = 1 case,
    triton_backend.py:808-833); torch_native itself cannot run SDPA on fp8 tensors, so this is the
    unmodified torch_native math applied to the de-quantised cache."""
    bs, hq, _ = q.shape
    dv = v_cache.shape[-1]
    o = torch.empty((bs, hq, dv), dtype=q.dtype)
    for b in range(bs):
        n = int(seq_lens[b])
        tok = req_to_token[int(req_pool_indices[b]), :n].long()
        k = k_cache[tok].float().transpose(0, 1) * k_scale   # [Hkv, n, D]
        v = v_cache[tok].float().transpose(0, 1) * v_scale
        out = _attend(q[b].float().unsqueeze(1), k, v, scale, None, logit_cap)  # [Hq, 1, Dv]
        o[b] = out.squeeze(1).to(q.dtype)
    return o


def extend_attention_oracle(q, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens,
                            extend_prefix_lens, extend_seq_lens, scale, causal=True,
                            logit_cap=0.0, k_scale=1.0, v_scale=1.0):
    """q [T, Hq, D] (new tokens of all requests, concatenated) -> o [T, Hq, Dv].
    k_scale / v_scale: de-quantisation scales of an FP8 pool (see decode_attention_oracle)."""
    t, hq, _ = q.shape
    dv = v_cache.shape[-1]
    o = torch.empty((t, hq, dv), dtype=q.dtype)
    start = 0
    for b in range(len(seq_lens)):
        n, pre, ext = int(seq_lens[b]), int(extend_prefix_lens[b]), int(extend_seq_lens[b])
        tok = req_to_token[int(req_pool_indices[b]), :n].long()
        k = k_cache[tok].float().transpose(0, 1) * k_scale
        v = v_cache[tok].float().transpose(0, 1) * v_scale
        qb = q[start:start + ext].float().transpose(0, 1)  # [Hq, ext, D]
        out = _attend(qb, k, v, scale, pre if causal else None, logit_cap)
        o[start:start + ext] = out.transpose(0, 1).to(q.dtype)
        start += ext
    return o
