"""Loader / wrappers for the reference's own CPU attention kernels built by oracle/build_ref.py
(oracle/_ref/*.so).  TEST INFRASTRUCTURE ONLY: used to cross-check the oracle restatement and as
the timed `cpu_baseline.kind == "reference"` leg of bench.py.

Calling convention follows the reference's CPU backend
(python/sglang/srt/layers/attention/intel_amx_backend.py:109-123, test/srt/cpu/test_decode.py:77-122):
decode_attention_cpu performs the KV write of the new token itself.
"""
from __future__ import annotations

import json
import os

import torch

from .build_ref import OUT, cpu_flags

_ops = None
_variant = None


def available() -> bool:
    return os.path.exists(os.path.join(OUT, "build_info.json"))


def load():
    global _ops, _variant
    if _ops is not None:
        return _ops
    info = json.load(open(os.path.join(OUT, "build_info.json")))
    have = cpu_flags()
    variant = "native" if set(info["native_required_flags"]) <= have else "generic"
    torch.ops.load_library(os.path.join(OUT, f"libsgl_cpu_ref_{variant}.so"))
    _ops = getattr(torch.ops, f"sgl_ref_{variant}")
    _variant = variant
    return _ops


def variant() -> str:
    load()
    return _variant


def decode_attention_ref(q, k_cache, v_cache, k_new, v_new, loc, req_to_token, req_pool_indices,
                         seq_lens, sm_scale, logit_cap=0.0, num_kv_splits=None):
    """q [bs, Hq, D] -> o [bs, Hq, Dv]; writes k_new/v_new into the caches at `loc` first."""
    ops = load()
    if num_kv_splits is None:
        # The reference backend always uses 8 splits (intel_amx_backend.py:30); its kernel returns
        # wrong rows for requests shorter than the split count (empty splits are merged in), so
        # short-sequence parity cases are checked with a single split.
        num_kv_splits = 8 if int(seq_lens.min()) >= 64 else 1
    bs, hq, _ = q.shape
    dv = v_cache.shape[-1]
    o = torch.empty((bs, hq, dv), dtype=q.dtype)
    attn_logits = torch.empty((bs, hq, num_kv_splits, dv + 1), dtype=torch.float32)
    ops.decode_attention_cpu(q.contiguous(), k_cache, v_cache, o, k_new, v_new, loc, attn_logits,
                             req_to_token, req_pool_indices, seq_lens, float(sm_scale),
                             float(logit_cap))
    return o


def fp8_linear_prepack(weight_fp8):
    """One-time VNNI pre-pack of an e4m3 weight [N, K] (convert_weight_packed, gemm.cpp:401-464);
    the reference CPU path does this at load time (is_vnni=True afterwards)."""
    return load().convert_weight_packed(weight_fp8)


def fp8_scaled_mm_ref(x, packed_weight, scales, block=(128, 128), bias=None):
    """The reference's CPU FP8 linear: x [M, K] bf16/f16 x e4m3 weight [N, K] with block scales
    [ceil(N/bn), ceil(K/bk)] fp32 -> [M, N] in x.dtype (fp8_scaled_mm_cpu, gemm_fp8.cpp:480-545:
    W8A16 — the activation stays 16-bit on the CPU path)."""
    return load().fp8_scaled_mm_cpu(x, packed_weight, scales, list(block), bias, x.dtype, True)
