"""Deterministic synthetic inputs shared by the tests and the golden-vector generator.

TEST INFRASTRUCTURE ONLY.  Inputs come from numpy's MT19937 (stable across library versions) so
fixtures only need to store reference OUTPUTS plus a digest of the inputs.
"""
from __future__ import annotations

import hashlib
from types import SimpleNamespace

import numpy as np
import torch


def _randn(rs, shape, dtype, mean=0.0, std=1.0):
    x = rs.standard_normal(size=shape).astype(np.float32) * std + mean
    return torch.from_numpy(x).to(dtype)


def make_layer(hq, hkv, d, layer_id=0, logit_cap=0.0, dv=None, attn_type=None):
    """Duck-typed RadixAttention attributes (layers/radix_attention.py:60-81)."""
    return SimpleNamespace(tp_q_head_num=hq, tp_k_head_num=hkv, tp_v_head_num=hkv, head_dim=d,
                           qk_head_dim=d, v_head_dim=dv or d, scaling=float(d) ** -0.5,
                           layer_id=layer_id, logit_cap=logit_cap, sliding_window_size=-1,
                           is_cross_attention=False, attn_type=attn_type, k_scale=None,
                           v_scale=None)


def make_attention_case(seed, hq, hkv, d, seq_lens, extend_lens=None, pool_size=None,
                        dtype=torch.bfloat16, shared_prefix=0, max_reqs=None, contiguous=False,
                        mean=0.0, std=1.0):
    """Builds one attention problem on CPU.

    decode  (extend_lens is None): every request has seq_lens[b] tokens in total, the last one is
            the new token (its K/V arrives in k_new/v_new and goes to slot out_cache_loc[b]).
    extend : the last extend_lens[b] tokens of request b are new; the first seq_lens[b] -
            extend_lens[b] are the cached radix prefix.  `shared_prefix` > 0 makes the first
            min(shared_prefix, prefix_b) slots of every request the SAME slot ids (what
            ScheduleBatch.prepare_for_extend does with prefix_indices, schedule_batch.py:1155-1158).
    """
    rs = np.random.RandomState(seed)
    bs = len(seq_lens)
    seq_lens = [int(x) for x in seq_lens]
    decode = extend_lens is None
    ext = [1] * bs if decode else [int(x) for x in extend_lens]
    pre = [s - e for s, e in zip(seq_lens, ext)]
    assert all(p >= 0 for p in pre)
    total_slots = sum(seq_lens)
    pool_size = pool_size or (total_slots + 7)
    max_reqs = max_reqs or (bs + 3)
    max_ctx = max(seq_lens) + 4
    order = np.arange(1, pool_size + 1) if contiguous else rs.permutation(pool_size) + 1
    req_to_token = torch.zeros((max_reqs, max_ctx), dtype=torch.int32)
    req_pool_indices = torch.from_numpy(rs.permutation(max_reqs)[:bs].astype(np.int64))
    cursor = 0
    shared = None
    if shared_prefix > 0:
        shared = order[:shared_prefix]
        cursor = shared_prefix
    out_loc = []
    for b in range(bs):
        n_shared = min(shared_prefix, pre[b]) if shared is not None else 0
        own = seq_lens[b] - n_shared
        slots = np.concatenate([shared[:n_shared], order[cursor:cursor + own]]) if n_shared else \
            order[cursor:cursor + own]
        cursor += own
        assert cursor <= pool_size, "pool too small for this case"
        req_to_token[req_pool_indices[b], :seq_lens[b]] = torch.from_numpy(slots.astype(np.int32))
        out_loc.append(slots[pre[b]:])
    out_cache_loc = torch.from_numpy(np.concatenate(out_loc).astype(np.int64))
    t = sum(ext)
    case = SimpleNamespace(
        hq=hq, hkv=hkv, d=d, dtype=dtype, bs=bs, pool_size=pool_size, decode=decode,
        q=_randn(rs, (t, hq * d), dtype, mean, std),
        k_new=_randn(rs, (t, hkv, d), dtype, mean, std),
        v_new=_randn(rs, (t, hkv, d), dtype, mean, std),
        k_cache=_randn(rs, (pool_size + 1, hkv, d), dtype, mean, std),
        v_cache=_randn(rs, (pool_size + 1, hkv, d), dtype, mean, std),
        req_to_token=req_to_token, req_pool_indices=req_pool_indices,
        seq_lens=torch.tensor(seq_lens, dtype=torch.int64),
        out_cache_loc=out_cache_loc,
        extend_prefix_lens=torch.tensor(pre, dtype=torch.int32),
        extend_seq_lens=torch.tensor(ext, dtype=torch.int32),
    )
    # slot 0 is the zero-initialised dummy row in the reference pool
    case.k_cache[0].zero_()
    case.v_cache[0].zero_()
    return case


def digest(case) -> str:
    h = hashlib.sha256()
    for name in ["q", "k_new", "v_new", "k_cache", "v_cache", "req_to_token", "req_pool_indices",
                 "seq_lens", "out_cache_loc"]:
        t = getattr(case, name).contiguous()
        h.update(t.view(torch.uint8).numpy().tobytes())
    return h.hexdigest()


# The parity cases: name -> kwargs of make_attention_case.  Shapes cover the geometries of the
# BASELINE configs (Llama-3.1-8B group 4 / D 128, Qwen2.5-0.5B group 7 / D 64, 70B@TP8 group 8,
# MHA) with ragged lengths around the kernel's tile sizes (16/32/64) and a shared radix prefix.
DECODE_CASES = {
    "decode_llama_g4_d128": dict(seed=11, hq=8, hkv=2, d=128, seq_lens=[1, 17, 64, 130, 257]),
    "decode_qwen_g7_d64": dict(seed=12, hq=14, hkv=2, d=64, seq_lens=[513, 3, 96, 1]),
    "decode_g8_d128": dict(seed=13, hq=8, hkv=1, d=128, seq_lens=[200, 33, 65]),
    "decode_mha_f16": dict(seed=14, hq=4, hkv=4, d=128, seq_lens=[31, 129], dtype=torch.float16),
    "decode_shared_prefix": dict(seed=15, hq=8, hkv=2, d=128, seq_lens=[140, 150, 129],
                                 shared_prefix=128),
}
EXTEND_CASES = {
    "extend_llama_g4_d128": dict(seed=21, hq=4, hkv=1, d=128, seq_lens=[40, 40, 193],
                                 extend_lens=[40, 7, 65]),
    "extend_qwen_g7_d64": dict(seed=22, hq=14, hkv=2, d=64, seq_lens=[70, 130],
                               extend_lens=[6, 130]),
    "extend_shared_prefix": dict(seed=23, hq=8, hkv=2, d=128, seq_lens=[80, 72, 65],
                                 extend_lens=[16, 8, 1], shared_prefix=64),
}
