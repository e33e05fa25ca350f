"""Generate tests/golden/attn_*.pt by running the UNMODIFIED reference backend
(python/sglang/srt/layers/attention/torch_native_backend.py, loaded by file path) on CPU.

Run in the authoring container only:  python -m oracle.gen_golden
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os
from types import SimpleNamespace

import torch

from . import ref_loader, scenarios
from .pools import OracleMHATokenToKVPool

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests",
                          "golden")


def run_reference(case, backend, attention_type):
    """Drive TorchNativeAttnBackend.forward_decode / forward_extend exactly as RadixAttention
    does (layers/radix_attention.py:83-109) with the reference-layout pools."""
    layer = scenarios.make_layer(case.hq, case.hkv, case.d, attn_type=attention_type.DECODER)
    pool = OracleMHATokenToKVPool(case.pool_size, 1, case.dtype, case.hkv, case.d, 1)
    pool.k_buffer[0].copy_(case.k_cache)
    pool.v_buffer[0].copy_(case.v_cache)
    fb = SimpleNamespace(
        token_to_kv_pool=pool,
        req_to_token_pool=SimpleNamespace(req_to_token=case.req_to_token),
        req_pool_indices=case.req_pool_indices, seq_lens=case.seq_lens,
        out_cache_loc=case.out_cache_loc, extend_prefix_lens=case.extend_prefix_lens,
        extend_seq_lens=case.extend_seq_lens)
    q = case.q.clone()
    if case.decode:
        o = backend.forward_decode(q, case.k_new.clone(), case.v_new.clone(), layer, fb)
    else:
        o = backend.forward_extend(q, case.k_new.clone(), case.v_new.clone(), layer, fb)
    return o, pool.k_buffer[0], pool.v_buffer[0]


def main():
    torch.set_num_threads(os.cpu_count() or 1)
    backend, attention_type = ref_loader.make_reference_backend()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    for name, kw in {**scenarios.DECODE_CASES, **scenarios.EXTEND_CASES}.items():
        case = scenarios.make_attention_case(**kw)
        o, kb, vb = run_reference(case, backend, attention_type)
        written = case.out_cache_loc
        torch.save({"name": name, "kwargs": {k: (str(v) if isinstance(v, torch.dtype) else v)
                                             for k, v in kw.items()},
                    "input_digest": scenarios.digest(case), "o": o,
                    "k_rows": kb[written].clone(), "v_rows": vb[written].clone(),
                    "source": "reference torch_native_backend.py on CPU, torch " + torch.__version__},
                   os.path.join(GOLDEN_DIR, f"attn_{name}.pt"))
        print(name, tuple(o.shape), o.dtype)


if __name__ == "__main__":
    main()
