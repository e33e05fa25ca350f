"""Test harness: drives the B200 backend (CUDA, through the C-ABI) and the oracle (CPU) on the
same scenario.  Mirrors how ModelRunner / RadixAttention call a backend in the reference
(model_runner.py:1139-1172, radix_attention.py:83-109)."""
from __future__ import annotations

from types import SimpleNamespace

import torch

from oracle import attention as oracle_attention
from oracle import scenarios
from oracle.pools import OracleMHATokenToKVPool


def oracle_run(case, logit_cap=0.0):
    """CPU oracle: write new K/V into the cache (set_kv_buffer), then attention."""
    layer = scenarios.make_layer(case.hq, case.hkv, case.d, logit_cap=logit_cap)
    pool = OracleMHATokenToKVPool(case.pool_size, 1, case.dtype, case.hkv, case.d, 1)
    pool.k_buffer[0].copy_(case.k_cache)
    pool.v_buffer[0].copy_(case.v_cache)
    pool.set_kv_buffer(layer, case.out_cache_loc, case.k_new.clone(), case.v_new.clone())
    q = case.q.view(-1, case.hq, case.d)
    if case.decode:
        o = oracle_attention.decode_attention_oracle(
            q, pool.k_buffer[0], pool.v_buffer[0], case.req_to_token, case.req_pool_indices,
            case.seq_lens, layer.scaling, logit_cap)
    else:
        o = oracle_attention.extend_attention_oracle(
            q, pool.k_buffer[0], pool.v_buffer[0], case.req_to_token, case.req_pool_indices,
            case.seq_lens, case.extend_prefix_lens, case.extend_seq_lens, layer.scaling, True,
            logit_cap)
    return o.reshape(o.shape[0], -1), pool.k_buffer[0], pool.v_buffer[0]


def make_model_runner(case, device="cuda", layer_num=1, num_kv_splits=0):
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool

    rtp = ReqToTokenPool(case.req_to_token.shape[0], case.req_to_token.shape[1], device)
    rtp.req_to_token.copy_(case.req_to_token)
    kvp = MHATokenToKVPool(case.pool_size, 1, case.dtype, case.hkv, case.d, layer_num, device)
    for l in range(layer_num):
        kvp.k_buffer[l].copy_(case.k_cache)
        kvp.v_buffer[l].copy_(case.v_cache)
    cfg = SimpleNamespace(num_attention_heads=case.hq, context_len=case.req_to_token.shape[1] - 4,
                          get_num_kv_heads=lambda tp: max(1, case.hkv // tp))
    return SimpleNamespace(device=device, req_to_token_pool=rtp, token_to_kv_pool=kvp,
                           model_config=cfg, tp_size=1, sliding_window_size=None,
                           server_args=SimpleNamespace(b200_attention_num_kv_splits=num_kv_splits))


def b200_run(case, logit_cap=0.0, num_kv_splits=0, strided_qkv=True, index_dtype=None,
             fused_tail=None, fp8_out=False, repeat=1, tail_in_kernel=None):
    """CUDA path: B200AttnBackend through RadixAttention.forward, KV write included."""
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.radix_attention import RadixAttention
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    dev = "cuda"
    mr = make_model_runner(case, dev, num_kv_splits=num_kv_splits)
    backend = B200AttnBackend(mr)
    if fused_tail is not None:
        backend.fused_tail = fused_tail
    if tail_in_kernel is not None:   # force the in-kernel (ticket) tail or the one-launch tail
        backend.tail_in_kernel_max_bs = 1 << 30 if tail_in_kernel else 0
    layer = RadixAttention(case.hq, case.d, float(case.d) ** -0.5, case.hkv, 0, logit_cap=logit_cap)
    t = case.q.shape[0]
    if strided_qkv:
        # q, k, v as slices of one fused qkv activation (models/llama.py:186-189): row stride
        # (Hq + 2 Hkv) * D, so k/v rows are NOT contiguous across tokens.
        qkv = torch.cat([case.q, case.k_new.reshape(t, -1), case.v_new.reshape(t, -1)], dim=1).to(dev)
        q, k, v = qkv.split([case.hq * case.d, case.hkv * case.d, case.hkv * case.d], dim=-1)
    else:
        q, k, v = case.q.to(dev), case.k_new.to(dev).reshape(t, -1), case.v_new.to(dev).reshape(t, -1)
    fb = ForwardBatch(
        forward_mode=ForwardMode.DECODE if case.decode else ForwardMode.EXTEND,
        batch_size=case.bs,
        # index_dtype=torch.int32 is what the reference's CudaGraphRunner hands the backend
        # (cuda_graph_runner.py:250-253); the default keeps the ForwardBatch's int64
        req_pool_indices=case.req_pool_indices.to(dev).to(index_dtype or torch.int64),
        seq_lens=case.seq_lens.to(dev).to(index_dtype or torch.int64),
        out_cache_loc=case.out_cache_loc.to(dev),
        seq_lens_sum=int(case.seq_lens.sum()), seq_lens_cpu=case.seq_lens,
        extend_seq_lens=case.extend_seq_lens.to(dev),
        extend_prefix_lens=case.extend_prefix_lens.to(dev),
        extend_seq_lens_cpu=case.extend_seq_lens.tolist(),
        extend_prefix_lens_cpu=case.extend_prefix_lens.tolist(),
        extend_num_tokens=t, req_to_token_pool=mr.req_to_token_pool,
        token_to_kv_pool=mr.token_to_kv_pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    kw = {"fp8_out": True} if fp8_out else {}
    for _ in range(repeat):  # repeats re-use the backend's ticket buffer (self-resetting counters)
        o = layer(q, k, v, fb, **kw)
    torch.cuda.synchronize()
    if fp8_out:
        return (o.q.cpu(), o.scale.cpu()), mr.token_to_kv_pool.k_buffer[0].cpu(), \
            mr.token_to_kv_pool.v_buffer[0].cpu()
    return o.cpu(), mr.token_to_kv_pool.k_buffer[0].cpu(), mr.token_to_kv_pool.v_buffer[0].cpu()


def error_stats(a, b):
    a, b = a.float(), b.float()
    diff = (a - b).abs()
    denom = b.abs().clamp_min(1e-6)
    return dict(max_abs=diff.max().item(),
                rel_l2=(diff.norm() / b.norm().clamp_min(1e-12)).item(),
                max_rel_at_large=(diff / denom)[b.abs() > 0.05 * b.abs().max()].max().item()
                if (b.abs() > 0.05 * b.abs().max()).any() else 0.0)
