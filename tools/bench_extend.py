#!/usr/bin/env python
"""Extend (prefill with a cached radix prefix) attention throughput on the tcgen05 kernel:
TFLOP/s = 4 * Hq * D * sum_b ext_b * (prefix_b + (ext_b + 1) / 2) / time  (SURVEY section 8d cfg 5),
against the measured dense bf16 peak.  Shapes: Llama-3.1-8B heads (Hq 32, Hkv 8, D 128) on one GPU
and the per-rank heads of Llama-3.1-70B at TP=8 (Hq 8, Hkv 1), radix hit-rate sweep
h = prefix / ctx in {0, 0.5, 0.9}.  Times the kernel with CUDA events through the library timing
registry (KV write excluded), L2 flushed between launches."""
import argparse
import json
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bench_extend.json"))
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--ab", action="store_true", help="also time the round-1 softmax instance")
    ap.add_argument("--headline", action="store_true", help="Llama-3.1-8B heads only (6 points)")
    args = ap.parse_args()
    from sglang_security_b200 import _lib
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.radix_attention import RadixAttention
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    dev = "cuda"
    peak = 1662.9
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    except Exception:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    shapes = [("llama8b_tp1", 32, 8, 128), ("llama70b_tp8_rank", 8, 1, 128)]
    if args.headline:
        shapes = shapes[:1]
    ctxs = [(8, 2048), (4, 8192)] if not args.quick else [(4, 2048)]
    for name, hq, hkv, d in shapes:
        for bs, ctx in ctxs:
            for hit in (0.0, 0.5, 0.9):
                prefix = int(ctx * hit)
                ext = ctx - prefix
                total = bs * ctx
                rtp = ReqToTokenPool(bs + 1, ctx + 4, dev)
                kvp = MHATokenToKVPool(total + 8, 1, torch.bfloat16, hkv, d, 1, dev)
                kvp.k_buffer[0].normal_()
                kvp.v_buffer[0].normal_()
                g = torch.Generator().manual_seed(0)
                slots = (torch.randperm(total, generator=g) + 1).to(torch.int32).view(bs, ctx)
                rtp.req_to_token[:bs, :ctx] = slots.to(dev)
                cfg = SimpleNamespace(num_attention_heads=hq, context_len=ctx,
                                      get_num_kv_heads=lambda t: hkv)
                backend = B200AttnBackend(SimpleNamespace(
                    device=dev, req_to_token_pool=rtp, token_to_kv_pool=kvp, model_config=cfg,
                    tp_size=1, server_args=SimpleNamespace()))
                layer = RadixAttention(hq, d, d ** -0.5, hkv, 0)
                t = bs * ext
                qkv = torch.randn(t, (hq + 2 * hkv) * d, device=dev).to(torch.bfloat16)
                q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
                fb = ForwardBatch(
                    forward_mode=ForwardMode.EXTEND, batch_size=bs,
                    req_pool_indices=torch.arange(bs, dtype=torch.int64, device=dev),
                    seq_lens=torch.full((bs,), ctx, dtype=torch.int64, device=dev),
                    out_cache_loc=slots[:, prefix:].reshape(-1).to(torch.int64).to(dev),
                    seq_lens_sum=total,
                    extend_seq_lens=torch.full((bs,), ext, dtype=torch.int32, device=dev),
                    extend_prefix_lens=torch.full((bs,), prefix, dtype=torch.int32, device=dev),
                    extend_seq_lens_cpu=[ext] * bs, extend_prefix_lens_cpu=[prefix] * bs,
                    extend_num_tokens=t, req_to_token_pool=rtp, token_to_kv_pool=kvp,
                    attn_backend=backend)
                backend.init_forward_metadata(fb)

                def timed(fast):
                    # the library reads B200_EXTEND_FAST_SOFTMAX per launch (0 = round-1 softmax)
                    os.environ["B200_EXTEND_FAST_SOFTMAX"] = "1" if fast else "0"
                    for _ in range(2):
                        layer(q, k, v, fb)
                    torch.cuda.synchronize()
                    _lib.timing_enable(True)
                    for _ in range(args.iters):
                        flush.zero_()
                        layer(q, k, v, fb)
                    torch.cuda.synchronize()
                    _lib.timing_enable(False)
                    n, ms = _lib.timing_read("attn_extend")
                    _lib.timing_read("")
                    return ms / max(n, 1) * 1e3

                us_r1 = timed(False) if args.ab else None
                us = timed(True)
                flops = 4.0 * hq * d * bs * ext * (prefix + (ext + 1) / 2)
                tf = flops / (us * 1e-6) / 1e12
                row = {"shape": name, "Hq": hq, "Hkv": hkv, "D": d, "bs": bs, "ctx": ctx,
                       "radix_hit": hit, "prefix": prefix, "extend": ext,
                       "us_per_launch": round(us, 1), "tflops": round(tf, 1),
                       "frac_of_measured_bf16_peak": round(tf / peak, 3),
                       "prefill_tokens_per_s_per_layer": round(t / (us * 1e-6))}
                if us_r1 is not None:
                    row["us_round1_softmax"] = round(us_r1, 1)
                    row["tflops_round1_softmax"] = round(flops / (us_r1 * 1e-6) / 1e12, 1)
                print(json.dumps(row), flush=True)
                rows.append(row)
                del kvp, rtp, backend
                torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
