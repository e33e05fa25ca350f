#!/usr/bin/env python
"""Tuning sweep of the decode-attention kernel on the cfg-2 geometry (run on the GPU box).

Times attn_decode_split (+ merge) per launch with the library's per-kernel CUDA events while
rotating over several layers' pools (each ~1 GB, so L2 never helps) and prints achieved GB/s of
the algorithmic bytes against MEASURED_PEAKS.json for every (slots, num_splits, bs, ctx) point."""
import argparse
import json
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from sglang_security_b200 import _lib
from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
from sglang_security_b200.layers.radix_attention import RadixAttention
from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode


def run_point(bs, ctx, hq, hkv, d, layers, slots_mode, splits, iters, ragged=False, kv_dtype="bf16"):
    dev = torch.device("cuda")
    size = bs * ctx
    rtp = ReqToTokenPool(bs + 1, ctx + 4, dev)
    fp8 = kv_dtype == "fp8"
    kvp = MHATokenToKVPool(size, 1, torch.float8_e4m3fn if fp8 else torch.bfloat16, hkv, d, layers, dev)
    for l in range(layers):
        if fp8:  # random e4m3 bytes, NaN patterns (0x7f / 0xff) remapped
            for buf in (kvp.k_buffer[l], kvp.v_buffer[l]):
                buf.random_(0, 256)
                buf[(buf & 0x7f) == 0x7f] = 0x38
        else:
            kvp.k_buffer[l].normal_()
            kvp.v_buffer[l].normal_()
    g = torch.Generator().manual_seed(0)
    slots = (torch.randperm(size, generator=g) + 1) if slots_mode == "random" else torch.arange(1, size + 1)
    rtp.req_to_token[:bs, :ctx] = slots.view(bs, ctx).to(torch.int32).to(dev)
    cfg = SimpleNamespace(num_attention_heads=hq, context_len=ctx, get_num_kv_heads=lambda t: hkv)
    mr = SimpleNamespace(device=dev, req_to_token_pool=rtp, token_to_kv_pool=kvp, model_config=cfg,
                         tp_size=1, server_args=SimpleNamespace(b200_attention_num_kv_splits=splits))
    backend = B200AttnBackend(mr)
    attn = [RadixAttention(hq, d, d ** -0.5, hkv, l) for l in range(layers)]
    q = torch.randn(layers, bs, hq * d, device=dev, dtype=torch.bfloat16)
    if ragged:
        lens = torch.randint(ctx // 2, ctx + 1, (bs,), generator=g).to(torch.int64)
    else:
        lens = torch.full((bs,), ctx, dtype=torch.int64)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs,
                      req_pool_indices=torch.arange(bs, device=dev), seq_lens=lens.to(dev),
                      out_cache_loc=torch.zeros(bs, dtype=torch.int64, device=dev),
                      seq_lens_sum=int(lens.sum()), req_to_token_pool=rtp, token_to_kv_pool=kvp,
                      attn_backend=backend)
    backend.init_forward_metadata(fb)
    for l in range(layers):
        attn[l](q[l], None, None, fb, save_kv_cache=False)
    torch.cuda.synchronize()
    _lib.timing_enable(True)
    for _ in range(iters):
        for l in range(layers):
            attn[l](q[l], None, None, fb, save_kv_cache=False)
    torch.cuda.synchronize()
    _lib.timing_enable(False)
    n, ms = _lib.timing_read("attn_decode_fp8kv" if fp8 else "attn_decode_split")
    nm, msm = _lib.timing_read("attn_merge")
    tokens = int(lens.sum())
    bytes_ = tokens * hkv * 2 * d * (1 if fp8 else 2) + bs * hq * d * 2 * 2 + tokens * 4
    us = ms / n * 1e3
    return dict(bs=bs, ctx=ctx, hq=hq, hkv=hkv, d=d, slots=slots_mode, ragged=ragged, kv_dtype=kv_dtype,
                splits=backend.forward_metadata.num_splits, split_us=round(us, 2),
                merge_us=round(msm / nm * 1e3, 2) if nm else 0.0,
                gbs=round(bytes_ / (us * 1e-6) / 1e9, 1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--tune", action="store_true", help="cfg-2 point only, a few split counts")
    ap.add_argument("--tp-shapes", action="store_true",
                    help="per-rank head geometry of Llama-3.1-8B at TP 2/4/8 over a range of split counts")
    ap.add_argument("--kv-dtype", default="bf16", choices=["bf16", "fp8"])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep_decode.json"))
    args = ap.parse_args()
    peak = 6541.5
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    rows = []
    points = []
    for slots in ("random", "contiguous"):
        for s in ([0, 1, 4, 8] if args.quick else [0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 16]):
            points.append(dict(bs=64, ctx=4096, hq=32, hkv=8, d=128, layers=4, slots_mode=slots, splits=s))
    if args.tune:
        sm = os.environ.get("SWEEP_SLOTS", "random")
        points = [dict(bs=64, ctx=4096, hq=32, hkv=8, d=128, layers=4, slots_mode=sm, splits=s)
                  for s in (0, 2, 3, 4, 5, 6, 8)]
        points.append(dict(bs=64, ctx=4096, hq=64, hkv=8, d=128, layers=4, slots_mode="random", splits=0))
        points.append(dict(bs=64, ctx=4096, hq=8, hkv=1, d=128, layers=16, slots_mode="random", splits=0))
    if args.tp_shapes:
        points = []
        for hq, hkv in ((16, 4), (8, 2), (4, 1)):
            for sp in (0, 2, 3, 4, 5, 6, 8, 9, 10, 12, 14, 16):
                points.append(dict(bs=64, ctx=4096, hq=hq, hkv=hkv, d=128, layers=16, slots_mode="random",
                                   splits=sp))
    if not args.quick and not args.tune and not args.tp_shapes:
        points += [
            dict(bs=64, ctx=4096, hq=32, hkv=8, d=128, layers=4, slots_mode="random", splits=0, ragged=True),
            dict(bs=1, ctx=4096, hq=32, hkv=8, d=128, layers=8, slots_mode="random", splits=0),
            dict(bs=8, ctx=4096, hq=32, hkv=8, d=128, layers=8, slots_mode="random", splits=0),
            dict(bs=256, ctx=1024, hq=32, hkv=8, d=128, layers=4, slots_mode="random", splits=0),
            dict(bs=64, ctx=4096, hq=8, hkv=1, d=128, layers=16, slots_mode="random", splits=0),   # 70B @ TP8
            dict(bs=64, ctx=4096, hq=4, hkv=1, d=128, layers=16, slots_mode="random", splits=0),   # 8B @ TP8
            dict(bs=8, ctx=528, hq=14, hkv=2, d=64, layers=24, slots_mode="random", splits=0),     # cfg 1
            dict(bs=16, ctx=16384, hq=32, hkv=8, d=128, layers=4, slots_mode="random", splits=0),
        ]
    for p in points:
        r = run_point(iters=5, kv_dtype=args.kv_dtype, **p)
        r["frac_of_measured_peak"] = round(r["gbs"] / peak, 3)
        rows.append(r)
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
