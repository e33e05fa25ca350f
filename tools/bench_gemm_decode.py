#!/usr/bin/env python
"""Decode-regime (M <= 256) GEMM timing the way a decode step runs them: many launches back to back
from a CUDA graph, rotating over enough distinct weight copies (> 2x the 126 MB L2) that every
launch streams its weights from HBM, as the 32 distinct layers of a model do.  Reports
us/launch (graph total / launches) and GB/s of algorithmic bytes for
  fp8 per-channel (b200_fp8_gemm_rowcol), fp8 block-128 (b200_fp8_gemm_block128),
  torch._scaled_mm row-wise (cuBLASLt, the library competitor) and optionally W4A16.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from sglang_security_b200.layers.quantization.fp8_kernel import (fp8_blockwise_scaled_mm,
                                                                fp8_gemm_smallm_config,
                                                                fp8_scaled_mm)

FP8 = torch.float8_e4m3fn
SHAPES = {"llama8b": [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)],
          "llama70b_tp8": [(1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)],
          "llama8b_tp2": [(3072, 4096), (4096, 2048), (14336, 4096), (4096, 7168)],
          "llama8b_tp4": [(1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584)],
          "llama8b_tp8": [(768, 4096), (4096, 512), (3584, 4096), (4096, 1792)]}


def graph_time(fns, reps):
    """fns: list of zero-arg callables (one per weight copy).  Captures reps passes over all of
    them into one CUDA graph and returns seconds per launch (best of 5 replays)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for f in fns:
            f()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            for f in fns:
                f()
    g.replay()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e-3 / (reps * len(fns))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ms", default="1,8,16,32,64,128,256")
    ap.add_argument("--model", default="llama8b")
    ap.add_argument("--shapes", default="", help="optional NxK,NxK filter overriding --model")
    ap.add_argument("--variants", default="auto", help="comma list: auto,big,s1,s2,s3,s4,s6,s8")
    ap.add_argument("--kinds", default="rowcol,block128,torch_rowwise")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bench_gemm_decode.json"))
    args = ap.parse_args()
    dev = "cuda"
    rows = []
    # launch-overhead floor of this method: an empty-ish kernel in the same graph pattern
    z = torch.zeros(1, device=dev)
    t0 = graph_time([lambda: z.add_(1)], 64)
    print(json.dumps({"graph_launch_floor_us": round(t0 * 1e6, 2)}), flush=True)
    rows.append({"graph_launch_floor_us": t0 * 1e6})
    kinds = args.kinds.split(",")
    shapes = ([tuple(int(v) for v in t.split('x')) for t in args.shapes.split(',')] if args.shapes
              else SHAPES[args.model])
    for n, k in shapes:
        copies = max(2, min(32, -(-400_000_000 // (n * k))))
        ws = [torch.randn(n, k, device=dev).to(FP8) for _ in range(copies)]
        sb = torch.rand(n, device=dev) * 1e-2
        bs = torch.rand((n + 127) // 128, (k + 127) // 128, device=dev) * 1e-2
        reps = max(1, 64 // copies)
        for m in [int(x) for x in args.ms.split(",")]:
            x = torch.randn(m, k, device=dev).to(FP8)
            sa = torch.rand(m, device=dev) * 1e-2
            a_s = torch.rand(m, (k + 127) // 128, device=dev) * 1e-2
            bytes_ = n * k + m * k + m * n * 2 + n * 4 + m * 4
            res = {"M": m, "N": n, "K": k, "weight_copies": copies,
                   "hbm_floor_us": round(bytes_ / 6548.5e9 * 1e6, 2)}
            for var in args.variants.split(","):
                if var != "auto" and m > 128:
                    continue
                if var == "auto":
                    fp8_gemm_smallm_config(-1, -1)
                elif var == "big":
                    fp8_gemm_smallm_config(0, -1)
                else:
                    fp8_gemm_smallm_config(128, int(var[1:]))
                for kind in kinds:
                    if kind == "torch_rowwise" and var != "auto":
                        continue
                    if kind == "rowcol":
                        fns = [(lambda w=w: fp8_scaled_mm(x, w.t(), sa, sb, torch.bfloat16)) for w in ws]
                    elif kind == "block128":
                        fns = [(lambda w=w: fp8_blockwise_scaled_mm(x, w, a_s, bs, torch.bfloat16))
                               for w in ws]
                    else:
                        fns = [(lambda w=w: torch._scaled_mm(x, w.t(), scale_a=sa.view(-1, 1),
                                                             scale_b=sb.view(1, -1),
                                                             out_dtype=torch.bfloat16)) for w in ws]
                    try:
                        t = graph_time(fns, reps)
                        key = kind if var == "auto" else f"{kind}_{var}"
                        res[key + "_us"] = round(t * 1e6, 2)
                        if var == "auto":
                            res[key + "_gbs"] = round(bytes_ / t / 1e9, 0)
                    except Exception as e:  # noqa: BLE001
                        res[f"{kind}_{var}_err"] = str(e)[:100]
            fp8_gemm_smallm_config(-1, -1)
            print(json.dumps(res), flush=True)
            rows.append(res)
        del ws
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
