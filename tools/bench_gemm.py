#!/usr/bin/env python
"""FP8 GEMM sweep on the cfg-3 shapes (Llama-3.1-8B): per-kernel CUDA-event times through the
library timing registry, L2 flushed between launches, reported as TFLOP/s (M >= 512, tensor
roofline) and GB/s of algorithmic bytes (M <= 256, HBM roofline).  Also measures the dense FP8
peak stand-in (torch._scaled_mm 8192^3) and torch._scaled_mm row-wise as the library competitor."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from sglang_security_b200 import _lib
from sglang_security_b200.layers.quantization.fp8_kernel import (fp8_blockwise_scaled_mm,
                                                                fp8_gemm_smallm_config,
                                                                fp8_scaled_mm)

FP8 = torch.float8_e4m3fn
SHAPES = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]


def time_events(fn, iters, flush):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in ev:
        flush.zero_()
        s.record()
        fn()
        e.record()
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in ev)
    return ts[len(ts) // 2] * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ms", default="1,16,64,256,1024,4096,8192,16384")
    ap.add_argument("--smallm-sweep", action="store_true",
                    help="for M <= 128 also time the big-tile kernel and forced cluster split-K factors")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bench_gemm.json"))
    args = ap.parse_args()
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    # dense FP8 peak stand-in
    a = torch.randn(8192, 8192, device=dev).to(FP8)
    b = torch.randn(8192, 8192, device=dev).to(FP8)
    one = torch.ones((), device=dev)
    f = lambda: torch._scaled_mm(a, b.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
    for _ in range(3):
        f()
    t = time_events(f, 10, flush)
    peak = 2 * 8192 ** 3 / t / 1e12
    print(json.dumps({"fp8_dense_peak_tflops_scaled_mm_8192": round(peak, 1)}), flush=True)
    rows.append({"fp8_dense_peak_tflops_scaled_mm_8192": peak})
    del a, b
    for n, k in SHAPES:
        w = torch.randn(n, k, device=dev).to(FP8)
        sb = torch.rand(n, device=dev) * 1e-2
        bs = torch.rand((n + 127) // 128, (k + 127) // 128, device=dev) * 1e-2
        for m in [int(x) for x in args.ms.split(",")]:
            x = torch.randn(m, k, device=dev).to(FP8)
            sa = torch.rand(m, device=dev) * 1e-2
            a_s = torch.rand(m, (k + 127) // 128, device=dev) * 1e-2
            flops = 2.0 * m * n * k
            bytes_ = n * k + m * k + m * n * 2 + n * 4 + m * 4
            res = {"M": m, "N": n, "K": k}
            for name, fn in (
                ("rowcol", lambda: fp8_scaled_mm(x, w.t(), sa, sb, torch.bfloat16)),
                ("block128", lambda: fp8_blockwise_scaled_mm(x, w, a_s, bs, torch.bfloat16)),
                ("torch_rowwise", lambda: torch._scaled_mm(x, w.t(), scale_a=sa.view(-1, 1),
                                                           scale_b=sb.view(1, -1),
                                                           out_dtype=torch.bfloat16)),
            ):
                try:
                    for _ in range(2):
                        fn()
                    t = time_events(fn, 7, flush)
                    res[name + "_us"] = round(t * 1e6, 1)
                    res[name + "_tflops"] = round(flops / t / 1e12, 1)
                    res[name + "_gbs"] = round(bytes_ / t / 1e9, 1)
                except Exception as e:
                    res[name + "_err"] = str(e)[:80]
            if args.smallm_sweep and m <= 128:
                variants = [("big", 0, -1)] + [(f"s{sp}", 128, sp) for sp in (1, 2, 3, 4, 6, 8)]
                for tag, max_m, sp in variants:
                    fp8_gemm_smallm_config(max_m, sp)
                    for name, fn in (
                        ("rowcol", lambda: fp8_scaled_mm(x, w.t(), sa, sb, torch.bfloat16)),
                        ("block128", lambda: fp8_blockwise_scaled_mm(x, w, a_s, bs, torch.bfloat16)),
                    ):
                        for _ in range(2):
                            fn()
                        t = time_events(fn, 7, flush)
                        res[f"{name}_{tag}_us"] = round(t * 1e6, 1)
                fp8_gemm_smallm_config(-1, -1)
            print(json.dumps(res), flush=True)
            rows.append(res)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
