#!/usr/bin/env python
"""BASELINE cfg 4: Qwen2.5-32B AWQ INT4 W4A16 g128 per-rank GEMMs at TP=2 / TP=4, GEMV -> GEMM
crossover sweep.  Every launch streams distinct weights (copies rotate over > 2x L2) from a CUDA
graph.  Columns: the HBM-bound GEMV (M <= 16), the tcgen05 dequant-GEMM, and the reference's own
approach on the same GPU (awq_dequantize to a 16-bit [K, N] tensor + library matmul, awq.py:203-215).
Roofline: bytes = K*N/2 + (K/g)*N*4 + M*K*2 + M*N*2 vs measured HBM peak; flops 2MNK vs bf16 peak."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch

from bench_gemm_decode import graph_time
from sglang_security_b200.layers.quantization.w4a16_kernel import (awq_dequantize, gemv_stream_repack,
                                                                  unpack_zeros, w4a16_gemm,
                                                                  w4a16_gemv_config, w4a16_gemv_v2,
                                                                  w4a16_gemv_v2_config)

SHAPES = {"tp2": [(3584, 5120), (5120, 2560), (27648, 5120), (5120, 13824)],
          "tp4": [(1792, 5120), (5120, 1280), (13824, 5120), (5120, 6912)]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", default="tp2,tp4")
    ap.add_argument("--ms", default="1,2,4,8,16,32,48,64,96,128,192,256,512,1024,2048")
    ap.add_argument("--splits", default="")
    ap.add_argument("--splits2", default="", help="extra forced split-K factors for the stream GEMV")
    ap.add_argument("--skip-ref", action="store_true")
    ap.add_argument("--dtype", default="float16")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bench_w4a16.json"))
    args = ap.parse_args()
    dev, dt, group = "cuda", getattr(torch, args.dtype), 128
    rows = []
    for tp in args.tp.split(","):
        for n, k in SHAPES[tp]:
            wbytes = k * n // 2
            copies = max(2, min(16, -(-300_000_000 // wbytes)))
            ws = []
            for i in range(copies):
                qw = torch.randint(-2 ** 31, 2 ** 31 - 1, (k, n // 8), device=dev, dtype=torch.int64).to(torch.int32)
                qz = torch.randint(-2 ** 31, 2 ** 31 - 1, (k // group, n // 8), device=dev, dtype=torch.int64).to(torch.int32)
                sc = (torch.rand(k // group, n, device=dev) * 0.02).to(dt)
                ws.append((qw, qz, unpack_zeros(qz, dt, True, False), sc, gemv_stream_repack(qw)))
            reps = max(1, 32 // copies)
            for m in [int(v) for v in args.ms.split(",")]:
                x = torch.randn(m, k, device=dev).to(dt)
                bytes_ = wbytes + (k // group) * n * 4 + m * k * 2 + m * n * 2
                res = {"tp": tp, "M": m, "N": n, "K": k, "hbm_floor_us": round(bytes_ / 6548.5e9 * 1e6, 2)}
                variants = [("gemm_tcgen05", 0, -1)]
                if m <= 16:
                    variants.append(("gemv", 16, 0))
                    for s in [int(v) for v in args.splits.split(",") if v]:
                        variants.append((f"gemv_s{s}", 16, s))
                for name, max_m, s in variants:
                    w4a16_gemv_config(max_m, s)
                    fns = [(lambda w=w: w4a16_gemm(x, w[0], w[2], w[3], group)) for w in ws]
                    t = graph_time(fns, reps)
                    res[name + "_us"] = round(t * 1e6, 2)
                w4a16_gemv_config(-1, -1)
                if m <= 16:
                    for s in [0] + [int(v) for v in args.splits2.split(",") if v]:
                        w4a16_gemv_v2_config(s)
                        fns = [(lambda w=w: w4a16_gemv_v2(x, w[4], w[2], w[3], group)) for w in ws]
                        t2 = graph_time(fns, reps)
                        key = "gemv2" if s == 0 else f"gemv2_s{s}"
                        res[key + "_us"] = round(t2 * 1e6, 2)
                        if s == 0:
                            res["gemv2_gbs"] = round(bytes_ / t2 / 1e9)
                            res["gemv2_frac_hbm"] = round(bytes_ / t2 / 6548.5e9, 3)
                    w4a16_gemv_v2_config(0)
                fns = [(lambda w=w: w4a16_gemm(x, w[0], w[2], w[3], group)) for w in ws]
                t = graph_time(fns, reps)
                res["dispatch_us"] = round(t * 1e6, 2)
                res["dispatch_gbs"] = round(bytes_ / t / 1e9)
                res["dispatch_tflops"] = round(2.0 * m * n * k / t / 1e12, 1)
                if not args.skip_ref:
                    fns = [(lambda w=w: torch.matmul(x, awq_dequantize(w[0], w[3], w[1]))) for w in ws]
                    res["ref_dequant_matmul_us"] = round(graph_time(fns, reps) * 1e6, 2)
                print(json.dumps(res), flush=True)
                rows.append(res)
            del ws
            torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
