#!/usr/bin/env python
"""CPU baseline of the AWQ INT4 (W4A16, group 128) linear path — SURVEY §8(d) "CPU baseline" item 3:
the reference's torch-native route `awq_dequantize_torch + matmul` (sgl-kernel/tests/
test_awq_dequant.py:9-57 restated in oracle/int4.py, pinned by tests/golden/awq_dequant.pt), on all
host cores, at the per-rank shapes of Qwen2.5-32B TP 2 (BASELINE cfg 4) for decode-size batches.
A reported baseline next to the GPU numbers of profiles/r02_bench_w4a16_gemv2.json — not a target.
Needs no GPU.  Lives under tests/ because it executes oracle/ code (only tests/, smoke() and bench.py's
cpu_baseline leg may); not collected by pytest.  Run: python tests/cpu_baseline_w4a16.py"""
import argparse
import json
import os
import platform
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from oracle import int4 as O


def median_ms(fn, warm=1, iters=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e3)
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_cpu_baseline_w4a16.json"))
    ap.add_argument("--gpu-json", default=os.path.join(ROOT, "profiles", "r02_bench_w4a16_gemv2.json"))
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    gpu = {}
    if os.path.exists(args.gpu_json):
        for r in json.load(open(args.gpu_json)):
            gpu[(r["M"], r["N"], r["K"])] = r.get("gemv2_us")   # M <= 16 dispatches to the stream GEMV
    g = torch.Generator().manual_seed(0)
    rows = []
    for n, k in [(3584, 5120), (5120, 2560), (27648, 5120), (5120, 13824)]:
        group = 128
        qweight = torch.randint(0, 2 ** 31 - 1, (k, n // 8), generator=g, dtype=torch.int32)
        qzeros = torch.randint(0, 2 ** 31 - 1, (k // group, n // 8), generator=g, dtype=torch.int32)
        scales = (torch.rand((k // group, n), generator=g) * 0.02).to(torch.float16)
        dq_ms = median_ms(lambda: O.awq_dequantize_oracle(qweight, scales, qzeros, group))
        w = O.awq_dequantize_oracle(qweight, scales, qzeros, group)
        for m in (1, 16):
            x = torch.randn((m, k), generator=g).to(torch.float16)
            mm_ms = median_ms(lambda: O.w4a16_linear_oracle(x, w))
            total = dq_ms + mm_ms           # the reference's apply() dequantises on every call (awq.py:197-215)
            row = {"M": m, "N": n, "K": k, "cpu_dequant_ms": round(dq_ms, 2), "cpu_matmul_ms": round(mm_ms, 2),
                   "cpu_total_ms": round(total, 2), "gpu_gemv2_us": gpu.get((m, n, k))}
            if row["gpu_gemv2_us"]:
                row["gpu_over_cpu"] = round(total * 1e3 / row["gpu_gemv2_us"], 0)
            print(json.dumps(row), flush=True)
            rows.append(row)
    out = {"what": "AWQ W4A16 linear on the CPU: oracle restatement of awq_dequantize_torch + matmul (kind = port)",
           "cores": os.cpu_count(), "machine": platform.processor() or platform.machine(),
           "where": "authoring container (no GPU); GPU column = B200 numbers of " + os.path.basename(args.gpu_json),
           "torch": torch.__version__, "rows": rows}
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
