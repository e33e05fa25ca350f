#!/usr/bin/env python
"""Tiny driver for ncu captures of the FP8 per-channel GEMM (one shape, a few launches over rotating weights)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6144
k = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
if len(sys.argv) > 5:   # force the swap-AB small-M kernel: max_m, splits
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_gemm_smallm_config

    fp8_gemm_smallm_config(int(sys.argv[4]), int(sys.argv[5]))
FP8 = torch.float8_e4m3fn
ws = [torch.randn(n, k, device="cuda").to(FP8) for _ in range(8)]
x = torch.randn(m, k, device="cuda").to(FP8)
sa = torch.rand(m, device="cuda") * 1e-2
sb = torch.rand(n, device="cuda") * 1e-2
for w in ws:
    y = fp8_scaled_mm(x, w.t(), sa, sb, torch.bfloat16)
torch.cuda.synchronize()
from sglang_security_b200 import _lib

print(y.float().abs().mean().item(), _lib.load().b200_last_variant())
