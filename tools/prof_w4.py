#!/usr/bin/env python
"""Tiny driver for ncu captures of the W4A16 tcgen05 GEMM (one shape, a few launches)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from sglang_security_b200.layers.quantization.w4a16_kernel import unpack_zeros, w4a16_gemm, w4a16_gemv_config

m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
n = int(sys.argv[2]) if len(sys.argv) > 2 else 27648
k = int(sys.argv[3]) if len(sys.argv) > 3 else 5120
dt = torch.bfloat16
qw = torch.randint(-2 ** 31, 2 ** 31 - 1, (k, n // 8), device="cuda", dtype=torch.int64).to(torch.int32)
qz = torch.randint(-2 ** 31, 2 ** 31 - 1, (k // 128, n // 8), device="cuda", dtype=torch.int64).to(torch.int32)
sc = (torch.rand(k // 128, n, device="cuda") * 0.02).to(dt)
z = unpack_zeros(qz, dt, True, False)
x = torch.randn(m, k, device="cuda").to(dt)
w4a16_gemv_config(0, -1)
for _ in range(4):
    y = w4a16_gemm(x, qw, z, sc, 128)
torch.cuda.synchronize()
print(y.float().abs().mean().item())
