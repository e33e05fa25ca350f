#!/bin/bash
# W4A16 GEMM pass: parity tests, then the cfg-4 crossover sweep (TP2 / TP4 per-rank shapes).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_w4a16_gpu.py -m gpu -x -q > gpurun_out/pytest_w4.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_w4.log
timeout 1200 python tools/bench_w4a16.py --tp ${W4_TP:-tp2,tp4} --ms ${W4_MS:-1,8,16,32,64,128,256,512,1024,2048,4096} --out gpurun_out/bench_w4a16.json > gpurun_out/bench_w4a16.log 2>&1; echo "sweep rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/bench_w4a16.log'):
    if l.startswith('{'):
        r=json.loads(l); print(r['tp'],r['M'],r['N'],r['K'],'gemm',r['gemm_tcgen05_us'],'disp',r['dispatch_us'],r['dispatch_tflops'],'TF',r['dispatch_gbs'],'GB/s ref',r['ref_dequant_matmul_us'])
P
