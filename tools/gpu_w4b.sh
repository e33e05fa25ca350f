#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_w4a16_gpu.py -m gpu -x -q > gpurun_out/pytest_w4.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_w4.log
for cfg in ${W4_CFGS:-1 2 0}; do
echo "== CFG=$cfg"
B200_W4_CFG=$cfg timeout 600 python tools/bench_w4a16.py --tp tp2 --ms ${W4_MS:-32,128,512,4096} --out gpurun_out/bench_w4a16_cfg$cfg.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); print(r['tp'],r['M'],r['N'],r['K'],'gemm',r['gemm_tcgen05_us'],r['dispatch_tflops'],'TF',r['dispatch_gbs'],'GB/s ref',r['ref_dequant_matmul_us'])
"
done
