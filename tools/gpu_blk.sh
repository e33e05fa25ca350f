#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_gpu.py -m gpu -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_fp8.log
timeout 900 python tools/bench_gemm.py --ms 512,2048,8192,16384 --out gpurun_out/bench_gemm_large.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l)
        if 'M' in r: print(r['M'],r['N'],r['K'],'rowcol',r.get('rowcol_tflops'),'block',r.get('block128_tflops'),'torch',r.get('torch_rowwise_tflops'))
        else: print(r)
"
