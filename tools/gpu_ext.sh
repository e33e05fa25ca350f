#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_attn_extend_gpu.py tests/test_fp8_kv_gpu.py -m gpu -x -q > gpurun_out/pytest_ext.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_ext.log
timeout 600 python tools/bench_extend.py --out gpurun_out/bench_extend.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); print(r['shape'],'bs',r['bs'],'ctx',r['ctx'],'hit',r['radix_hit'],'us',r['us_per_launch'],'TF',r['tflops'])
"
