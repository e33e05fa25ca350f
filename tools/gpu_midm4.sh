#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_gpu.py tests/test_comm_gpu.py -m gpu -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_fp8.log | cut -c1-300
timeout 900 python tools/bench_gemm_decode.py --ms 48,64,128,256 --kinds rowcol,torch_rowwise --variants auto --out gpurun_out/bench_gemm_midm_ks2.json 2>&1 | grep "^{" | cut -c1-260
timeout 600 python tools/bench_decode_model.py --bs 64,128,256 --ctx 1024 --out gpurun_out/tmp_model.json 2>&1 | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    r=json.loads(l); print('bs',r['bs'],'ctx',r['ctx'],'ms',r['ms_per_step'],'tok/s',r['tokens_per_s'])
"
