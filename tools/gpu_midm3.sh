#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_gpu.py tests/test_fp8_kv_gpu.py tests/test_attn_extend_gpu.py -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_fp8.log | cut -c1-300
timeout 900 python tools/bench_gemm_decode.py --ms 48,64,96,128,192,256 --kinds rowcol,torch_rowwise --variants auto,big --out gpurun_out/bench_gemm_midm_splitk.json > gpurun_out/bench_gemm_midm_splitk.log 2>&1; echo "midm rc=$?"; cut -c1-330 gpurun_out/bench_gemm_midm_splitk.log
