#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_gpu.py -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_fp8.log
timeout 900 python tools/bench_gemm.py --ms 1,8,16,32,64,128,256 --smallm-sweep --out gpurun_out/bench_gemm_smallm.json > gpurun_out/bench_gemm_smallm.log 2>&1; echo "bench rc=$?"; tail -30 gpurun_out/bench_gemm_smallm.log
