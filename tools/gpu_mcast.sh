#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_gpu.py -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_fp8.log | cut -c1-300
for MC in 1 0; do
B200_GEMM_MCAST=$MC timeout 900 python tools/bench_gemm.py --ms 2048,4096,8192,16384 --out gpurun_out/bench_gemm_large_mc$MC.json > gpurun_out/bench_gemm_large_mc$MC.log 2>&1; echo "gemm MC=$MC rc=$?"; cut -c1-360 gpurun_out/bench_gemm_large_mc$MC.log | tail -20
done
