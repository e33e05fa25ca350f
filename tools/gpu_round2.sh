#!/bin/bash
# full parity suite + mid-M sweep after the split-K / dispatch changes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python tools/bench_gemm_decode.py --ms 16,32,48,64,96,128,192,256 --kinds rowcol,torch_rowwise --variants auto,big --out gpurun_out/bench_gemm_midm_splitk.json > gpurun_out/bench_gemm_midm_splitk.log 2>&1; echo "midm rc=$?"; cut -c1-330 gpurun_out/bench_gemm_midm_splitk.log
