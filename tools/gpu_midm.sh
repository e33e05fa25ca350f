#!/bin/bash
# 1-GPU pass: full parity suite, mid-M (64..256) FP8 GEMM sweep over dispatch variants, ncu of both kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python tools/bench_gemm_decode.py --ms 48,64,96,128,192,256 --kinds rowcol,torch_rowwise --variants auto,big,s1,s2,s4,s8 --out gpurun_out/bench_gemm_midm.json > gpurun_out/bench_gemm_midm.log 2>&1; echo "midm rc=$?"; cut -c1-400 gpurun_out/bench_gemm_midm.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_smallm -s 20 -c 1 -f -o gpurun_out/smallm_m64_full python tools/bench_gemm_decode.py --ms 64 --shapes 6144x4096 --kinds rowcol --variants auto --out gpurun_out/tmp.json > gpurun_out/ncu_smallm.log 2>&1; echo "ncu smallm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tn_kernel -s 20 -c 1 -f -o gpurun_out/gemmtn_m128_full python tools/bench_gemm_decode.py --ms 128 --shapes 6144x4096 --kinds rowcol --variants big --out gpurun_out/tmp.json > gpurun_out/ncu_gemmtn.log 2>&1; echo "ncu gemm_tn rc=$?"
