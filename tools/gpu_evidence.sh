#!/bin/bash
# tensor-roofline evidence: extend attention + FP8 GEMM sweeps, then ncu --set full of each
mkdir -p gpurun_out
timeout 900 python tools/bench_extend.py > gpurun_out/bench_extend.log 2>&1; echo "extend rc=$?"; cat gpurun_out/bench_extend.log | cut -c1-260
timeout 900 python tools/bench_gemm.py --ms 512,2048,8192,16384 --out gpurun_out/bench_gemm_large.json > gpurun_out/bench_gemm_large.log 2>&1; echo "gemm rc=$?"; cut -c1-330 gpurun_out/bench_gemm_large.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_extend -s 2 -c 1 -f -o gpurun_out/extend_full python tools/bench_extend.py --quick --iters 2 --out gpurun_out/tmp.json > gpurun_out/ncu_extend.log 2>&1; echo "ncu extend rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tn_kernel -s 4 -c 1 -f -o gpurun_out/fp8gemm_full python tools/bench_gemm.py --ms 8192 --out gpurun_out/tmp.json > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
