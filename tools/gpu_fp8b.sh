#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/bench_gemm_decode.py --ms 1,16,64,128,256 --variants auto,big,s1,s2,s4,s8 > gpurun_out/bench_gemm_decode.log 2>&1; echo "bench rc=$?"; cat gpurun_out/bench_gemm_decode.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:smallm -c 300 --csv --log-file gpurun_out/ncu_smallm.csv python tools/bench_gemm_decode.py --ms 16,64 --variants auto --kinds rowcol --out gpurun_out/tmp.json > gpurun_out/ncu_smallm.log 2>&1; echo "ncu rc=$?"
