#!/bin/bash
# W4A16 pass (BASELINE cfg 4): parity tests, the GEMV->GEMM crossover sweep over the Qwen2.5-32B AWQ
# per-rank shapes at TP=2 / TP=4, and one ncu full capture of the tcgen05 dequant-GEMM.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_w4a16_gpu.py -m gpu -x -q > gpurun_out/pytest_w4.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_w4.log
timeout 1500 python tools/bench_w4a16.py --tp tp2,tp4 --ms 1,2,4,8,16,32,48,64,96,128,192,256,512,1024,2048,4096 --out gpurun_out/bench_w4a16.json > gpurun_out/bench_w4a16.log 2>&1; echo "sweep rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm_kernel -s 2 -c 1 -f -o gpurun_out/w4_m4096 python tools/prof_w4.py 4096 27648 5120 > gpurun_out/ncu_w4.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm_kernel -s 2 -c 1 -f -o gpurun_out/w4_m64 python tools/prof_w4.py 64 5120 13824 >> gpurun_out/ncu_w4.log 2>&1; echo "ncu rc=$?"
