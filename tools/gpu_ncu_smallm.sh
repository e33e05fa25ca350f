#!/bin/bash
mkdir -p gpurun_out
for mt in 16 128; do
B200_SMALLM_FORCE_MT=$mt timeout 600 ncu --set full --clock-control none --import-source on -k regex:smallm -s 6 -c 1 -f -o gpurun_out/smallm_mt${mt} python tools/bench_gemm_decode.py --ms 1 --variants auto --kinds rowcol --shapes 28672x4096 --out gpurun_out/tmp.json > gpurun_out/ncu_smallm_mt${mt}.log 2>&1; echo "ncu mt$mt rc=$?"
done
