#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: full verification (tools/gpu_session.sh verify)
# plus the A/B of the cluster-split RMSNorm.
set -u
out=gpurun_out
mkdir -p $out
bash tools/gpu_session.sh verify
echo "== bench with one CTA per RMSNorm row"
B200_RMSNORM_TARGET_CTAS=0 timeout 300 python bench.py --no-cpu-baseline > $out/r02_bench_rmsnorm1.json 2> $out/r02_bench_rmsnorm1.err; echo "rc=$?"; cut -c1-330 $out/r02_bench_rmsnorm1.json
echo "== w4a16 / gemm sanity after the KSUB change"
timeout 300 python tools/bench_gemm_decode.py --ms 16,32,64 --variants auto --kinds rowcol --out $out/r02_gemm_auto.json 2>&1 | cut -c1-200
