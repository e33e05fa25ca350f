#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: large-M FP8 GEMM re-measure after the epilogue
# staging, FP8-KV decode attention with the single-wave split rule, then the full verification.
set -u
out=gpurun_out
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
stamp "large-M FP8 GEMM vs cuBLASLt"
timeout 400 python tools/bench_gemm.py --ms 512,2048,8192 --out $out/r02_bench_gemm_large.json 2>&1 | cut -c1-600 | tail -16
stamp "fp8-KV decode attention"
timeout 300 python tools/sweep_decode.py --tune --kv-dtype fp8 --out $out/r02_sweep_decode_fp8kv.json 2>&1 | cut -c1-260
stamp "verify"
bash tools/gpu_session.sh verify
stamp "done"
