#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: after staging the epilogue scales in shared memory.
set -u
out=gpurun_out
tag=${GPU_TAG:-r02g}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
stamp "pytest gpu"
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $out/${tag}_pytest_gpu.log | cut -c1-300
stamp "mid-M sweep llama8b"
timeout 400 python tools/bench_gemm_decode.py --ms 16,32,48,64,96,128 --variants auto,big,s1,s2,s4,s8 --kinds rowcol --out $out/${tag}_gemm_midm.json 2>&1 | cut -c1-420
for tp in tp2 tp4 tp8; do
stamp "mid-M sweep llama8b_$tp"
timeout 300 python tools/bench_gemm_decode.py --model llama8b_$tp --ms 32,64,128 --variants auto,big,s1,s2,s4,s8 --kinds rowcol --out $out/${tag}_gemm_midm_$tp.json 2>&1 | cut -c1-420
done
stamp "bench N=1"
timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; cut -c1-330 $out/${tag}_bench_n1.json; grep -o '"kernel_split_eager.*"comm"' $out/${tag}_bench_n1.json | cut -c1-900
stamp "bench N=1 --splits 1"
timeout 300 python bench.py --no-cpu-baseline --splits 1 > $out/${tag}_bench_n1_s1.json 2> $out/${tag}_bench_n1_s1.err; echo "rc=$?"; cut -c1-330 $out/${tag}_bench_n1_s1.json; grep -o '"attn_decode_split": {[^}]*}' $out/${tag}_bench_n1_s1.json
stamp "w4a16 gemv2 (bias hoist does not apply here; sanity)"
timeout 300 python tools/bench_w4a16.py --tp tp2 --ms 1,8,16 --skip-ref --out $out/${tag}_w4a16.json 2>&1 | cut -c1-330
stamp "done"
