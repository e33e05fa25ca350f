#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: source-level ncu of the mid-M FP8 GEMMs,
# attention split sweep on the TP per-rank head geometries, bench with the cluster merge+quant.
set -u
out=gpurun_out
tag=${GPU_TAG:-r02f}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
NCU="ncu --set full --clock-control none --import-source on -f"
stamp "ncu smallm M=64 s4"
timeout 200 $NCU -k regex:fp8_gemm_smallm -s 4 -c 1 -o $out/${tag}_smallm_m64 python tools/prof_fp8.py 64 6144 4096 128 4 > $out/${tag}_ncu1.log 2>&1; echo "rc=$?"; tail -1 $out/${tag}_ncu1.log | cut -c1-200
stamp "ncu smallm M=32 s4"
timeout 200 $NCU -k regex:fp8_gemm_smallm -s 4 -c 1 -o $out/${tag}_smallm_m32 python tools/prof_fp8.py 32 6144 4096 128 4 > $out/${tag}_ncu2.log 2>&1; echo "rc=$?"; tail -1 $out/${tag}_ncu2.log | cut -c1-200
stamp "ncu gemm_tn M=64"
timeout 200 $NCU -k regex:gemm_tn -s 4 -c 1 -o $out/${tag}_gemmtn_m64 python tools/prof_fp8.py 64 6144 4096 > $out/${tag}_ncu3.log 2>&1; echo "rc=$?"; tail -1 $out/${tag}_ncu3.log | cut -c1-200
for f in smallm_m64 smallm_m32 gemmtn_m64; do python tools/ncu_summarize.py full $out/${tag}_$f.ncu-rep $out/${tag}_ncu_${f}_summary.json | grep -E "duration|dram__bytes_read.sum\"|issue_active|warps_active|registers|occupancy_limit|grid_size|tensor_cycles" ; done
stamp "attention split sweep on TP shapes"
timeout 400 python tools/sweep_decode.py --tp-shapes --out $out/${tag}_sweep_tp_shapes.json 2>&1 | cut -c1-250
stamp "bench N=1 default / unfused"
timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; cut -c1-330 $out/${tag}_bench_n1.json
B200_ATTN_FUSED_TAIL=0 timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1_unfused.json 2> $out/${tag}_bench_n1_unfused.err; echo "rc=$?"; cut -c1-330 $out/${tag}_bench_n1_unfused.json
stamp "done"
