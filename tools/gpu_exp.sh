#!/bin/bash
# Scratch experiment list of the CURRENT GPU session (rewritten per session; results land in
# gpurun_out/, what is worth keeping is copied to profiles/ by hand).  Run: gpurun -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
tag=${GPU_TAG:-r02b}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }

stamp "pytest gpu"
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 $out/${tag}_pytest_gpu.log | cut -c1-300

stamp "bench fused tail (default)"
timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_fused.json 2> $out/${tag}_bench_fused.err; echo "rc=$?"
python - <<'EOF'
import json,sys
for f in ["gpurun_out/%s_bench_fused.json" % __import__("os").environ.get("GPU_TAG","r02b")]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], json.dumps(d.get("kernel_split_eager")), d["roofline"]["us_per_launch"])
    except Exception as e: print("ERR", f, e)
EOF
tail -3 $out/${tag}_bench_fused.err

stamp "bench unfused tail"
B200_ATTN_FUSED_TAIL=0 timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_unfused.json 2> $out/${tag}_bench_unfused.err; echo "rc=$?"
cut -c1-400 $out/${tag}_bench_unfused.json

stamp "bench fused tail, silu one CTA per row"
B200_SILU_TARGET_CTAS=64 timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_silu1.json 2> $out/${tag}_bench_silu1.err; echo "rc=$?"
cut -c1-400 $out/${tag}_bench_silu1.json

stamp "w4a16 gemv2"
timeout 600 python tools/bench_w4a16.py --tp tp2 --ms 1,4,8,16 --skip-ref --splits2 2,4,8 --out $out/${tag}_bench_w4a16_gemv2.json > $out/${tag}_bench_w4a16.log 2>&1; echo "rc=$?"
cat $out/${tag}_bench_w4a16.log | cut -c1-600

stamp "fp8 small-M MT=64 after the smem fix"
timeout 600 python tools/bench_gemm_decode.py --ms 32,48,64 --variants auto,big,s2,s4,s8 --kinds rowcol,torch_rowwise --out $out/${tag}_bench_gemm_midm.json > $out/${tag}_bench_gemm_midm.log 2>&1; echo "rc=$?"
cat $out/${tag}_bench_gemm_midm.log | cut -c1-600

stamp "ncu launch list (fused)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/ncu_summarize.py launches $out/${tag}_launches.csv $out/${tag}_ncu_launches.json | tail -40

stamp "ncu full: gemv2 + smallm mt64"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:w4a16_gemv2_kernel -s 440 -c 2 -f -o $out/${tag}_gemv2_full python tools/bench_w4a16.py --tp tp2 --ms 1 --skip-ref --out $out/${tag}_scratch.json > $out/${tag}_ncu_gemv2.log 2>&1; echo "ncu gemv2 rc=$?"
python tools/ncu_summarize.py full $out/${tag}_gemv2_full.ncu-rep $out/${tag}_ncu_gemv2_summary.json | tail -45
stamp "done"
