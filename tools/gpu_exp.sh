#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: in-tree argmax kernel; batch-1 launch list.
set -u
out=gpurun_out
tag=${GPU_TAG:-r02k}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
stamp "pytest gpu"
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $out/${tag}_pytest_gpu.log | cut -c1-300
stamp "bench"
timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; grep -o '"ms_per_step": [0-9.]*' $out/${tag}_bench_n1.json | head -2
stamp "small batches"
timeout 300 python tools/bench_decode_model.py --bs 1,8,32 --ctx 1024,4096 --out $out/${tag}_decode_small.json 2>&1 | grep -o '"bs": [0-9]*, "ctx": [0-9]*\|"ms_per_step": [0-9.]*' | paste - -
stamp "ncu launch list batch 1"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/${tag}_launches_bs1.csv python tools/bench_decode_model.py --bs 1 --ctx 1024 --steps 1 > $out/${tag}_ncu_bs1.log 2>&1; echo "rc=$?"
python tools/ncu_summarize.py launches $out/${tag}_launches_bs1.csv $out/${tag}_ncu_decode_model_bs1_launches.json | grep -v "torch:" | tail -16 | cut -c1-220
stamp "done"
