#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: PDL retest at batch 64, whole-model sweeps for the
# DESIGN tables, cfg-4 (AWQ) and cfg-5 (SafeKV multi-turn) tools on one GPU, round-2 ncu captures.
set -u
out=gpurun_out
tag=${GPU_TAG:-r02h}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
brief() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ", sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"],3), "tok/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "splits", d["config"]["num_kv_splits"], "attn us", d["roofline"]["us_per_launch"], "frac", round(d["roofline"]["frac"],3), "step frac", round(d["roofline_step"]["frac"],3))
except Exception as e: print("ERR", sys.argv[1], e)
PY
}
stamp "bench default"
timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; brief $out/${tag}_bench_n1.json
for pm in 64 256; do
stamp "bench PDL up to $pm tokens"
B200_PDL_MAX_M=$pm timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_pdl$pm.json 2> $out/${tag}_bench_pdl$pm.err; echo "rc=$?"; brief $out/${tag}_bench_pdl$pm.json
done
stamp "whole-model sweep (bf16 KV)"
timeout 400 python tools/bench_decode_model.py --bs 1,8,16,32,64,128,256 --ctx 1024,4096 --out $out/${tag}_decode_model_llama8b_fp8_tp1.json 2>&1 | cut -c1-230
stamp "whole-model sweep (fp8 KV)"
timeout 400 python tools/bench_decode_model.py --bs 1,16,64,128 --ctx 1024,4096 --kv-dtype fp8 --out $out/${tag}_decode_model_llama8b_fp8_tp1_kvfp8.json 2>&1 | cut -c1-230
stamp "cfg 4: Qwen2.5-32B AWQ on one GPU"
timeout 500 python tools/bench_decode_model.py --model qwen2.5-32b --quant awq --bs 1,8,16,64 --ctx 1024 --out $out/${tag}_decode_model_qwen32b_awq_tp1.json 2>&1 | cut -c1-330
stamp "cfg 5 tool on one GPU (Llama-3.1-8B)"
timeout 400 python tools/bench_safekv_multiturn.py --users 16 --turns 3 --out $out/${tag}_safekv_llama8b_tp1.json 2>&1 | tail -5 | cut -c1-500
stamp "ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "rc=$?"
python tools/ncu_summarize.py launches $out/${tag}_launches.csv $out/${tag}_ncu_launches.json | grep -v "torch:" | tail -14 | cut -c1-200
stamp "ncu full attention"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_mma -s 40 -c 2 -f -o $out/${tag}_decode_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "rc=$?"
python tools/ncu_summarize.py full $out/${tag}_decode_full.ncu-rep $out/${tag}_ncu_decode_full_summary.json | grep -E "duration|dram__bytes|registers|issue_active" | head -8
stamp "done"
