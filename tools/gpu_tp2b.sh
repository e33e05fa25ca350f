#!/bin/bash
# 2-GPU refresh: whole-model TP=2 decode for Llama-3.1-8B FP8 (FP8 KV) and Qwen2.5-32B AWQ
mkdir -p gpurun_out
rm -f gpurun_out/dm_*.log
run() { timeout "$1" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$2" --master-addr 127.0.0.1 --master-port "$3" "${@:4}"; }
run 400 2 29514 tools/bench_decode_model.py --model qwen2.5-32b --quant awq --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_qwen32b_awq_tp2.json > gpurun_out/dm_q32_tp2.log 2>&1; echo "q32 tp2 rc=$?"
run 400 2 29515 tools/bench_decode_model.py --model llama3.1-8b --quant fp8 --kv-dtype fp8 --bs 1,16,64,128 --ctx 1024,4096 --out gpurun_out/decode_model_llama8b_fp8_tp2_kvfp8.json > gpurun_out/dm_l8_tp2.log 2>&1; echo "l8 tp2 rc=$?"
for f in gpurun_out/dm_q32_tp2.log gpurun_out/dm_l8_tp2.log; do echo "== $f"; grep -v "^{" $f | tail -3; python - "$f" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if not l.startswith("{"): continue
    r = json.loads(l); print("tp", r["tp"], "bs", r["bs"], "ctx", r["ctx"], "ms", r["ms_per_step"], "tok/s", r["tokens_per_s"])
PY
done
