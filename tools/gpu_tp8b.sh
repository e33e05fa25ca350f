#!/bin/bash
# 8-GPU refresh: Llama-3.1-70B FP8 TP=8 whole-model decode (bf16 KV) after the small-token elementwise changes
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tools/bench_decode_model.py --model llama3.1-70b --quant fp8 --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_llama70b_fp8_tp8.json > gpurun_out/dm_l70_tp8.log 2>&1; echo "l70 tp8 rc=$?"
grep -v "^{" gpurun_out/dm_l70_tp8.log | tail -3
python - <<'PY'
import json
for l in open("gpurun_out/dm_l70_tp8.log"):
    if not l.startswith("{"): continue
    r = json.loads(l); print("tp", r["tp"], "bs", r["bs"], "ctx", r["ctx"], "ms", r["ms_per_step"], "tok/s", r["tokens_per_s"])
PY
