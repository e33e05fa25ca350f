#!/bin/bash
# whole-model decode (Llama-3.1-8B FP8 W8A8), 1 GPU: bf16 KV pool and fp8 KV pool
mkdir -p gpurun_out
for KV in bf16 fp8; do
timeout 900 python tools/bench_decode_model.py --bs 1,8,16,32,64,128,256 --ctx 1024,4096 --kv-dtype $KV --out gpurun_out/decode_model_fp8_tp1_kv$KV.json > gpurun_out/decode_model_tp1_kv$KV.log 2>&1; echo "model kv=$KV rc=$?"
grep -v "^{" gpurun_out/decode_model_tp1_kv$KV.log | tail -3
python - <<PY
import json
for l in open("gpurun_out/decode_model_tp1_kv$KV.log"):
    if not l.startswith("{"): continue
    r=json.loads(l); print("kv","$KV","bs",r["bs"],"ctx",r["ctx"],"ms",r["ms_per_step"],"tok/s",r["tokens_per_s"],"frac",r["frac_of_hbm_floor"])
PY
done
