#!/bin/bash
# 8-GPU all-reduce latency (peer-memory kernel vs NCCL) and 70B TP8 decode with the custom all-reduce off.
N=8
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/bench_allreduce.py --out gpurun_out/bench_allreduce_tp$N.json > gpurun_out/bench_allreduce_tp$N.log 2>&1; echo "ar bench rc=$?"; grep "^{" gpurun_out/bench_allreduce_tp$N.log | cut -c1-300; grep -v "^{" gpurun_out/bench_allreduce_tp$N.log | tail -3
B200_CUSTOM_AR=0 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29530 tools/bench_decode_model.py --model llama3.1-70b --quant fp8 --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_llama70b_fp8_tp8_nccl.json > gpurun_out/dm_l70_tp8_nccl.log 2>&1; echo "model nccl rc=$?"
python - <<PY
import json
for l in open("gpurun_out/dm_l70_tp8_nccl.log"):
    if not l.startswith("{"): continue
    r=json.loads(l); print("nccl tp",r["tp"],"bs",r["bs"],"ctx",r["ctx"],"ms",r["ms_per_step"],"tok/s",r["tokens_per_s"])
PY
