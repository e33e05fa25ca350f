#!/bin/bash
N=8
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -x -q -k "multi_gpu" > gpurun_out/pytest_comm_n$N.log 2>&1; echo "comm tests rc=$?"; tail -2 gpurun_out/pytest_comm_n$N.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29530 tools/bench_decode_model.py --model llama3.1-70b --quant fp8 --bs 1,4,16,32,64,256 --ctx 1024 --out gpurun_out/decode_model_llama70b_fp8_tp8.json > gpurun_out/dm_l70_tp8.log 2>&1; echo "model rc=$?"
python - <<PY
import json
for l in open("gpurun_out/dm_l70_tp8.log"):
    if not l.startswith("{"): continue
    r=json.loads(l); print("tp",r["tp"],"bs",r["bs"],"ctx",r["ctx"],"ms",r["ms_per_step"],"tok/s",r["tokens_per_s"])
PY
