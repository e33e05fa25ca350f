#!/bin/bash
# multi-GPU pass: bench.py (attention, heads sharded) and the whole-model TP decode bench
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n$N.err; cut -c1-400 gpurun_out/bench_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/bench_decode_model.py --bs 1,16,64,128 --ctx 1024,4096 --out gpurun_out/decode_model_fp8_tp$N.json > gpurun_out/decode_model_tp$N.log 2>&1; echo "model rc=$?"
grep -v "^{" gpurun_out/decode_model_tp$N.log | tail -5
python - <<PY
import json
for l in open("gpurun_out/decode_model_tp$N.log"):
    if not l.startswith("{"): continue
    r=json.loads(l); print("tp",r["tp"],"bs",r["bs"],"ctx",r["ctx"],"ms",r["ms_per_step"],"tok/s",r["tokens_per_s"],"frac",r["frac_of_hbm_floor"])
PY
