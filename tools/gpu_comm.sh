#!/bin/bash
# N-GPU pass for the peer-memory communicator: parity tests (local + cudaIpc multi-process),
# all-reduce / fused row-parallel GEMM latency vs NCCL, whole-model TP decode with and without it.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_comm_gpu.py -q --durations=8 > gpurun_out/pytest_comm_n$N.log 2>&1; echo "pytest comm rc=$?"; grep -E "passed|failed|FAILED|s call" gpurun_out/pytest_comm_n$N.log | head -30
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/bench_allreduce.py --out gpurun_out/bench_allreduce_tp$N.json > gpurun_out/bench_allreduce_tp$N.log 2>&1; echo "ar bench rc=$?"; grep "^{" gpurun_out/bench_allreduce_tp$N.log | cut -c1-250; grep -v "^{" gpurun_out/bench_allreduce_tp$N.log | tail -5
for AR in 1 0; do
B200_CUSTOM_AR=$AR timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$AR tools/bench_decode_model.py --bs 1,16,64,128 --ctx 1024,4096 --out gpurun_out/decode_model_fp8_tp${N}_ar$AR.json > gpurun_out/decode_model_tp${N}_ar$AR.log 2>&1; echo "model AR=$AR rc=$?"
grep -v "^{" gpurun_out/decode_model_tp${N}_ar$AR.log | tail -3
python - <<PY
import json
for l in open("gpurun_out/decode_model_tp${N}_ar$AR.log"):
    if not l.startswith("{"): continue
    r=json.loads(l); print("custom_ar",$AR,"tp",r["tp"],"bs",r["bs"],"ctx",r["ctx"],"ms",r["ms_per_step"],"tok/s",r["tokens_per_s"],"frac",r["frac_of_hbm_floor"])
PY
done
