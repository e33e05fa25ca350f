#!/bin/bash
# N-GPU pass (N = 4 or 8): communicator parity tests, bench.py scaling point, whole-model TP decode
# for the BASELINE configs that name this TP degree.
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
run() { timeout "$1" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$2" --master-addr 127.0.0.1 --master-port "$3" "${@:4}"; }
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -x -q > gpurun_out/pytest_comm_n$N.log 2>&1; echo "comm tests rc=$?"; tail -2 gpurun_out/pytest_comm_n$N.log
run 600 $N 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n$N.json
if [ "$N" = "4" ]; then
  run 500 4 29512 tools/bench_decode_model.py --model llama3.1-8b --quant fp8 --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_llama8b_fp8_tp4.json > gpurun_out/dm_l8_tp4.log 2>&1; echo "l8 tp4 rc=$?"
  run 500 4 29513 tools/bench_decode_model.py --model qwen2.5-32b --quant awq --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_qwen32b_awq_tp4.json > gpurun_out/dm_q32_tp4.log 2>&1; echo "q32 tp4 rc=$?"
  run 500 2 29514 tools/bench_decode_model.py --model qwen2.5-32b --quant awq --bs 1,16,64,256 --ctx 1024 --out gpurun_out/decode_model_qwen32b_awq_tp2.json > gpurun_out/dm_q32_tp2.log 2>&1; echo "q32 tp2 rc=$?"
else
  run 600 8 29512 tools/bench_decode_model.py --model llama3.1-70b --quant fp8 --bs 1,16,64,256 --ctx 1024,4096 --out gpurun_out/decode_model_llama70b_fp8_tp8.json > gpurun_out/dm_l70_tp8.log 2>&1; echo "l70 tp8 rc=$?"
  run 400 8 29513 tools/bench_decode_model.py --model llama3.1-8b --quant fp8 --bs 1,64,256 --ctx 1024 --out gpurun_out/decode_model_llama8b_fp8_tp8.json > gpurun_out/dm_l8_tp8.log 2>&1; echo "l8 tp8 rc=$?"
fi
for f in gpurun_out/dm_*.log; do echo "== $f"; grep -v "^{" $f | tail -3; python - "$f" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if not l.startswith("{"): continue
    r = json.loads(l); print("tp", r["tp"], "bs", r["bs"], "ctx", r["ctx"], "ms", r["ms_per_step"], "tok/s", r["tokens_per_s"], "frac", r.get("frac_of_hbm_floor"))
PY
done
