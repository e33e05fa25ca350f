#!/bin/bash
# Scratch experiment list of the CURRENT GPU session (rewritten per session; results land in
# gpurun_out/, what is worth keeping is copied to profiles/ by hand).
# This one: 2-GPU session — NVLS all-reduce bring-up + TP=2 whole-model bench.  Run:
#   gpurun --gpus 2 -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
tag=${GPU_TAG:-r02d}
N=${GPU_N:-2}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
trun() { port=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }

nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
stamp "comm tests (ipc + nvls)"
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -q -x -k "multi_gpu and ($N)" -rs > $out/${tag}_pytest_comm.log 2>&1; echo "rc=$?"; tail -15 $out/${tag}_pytest_comm.log | cut -c1-400
stamp "attention + fp8 + w4a16 tests"
timeout 600 python -m pytest tests/test_attn_decode_gpu.py tests/test_fp8_kv_gpu.py tests/test_fp8_gpu.py -m gpu -q > $out/${tag}_pytest_attn.log 2>&1; echo "rc=$?"; tail -5 $out/${tag}_pytest_attn.log | cut -c1-300
stamp "all-reduce latency"
timeout 600 bash -c "$(declare -f trun); N=$N; trun 29521 tools/bench_allreduce.py --ar-only --out $out/${tag}_bench_allreduce_tp$N.json" 2>&1 | grep -v "^W\|^\*\*\*\|^$" | cut -c1-300 | tail -20
stamp "bench.py N=$N"
timeout 900 bash -c "$(declare -f trun); N=$N; trun 29522 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "rc=$?"; cut -c1-3000 $out/${tag}_bench_n$N.json; tail -3 $out/${tag}_bench_n$N.err
stamp "bench.py N=$N without NVLS"
B200_COMM_NVLS=0 timeout 900 bash -c "$(declare -f trun); N=$N; trun 29523 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n${N}_nonvls.json 2> $out/${tag}_bench_n${N}_nonvls.err; echo "rc=$?"; cut -c1-600 $out/${tag}_bench_n${N}_nonvls.json
stamp "bench.py N=1 (merge+quant cluster kernel)"
timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; cut -c1-400 $out/${tag}_bench_n1.json
B200_ATTN_FUSED_TAIL=0 timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1_unfused.json 2> $out/${tag}_bench_n1_unfused.err; echo "rc=$?"; cut -c1-400 $out/${tag}_bench_n1_unfused.json
stamp "done"
