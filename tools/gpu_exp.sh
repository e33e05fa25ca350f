#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: N-GPU run of the final code (both bench arms).
#   GPU_N=2 gpurun --gpus 2 -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
N=${GPU_N:-2}
tag=${GPU_TAG:-r02final}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
trun() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
stamp "bench.py N=$N"
timeout 120 bash -c "$(declare -f trun); trun $N 29522 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "rc=$?"
grep -o '"ms_per_step": [0-9.]*\|"value": [0-9.]*\|"nvls": [a-z]*\|"ranks_sampled_identical_ids": [a-z]*' $out/${tag}_bench_n$N.json | head -6 | tr '\n' ' '; echo
grep -o '"kernel_split_eager.*"comm"' $out/${tag}_bench_n$N.json | cut -c1-900
tail -2 $out/${tag}_bench_n$N.err | cut -c1-200
stamp "reference arm N=$N"
timeout 120 bash -c "$(declare -f trun); trun $N 29524 bench.py --impl reference --gpus $N --steps 2 --warmup 1" > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "rc=$?"; cut -c1-200 $out/${tag}_bench_ref_n$N.json
stamp "done"
