#!/bin/bash
# Scratch experiment list of a GPU session (rewritten per session; results land in gpurun_out/, what is
# worth keeping is copied to profiles/ by hand).  Default content: the N-GPU run of both bench arms.
#   gpurun --gpus N -- env GPU_N=N bash tools/gpu_exp.sh        (1-GPU verification: tools/gpu_session.sh verify)
set -u
out=gpurun_out
N=${GPU_N:-2}
tag=${GPU_TAG:-r02final}
mkdir -p $out
trun() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
timeout 120 bash -c "$(declare -f trun); trun $N 29522 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench rc=$?"
cut -c1-400 $out/${tag}_bench_n$N.json
timeout 120 bash -c "$(declare -f trun); trun $N 29524 bench.py --impl reference --gpus $N --steps 2 --warmup 1" > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "reference rc=$?"
cut -c1-200 $out/${tag}_bench_ref_n$N.json
