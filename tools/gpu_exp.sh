#!/bin/bash
# Scratch experiment list of the CURRENT GPU session.  This one: N-GPU check of the bench teardown fix
# + NVLS test; every step under a short timeout.   gpurun --gpus N -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
tag=${GPU_TAG:-r02e}
N=${GPU_N:-2}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
trun() { port=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
stamp "bench.py N=$N"
timeout 200 bash -c "$(declare -f trun); N=$N; trun 29522 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "rc=$?"; cut -c1-300 $out/${tag}_bench_n$N.json; grep -o '"kernel_split_eager.*' $out/${tag}_bench_n$N.json | cut -c1-1500; tail -2 $out/${tag}_bench_n$N.err | cut -c1-200
stamp "nvls test"
timeout 150 python -m pytest tests/test_comm_gpu.py -m gpu -q -x -k "nvls and $N" -rs > $out/${tag}_pytest_nvls.log 2>&1; echo "rc=$?"; tail -4 $out/${tag}_pytest_nvls.log | cut -c1-300
stamp "reference arm N=$N"
timeout 200 bash -c "$(declare -f trun); N=$N; trun 29524 bench.py --impl reference --gpus $N --steps 2 --warmup 1" > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "rc=$?"; cut -c1-400 $out/${tag}_bench_ref_n$N.json
stamp "done"
