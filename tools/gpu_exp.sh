#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: the 8-GPU session (every second costs 8x).
#   gpurun --gpus 8 -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
tag=${GPU_TAG:-r02i}
N=${GPU_N:-8}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
trun() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
brief() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith("{")][-1])
    ks=d.get("kernel_split_eager",{})
    print("  ", sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"],3), "tok/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "splits", d["config"]["num_kv_splits"], "attn us", d["roofline"]["us_per_launch"], "comm", d.get("comm"))
    print("     ", {k:(v["launches_per_step"], v["us_per_launch"]) for k,v in ks.items()})
except Exception as e: print("ERR", sys.argv[1], e)
PY
}
nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
stamp "bench.py N=$N"
timeout 100 bash -c "$(declare -f trun); trun $N 29522 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "rc=$?"; brief $out/${tag}_bench_n$N.json; tail -2 $out/${tag}_bench_n$N.err | cut -c1-200
stamp "all-reduce latency N=$N (NVLS on)"
B200_COMM_NVLS=1 timeout 120 bash -c "$(declare -f trun); trun $N 29521 tools/bench_allreduce.py --ar-only --out $out/${tag}_bench_allreduce_tp$N.json" 2>&1 | grep "^{" | cut -c1-200
stamp "bench.py N=$N NVLS"
B200_COMM_NVLS=1 timeout 100 bash -c "$(declare -f trun); trun $N 29523 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n${N}_nvls.json 2> $out/${tag}_bench_n${N}_nvls.err; echo "rc=$?"; brief $out/${tag}_bench_n${N}_nvls.json
stamp "bench.py N=$N PDL up to 512 tokens"
B200_PDL_MAX_M=512 timeout 100 bash -c "$(declare -f trun); trun $N 29524 bench.py --gpus $N --no-cpu-baseline" > $out/${tag}_bench_n${N}_pdl.json 2> $out/${tag}_bench_n${N}_pdl.err; echo "rc=$?"; brief $out/${tag}_bench_n${N}_pdl.json
stamp "bench.py N=4"
timeout 100 bash -c "$(declare -f trun); trun 4 29525 bench.py --gpus 4 --no-cpu-baseline" > $out/${tag}_bench_n4.json 2> $out/${tag}_bench_n4.err; echo "rc=$?"; brief $out/${tag}_bench_n4.json
stamp "nvls + ipc tests world $N"
B200_COMM_NVLS=1 timeout 120 python -m pytest tests/test_comm_gpu.py -m gpu -q -x -k "multi_gpu and $N" -rs > $out/${tag}_pytest_comm.log 2>&1; echo "rc=$?"; tail -4 $out/${tag}_pytest_comm.log | cut -c1-300
stamp "cfg 5: Llama-3.1-70B FP8 TP=$N, SafeKV multi-turn"
timeout 150 bash -c "$(declare -f trun); trun $N 29526 tools/bench_safekv_multiturn.py --users 32 --turns 3 --decode-steps 4 --out $out/${tag}_safekv_llama70b_tp$N.json" 2>&1 | grep "^{" | cut -c1-600
stamp "done"
