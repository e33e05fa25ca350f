#!/bin/bash
# One parametrised runner for every GPU session (run on the B200 box through gpurun):
#   tools/gpu_session.sh verify            parity suite + bench.py (both arms) + ncu launch list + ncu --set full of the attention kernel
#   tools/gpu_session.sh tests [pytest args]   parity suite only (extra args go to pytest)
#   tools/gpu_session.sh bench [bench args]    bench.py only
#   tools/gpu_session.sh tp N [bench args]     N-GPU: communicator tests + bench.py under torchrun
#   tools/gpu_session.sh run <cmd...>          any command, logged
# Everything lands under gpurun_out/ (scratch); summaries worth keeping are copied to profiles/ by hand.
set -u
mode=${1:-verify}; shift || true
tag=${GPU_TAG:-r02}
out=gpurun_out
mkdir -p $out
torchrun_n() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
pytest_gpu() { timeout 1500 python -m pytest tests -m gpu -q "$@" > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $out/${tag}_pytest_gpu.log | cut -c1-400; }
bench_one() { timeout 900 python bench.py "$@" > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "bench rc=$?"; cut -c1-2500 $out/${tag}_bench_n1.json; tail -5 $out/${tag}_bench_n1.err; }
case "$mode" in
  verify)
    nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_gpu.txt 2>&1
    pytest_gpu -x
    bench_one "$@"
    timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2>$out/${tag}_bench_reference.err; echo "reference rc=$?"; cut -c1-700 $out/${tag}_bench_reference.json
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
    python tools/ncu_summarize.py launches $out/${tag}_launches.csv $out/${tag}_ncu_launches.json | tail -30
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode_mma -s 40 -c 2 -f -o $out/${tag}_decode_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
    python tools/ncu_summarize.py full $out/${tag}_decode_full.ncu-rep $out/${tag}_ncu_decode_full_summary.json | tail -40
    ;;
  tests) pytest_gpu "$@" ;;
  bench) bench_one "$@" ;;
  tp)
    n=$1; shift
    nvidia-smi topo -m > $out/${tag}_topo_n$n.txt 2>&1
    timeout 900 python -m pytest tests/test_comm_gpu.py -m gpu -x -q > $out/${tag}_pytest_comm_n$n.log 2>&1; echo "comm tests rc=$?"; tail -3 $out/${tag}_pytest_comm_n$n.log
    timeout 900 bash -c "$(declare -f torchrun_n); torchrun_n $n 29511 bench.py --gpus $n $*" > $out/${tag}_bench_n$n.json 2> $out/${tag}_bench_n$n.err; echo "bench n$n rc=$?"; cut -c1-2500 $out/${tag}_bench_n$n.json; tail -5 $out/${tag}_bench_n$n.err
    ;;
  run) "$@" 2>&1 | tee $out/${tag}_run.log | tail -60 ;;
  *) echo "unknown mode $mode"; exit 2 ;;
esac
