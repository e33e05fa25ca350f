#!/bin/bash
# Scratch experiment list of the CURRENT GPU session (rewritten per session; results land in
# gpurun_out/, what is worth keeping is copied to profiles/ by hand).  Run: gpurun -- bash tools/gpu_exp.sh
set -u
out=gpurun_out
tag=${GPU_TAG:-r02c}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ks=d.get("kernel_split_eager",{})
    print(sys.argv[1], "ms/step", round(d["ms_per_step"],3), "tok/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "attn us", d["roofline"]["us_per_launch"], "frac_step", round(d["roofline_step"]["frac"],3))
    print("   ", {k:(v["launches_per_step"], v["us_per_launch"]) for k,v in ks.items()})
except Exception as e: print("ERR", sys.argv[1], e)
PY
}

stamp "pytest gpu"
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 $out/${tag}_pytest_gpu.log | cut -c1-300

stamp "bench default (merge+quant launch at bs 64)"
timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "rc=$?"
show $out/${tag}_bench_default.json; tail -2 $out/${tag}_bench_default.err

stamp "bench unfused"
B200_ATTN_FUSED_TAIL=0 timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_unfused.json 2> $out/${tag}_bench_unfused.err; echo "rc=$?"
show $out/${tag}_bench_unfused.json

stamp "small-batch sweep: tail as one launch"
B200_ATTN_TAIL_IN_KERNEL_MAX_BS=0 timeout 600 python tools/bench_decode_model.py --bs 1,4,8,16,32 --ctx 1024,4096 --out $out/${tag}_decode_tail_launch.json 2>&1 | cut -c1-200
stamp "small-batch sweep: tail in kernel"
B200_ATTN_TAIL_IN_KERNEL_MAX_BS=100000 timeout 600 python tools/bench_decode_model.py --bs 1,4,8,16,32 --ctx 1024,4096 --out $out/${tag}_decode_tail_kernel.json 2>&1 | cut -c1-200
stamp "small-batch sweep: unfused"
B200_ATTN_FUSED_TAIL=0 timeout 600 python tools/bench_decode_model.py --bs 1,4,8,16,32 --ctx 1024,4096 --out $out/${tag}_decode_tail_off.json 2>&1 | cut -c1-200
stamp "done"
