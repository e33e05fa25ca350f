#!/bin/bash
# Scratch experiment list of the CURRENT GPU session: weight prefetch before the dependency wait in
# gemm_tn (PDL), smoke(), PDL limit at large batches.
set -u
out=gpurun_out
tag=${GPU_TAG:-r02j}
mkdir -p $out
t0=$(date +%s)
stamp() { echo "== $1 (+$(( $(date +%s) - t0 )) s)"; }
stamp "smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | cut -c1-300
stamp "pytest gpu"
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $out/${tag}_pytest_gpu.log | cut -c1-300
stamp "mid-M sweep with PDL (default limits)"
timeout 400 python tools/bench_gemm_decode.py --ms 32,48,64,128 --variants auto,big,s2,s4,s8 --kinds rowcol,torch_rowwise --out $out/${tag}_gemm_midm.json 2>&1 | cut -c1-420
timeout 300 python tools/bench_gemm_decode.py --model llama8b_tp8 --ms 64 --variants auto,s4,s8 --kinds rowcol --out $out/${tag}_gemm_midm_tp8.json 2>&1 | cut -c1-300
stamp "bench x2"
for i in 1 2; do timeout 300 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1_$i.json 2> $out/${tag}_bench_n1_$i.err; echo "rc=$?"; cut -c1-330 $out/${tag}_bench_n1_$i.json | grep -o '"ms_per_step": [0-9.]*'; done
stamp "large batches: PDL limit 64 vs 512"
timeout 300 python tools/bench_decode_model.py --bs 32,128,256 --ctx 1024,4096 --out $out/${tag}_decode_pdl64.json 2>&1 | grep -o '"bs": [0-9]*, "ctx": [0-9]*\|"ms_per_step": [0-9.]*' | paste - - 
B200_PDL_MAX_M=512 timeout 300 python tools/bench_decode_model.py --bs 32,128,256 --ctx 1024,4096 --out $out/${tag}_decode_pdl512.json 2>&1 | grep -o '"bs": [0-9]*, "ctx": [0-9]*\|"ms_per_step": [0-9.]*' | paste - -
stamp "done"
