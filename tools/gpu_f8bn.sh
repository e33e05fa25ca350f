#!/bin/bash
# FP8 per-channel GEMM, weight-streaming mid-M regime: 128- vs 256-wide tiles (B200_GEMM_BN) vs the dispatch rule
set -u
mkdir -p gpurun_out
for bn in ${F8_BNS:-0 128 256}; do
echo "== BN=$bn"
B200_GEMM_BN=$bn timeout 600 python tools/bench_gemm_decode.py --shapes ${F8_SHAPES:-28672x4096,6144x4096,4096x14336} --ms ${F8_MS:-48,64,128,192,256,384,512} --kinds rowcol,torch_rowwise --out gpurun_out/bench_gemm_bn$bn.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l)
        if 'M' in r: print(r['M'],r['N'],r['K'],'ours',r.get('rowcol_us'),'lt',r.get('torch_rowwise_us'),'floor',r.get('hbm_floor_us'))
"
done
