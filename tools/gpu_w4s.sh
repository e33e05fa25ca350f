#!/bin/bash
# forced split-K factors of the swap-AB W4A16 kernels on the TP=2 shapes
set -u
mkdir -p gpurun_out
for s in ${W4_SPLITS:-0 2 3 4 6}; do
echo "== SPLITS=$s"
B200_W4_SPLITS=$s timeout 600 python tools/bench_w4a16.py --tp tp2 --ms ${W4_MS:-32,64,128} --out gpurun_out/bench_w4a16_s$s.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); print(r['M'],r['N'],r['K'],'gemm',r['gemm_tcgen05_us'],r['dispatch_gbs'],'GB/s ref',r['ref_dequant_matmul_us'])
"
done
