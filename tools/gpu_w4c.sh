#!/bin/bash
for d in 0 4 16 20 28 31 30 2 3; do
echo "== DEBUG=$d"
B200_W4_DEBUG=$d timeout 300 python tools/bench_w4a16.py --tp tp2 --ms 32,4096 --out gpurun_out/tmp.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l)
        if r['N']==3584: print(r['M'],r['N'],r['K'],'gemm',r['gemm_tcgen05_us'])
"
done
