#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python tools/bench_decode_model.py --bs 1,8,16,32,64,128,256 --ctx 1024,4096 --out gpurun_out/decode_model_fp8_tp1.json 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print('bs',r['bs'],'ctx',r['ctx'],'ms',r['ms_per_step'],'tok/s',r['tokens_per_s'],'frac',r['frac_of_hbm_floor'])
"
