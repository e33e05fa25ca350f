#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_kv_gpu.py -x -q > gpurun_out/pytest_fp8kv.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_fp8kv.log | cut -c1-300
timeout 600 python tools/sweep_decode.py --tune --kv-dtype fp8 --out gpurun_out/sweep_decode_fp8kv.json 2>&1 | cut -c1-300
timeout 600 python tools/sweep_decode.py --tune --out gpurun_out/sweep_decode_bf16.json 2>&1 | head -3 | cut -c1-300
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_fp8kv -s 8 -c 1 -f -o gpurun_out/decode_fp8kv_full python tools/sweep_decode.py --tune --kv-dtype fp8 --out gpurun_out/tmp.json > gpurun_out/ncu_fp8kv.log 2>&1; echo "ncu rc=$?"
