#!/bin/bash
# One GPU-box pass: parity tests, bench (both arms), ncu launch list + full capture of the dominant kernel.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode_mma_kernel -s 40 -c 2 -f -o gpurun_out/decode_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_n1.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_extend_kernel -s 30 -c 1 -f -o gpurun_out/extend_full python tools/bench_extend.py --out gpurun_out/tmp_ext.json > gpurun_out/ncu_extend.log 2>&1; echo "ncu extend rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fp8_block_gemm_kernel -s 4 -c 1 -f -o gpurun_out/blockgemm_full python tools/bench_gemm.py --ms 8192 --out gpurun_out/tmp_gemm.json > gpurun_out/ncu_blockgemm.log 2>&1; echo "ncu block gemm rc=$?"
