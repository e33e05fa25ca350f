mkdir -p gpurun_out
timeout 600 python tools/bench_decode_model.py --bs 1,8,16,32,64 --ctx 1024 --out gpurun_out/tmp_model.json 2>&1 | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    r=json.loads(l); print('bs',r['bs'],'ctx',r['ctx'],'ms',r['ms_per_step'],'tok/s',r['tokens_per_s'], {k:(v['us_total']) for k,v in r['kernel_split_eager'].items()})
"
timeout 600 python tools/bench_gemm_decode.py --ms 1,16,32 --kinds rowcol --variants auto --out gpurun_out/tmp.json 2>&1 | cut -c1-200
