#!/bin/bash
# round 2: step time of the tensor-core parity mode (bf16x3) beside fp32 (SIMT) at the benchmark configuration
mkdir -p gpurun_out
timeout 600 python bench.py --conv-mode bf16x3 --no-cpu-baseline --no-extras --steps 10 --warmup 3 > gpurun_out/r2_bench_50_bf16x3.json 2> gpurun_out/r2_bench_50.err
timeout 900 python bench.py --conv-mode fp32 --no-cpu-baseline --no-extras --steps 5 --warmup 3 > gpurun_out/r2_bench_50_fp32.json 2>> gpurun_out/r2_bench_50.err
python - <<'PY'
import json
for m in ("bf16x3","fp32"):
    try:
        d=json.load(open('gpurun_out/r2_bench_50_%s.json'%m)); print(m, d['ms_per_step'], d['value'], d['e2e']['value'], d['dtype'])
    except Exception as e: print(m,'ERR',e)
PY
tail -5 gpurun_out/r2_bench_50.err
