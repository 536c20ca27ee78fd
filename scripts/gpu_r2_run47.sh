#!/bin/bash
# round 2: default bench line with the e2e loss read one step behind
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r2_bench_47.json 2> gpurun_out/r2_bench_47.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_47.json')); r=d['roofline']
print(d['ms_per_step'], d['value'], 'e2e', d['e2e'], 'frac', r['frac'])
PY
tail -3 gpurun_out/r2_bench_47.err
timeout 300 python -m pytest tests -m gpu -x -q -k "bench" 2>&1 | tail -2
