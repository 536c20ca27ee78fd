#!/bin/bash
# round 2: wgrad dynamic K-tile scheduling (claims from per-job counters) against the static split-K ranges:
# parity tests, per-layer timings, whole step A/B on the same box
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "wgrad or train_step or bench_layer or golden or stem or graph" ) > gpurun_out/r2_pytest_31.log 2>&1
tail -2 gpurun_out/r2_pytest_31.log
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_STATIC=1 OF_WG_CLAIM=1 OF_WG_CLAIM=8 2>&1 | tee gpurun_out/r2_wgrad_layers_31.txt
for v in default OF_WG_STATIC=1 OF_WG_CLAIM=1 OF_WG_CLAIM=2; do
  opt=""; [ "$v" != default ] && opt="--option $v"
  timeout 300 python bench.py --no-cpu-baseline $opt > gpurun_out/r2_bench_31_$v.json 2>> gpurun_out/r2_bench_31.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_bench_31_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:3]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_31.err
