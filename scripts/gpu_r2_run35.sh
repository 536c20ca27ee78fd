#!/bin/bash
# round 2: wgrad bias gradient as SIMT column sums by the epilogue warps (default) against the ones-MMAs (OF_WG_MMA_BIAS)
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "wgrad or train_step or bench_layer or golden or stem or graph or fp32_step or residual" ) > gpurun_out/r2_pytest_35.log 2>&1
tail -3 gpurun_out/r2_pytest_35.log
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_MMA_BIAS=1 NOBIAS 2>&1 | tee gpurun_out/r2_wgrad_layers_35.txt
for v in default OF_WG_MMA_BIAS=1; do
  opt=""; [ "$v" != default ] && opt="--option $v"
  timeout 300 python bench.py --no-cpu-baseline $opt > gpurun_out/r2_bench_35_$v.json 2>> gpurun_out/r2_bench_35.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_bench_35_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'])
        for k,v in list(d.get('kernels',{}).items())[:12]:
            if 'wgrad' in k: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_35.err
