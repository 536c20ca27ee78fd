#!/bin/bash
# round 2: transposed bias MMAs (OF_WG_BIAS_T) on the whole step, A/B on one box
mkdir -p gpurun_out
for v in default OF_WG_BIAS_T=1 default2 OF_WG_BIAS_T=1,x; do
  opt=""; case $v in OF_WG_BIAS_T*) opt="--option OF_WG_BIAS_T=1";; esac
  timeout 300 python bench.py --no-cpu-baseline $opt > gpurun_out/r2_bench_42_$v.json 2>> gpurun_out/r2_bench_42.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_bench_42_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:12]:
            if 'wgrad' in k: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_42.err
