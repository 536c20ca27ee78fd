#!/bin/bash
# round 2: GPU gate + default bench (residual fusions opt-in, bit-mask act_bwd, warp backward at 2 CTAs / SM, zero K steps skipped in conv_tc3)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_17.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_17.log
tail -3 gpurun_out/r2_pytest_17.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_17.json > gpurun_out/r2_bench_17.json 2> gpurun_out/r2_bench_17.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_17.json',):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'], 'loss', r['loss_fused']['frac'])
        for k,v in d.get('kernels',{}).items(): print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_17.err
