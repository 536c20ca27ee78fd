#!/bin/bash
# round 2: specialised epilogues in the general conv kernel (stem, strided, 1x1, Cout = 256 layers) + cvtc_bwd L1 prefetch:
# GPU gate, default bench, same-box A/B (option OF_TC_NO_SPEC)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_22.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_22.log
tail -3 gpurun_out/r2_pytest_22.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_22.json > gpurun_out/r2_bench_22.json 2> gpurun_out/r2_bench_22.err
timeout 300 python bench.py --no-cpu-baseline --option OF_TC_NO_SPEC=1 > gpurun_out/r2_bench_22_nospec.json 2>> gpurun_out/r2_bench_22.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_22.json','gpurun_out/r2_bench_22_nospec.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'], r['excluded_layers']['kernels'])
        for k,v in list(d.get('kernels',{}).items())[:12]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_22.err
