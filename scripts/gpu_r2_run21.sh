#!/bin/bash
# round 2: conv_tc3 with compile-time specialised epilogues: GPU gate, default bench, same-box A/B (option OF_T3_NO_SPEC), per-layer timings
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_21.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_21.log
tail -3 gpurun_out/r2_pytest_21.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_21.json > gpurun_out/r2_bench_21.json 2> gpurun_out/r2_bench_21.err
timeout 300 python bench.py --no-cpu-baseline --option OF_T3_NO_SPEC=1 > gpurun_out/r2_bench_21_nospec.json 2>> gpurun_out/r2_bench_21.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_21.json','gpurun_out/r2_bench_21_nospec.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'])
        for k,v in list(d.get('kernels',{}).items())[:6]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_21.err
timeout 120 python scripts/prof_conv_pair.py 2>&1 | tail -3
