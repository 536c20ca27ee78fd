#!/bin/bash
# round 2, call 13: GPU gate (driver form) + smoke + default bench line with the SURVEY-8d conv figure (stem / heads listed apart),
# timeline dump, launch list under ncu
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_13.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_13.log
tail -3 gpurun_out/r2_pytest_13.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" ) > gpurun_out/r2_smoke_13.log 2>&1
tail -2 gpurun_out/r2_smoke_13.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_13.json > gpurun_out/r2_bench_13.json 2> gpurun_out/r2_bench_13.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_13.json')); r=d['roofline']
    print(d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'], r['excluded_layers']['kernels'])
    for k,v in d.get('kernels',{}).items(): print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_13.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_13.json 2>> gpurun_out/r2_bench_13.err
head -c 300 gpurun_out/r2_bench_ref_13.json; echo
