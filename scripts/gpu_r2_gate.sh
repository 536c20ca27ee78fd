#!/bin/bash
# round 2: full GPU gate + smoke + default bench (both arms) on the current tree
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_46.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_46.log
tail -3 gpurun_out/r2_pytest_46.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" ) > gpurun_out/r2_smoke_46.log 2>&1
tail -1 gpurun_out/r2_smoke_46.log
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_46.json 2> gpurun_out/r2_bench_46.err
timeout 600 python bench.py > gpurun_out/r2_bench_46.json 2>> gpurun_out/r2_bench_46.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_46.json')); r=d['roofline']
    print(d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'], 'loss', r['loss_fused']['frac'], 'cpu', d['cpu_baseline']['value'], d['clocks'])
    print('ref', json.load(open('gpurun_out/r2_bench_ref_46.json'))['value'])
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_46.err
