#!/bin/bash
# round 2: warp backward same-row fast path: GPU gate, default bench (kernel shares), kernels leg
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_28.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_28.log
tail -3 gpurun_out/r2_pytest_28.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_28.json 2> gpurun_out/r2_bench_28.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_28.json')); r=d['roofline']
    print(d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
    for k,v in list(d.get('kernels',{}).items())[:12]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_28.err
timeout 300 python bench.py --workload kernels --steps 20 --no-cpu-baseline > gpurun_out/r2_kernels_28.json 2>> gpurun_out/r2_bench_28.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_28.json'))
    for k,v in d['roofline']['kernels'].items():
        if 'K3' in k: print("%-55s %8.1f us %.3f"%(k,v['us'],v['frac']))
except Exception as e: print('ERR',e)
PY
