#!/bin/bash
# round 2: quick check of the cvtc_bwd builder handshake change (cost-volume parity tests + default bench)
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "cost_volume or costvol or cv" ) > gpurun_out/r2_pytest_23.log 2>&1
tail -2 gpurun_out/r2_pytest_23.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_23.json 2> gpurun_out/r2_bench_23.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_23.json')); r=d['roofline']
    print(d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
    for k,v in list(d.get('kernels',{}).items())[:12]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_23.err
