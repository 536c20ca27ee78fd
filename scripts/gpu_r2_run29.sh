#!/bin/bash
# round 2: wgrad jobs launched as thread-block clusters (L2 sharing of the pixel range's tiles): parity tests, per-layer
# timings and whole step, A/B on the same box against option OF_WG_NO_CLUSTER
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "wgrad or train_step or bench_layer or golden or stem" ) > gpurun_out/r2_pytest_29.log 2>&1
tail -2 gpurun_out/r2_pytest_29.log
echo "== cluster"; timeout 200 python scripts/bench_wgrad_layer.py 2>&1 | grep default
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_29.json 2>> gpurun_out/r2_bench_29.err
timeout 300 python bench.py --no-cpu-baseline --option OF_WG_NO_CLUSTER=1 > gpurun_out/r2_bench_29_nocluster.json 2>> gpurun_out/r2_bench_29.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_29.json','gpurun_out/r2_bench_29_nocluster.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:3]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_29.err
