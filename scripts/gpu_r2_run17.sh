#!/bin/bash
# round 2: single-thread warp roles guarded by elect.sync instead of lane == 0 (no per-instruction waterfall loop around
# tcgen05.mma / TMA): GPU gate, default bench, same-box A/B against the lane-0 build, per-layer conv timings
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_18.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_18.log
tail -3 gpurun_out/r2_pytest_18.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_18.json > gpurun_out/r2_bench_18.json 2> gpurun_out/r2_bench_18.err
OF_PROFILE_LIB=lane0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_18_lane0.json 2>> gpurun_out/r2_bench_18.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_18.json','gpurun_out/r2_bench_18_lane0.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'])
        for k,v in list(d.get('kernels',{}).items())[:8]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_18.err
echo "== elect"; timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -9
echo "== lane0"; OF_PROFILE_LIB=lane0 timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -9
