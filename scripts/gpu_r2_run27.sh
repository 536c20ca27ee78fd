#!/bin/bash
# round 2: wgrad bias MMAs rotated per K tile (+ interleaved split-K tiles): parity tests, A/B on the same box against -DWG_BIAS_PER_RANGE
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "wgrad or train_step or bench_layer or golden" ) > gpurun_out/r2_pytest_27.log 2>&1
tail -2 gpurun_out/r2_pytest_27.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_27.json 2>> gpurun_out/r2_bench_27.err
OF_PROFILE_LIB=biasrange timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_27_biasrange.json 2>> gpurun_out/r2_bench_27.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_27.json','gpurun_out/r2_bench_27_biasrange.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:4]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_27.err
