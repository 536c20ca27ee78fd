#!/bin/bash
# round 2: fused bias gradient also for Cout = 256 (transposed bias MMAs need 32 TMEM columns) against the colsum pass
mkdir -p gpurun_out
OF_EXP_FUSE_DB256=1 timeout 200 python -m pytest tests -m gpu -x -q -k "wgrad" 2>&1 | tail -2
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_44_default.json 2>> gpurun_out/r2_bench_44.err
OF_EXP_FUSE_DB256=1 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_44_fuse256.json 2>> gpurun_out/r2_bench_44.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_bench_44_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:12]:
            if 'wgrad' in k or 'colsum' in k: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_44.err
