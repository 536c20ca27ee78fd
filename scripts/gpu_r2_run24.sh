#!/bin/bash
# round 2: residual fusions with their own epilogue specialisations - A/B on the same box, fusion + conv parity tests, C5-shaped step
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_b2_gpu_conv_tc.py tests/test_b3_gpu_net.py -m gpu -x -q ) > gpurun_out/r2_pytest_24.log 2>&1
tail -2 gpurun_out/r2_pytest_24.log
for v in "" "--fuse-residual"; do
  timeout 300 python bench.py --no-cpu-baseline $v > gpurun_out/r2_bench_24${v:+_fused}.json 2>> gpurun_out/r2_bench_24.err
done
timeout 300 python bench.py --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_bench_24_c5.json 2>> gpurun_out/r2_bench_24.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_24.json','gpurun_out/r2_bench_24_fused.json','gpurun_out/r2_bench_24_c5.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:12]: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_24.err
