#!/bin/bash
# round 2: residual-block fusions (conv2 + add + ReLU in one epilogue, ReLU masks and the decoder gradient in the dgrad epilogues):
# GPU gate, default bench, A/B with the fusions switched off, warp backward occupancy variants
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_15.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_15.log
tail -3 gpurun_out/r2_pytest_15.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_15.json > gpurun_out/r2_bench_15.json 2> gpurun_out/r2_bench_15.err
OF_NO_POST_ADD=1 OF_NO_MASK_FUSE=1 timeout 300 python bench.py --no-extras --no-cpu-baseline > gpurun_out/r2_bench_15_nofuse.json 2>> gpurun_out/r2_bench_15.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_15.json','gpurun_out/r2_bench_15_nofuse.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'])
        for k,v in d.get('kernels',{}).items(): print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_15.err
for v in "" wb3 wb4; do
  OF_PROFILE_LIB=$v timeout 200 python bench.py --workload kernels --steps 20 --no-cpu-baseline > gpurun_out/r2_kernels_15_${v:-wb1}.json 2>> gpurun_out/r2_bench_15.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_15_${v:-wb1}.json'))
    for k,v in d['roofline']['kernels'].items():
        if 'K3' in k: print("${v:-wb1} %-55s %8.1f us %.3f"%(k,v['us'],v['frac']))
except Exception as e: print('ERR',e)
PY
done
