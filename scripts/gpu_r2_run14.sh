#!/bin/bash
# round 2: warp backward occupancy variants (launch bounds 256 x {1 (=2 resident), 3, 4}) on the kernels leg
mkdir -p gpurun_out
for v in "" wb3 wb4; do
  OF_PROFILE_LIB=$v timeout 200 python bench.py --workload kernels --steps 20 --no-cpu-baseline > gpurun_out/r2_kernels_14_${v:-wb1}.json 2>> gpurun_out/r2_bench_14.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_14_${v:-wb1}.json'))
    for k,v in d['roofline']['kernels'].items():
        if 'bwd' in k: print("${v:-wb1} %-55s %8.1f us %.3f"%(k,v['us'],v['frac']))
except Exception as e: print('ERR',e)
PY
done
tail -2 gpurun_out/r2_bench_14.err
