#!/bin/bash
# round 2, call 10: GPU gate without -x (see every failure), loss kernel isolation A/B (gather wait mode, scalar / packed build)
# and the profile-hooks dissection of the current kernel
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/r2_pytest_10.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_10.log
tail -6 gpurun_out/r2_pytest_10.log
for v in "" w0 w2 w1sc w0scnf; do
  OF_PROFILE_LIB=$v timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf24_${v:-w1}_smooth.json 2>> gpurun_out/r2_bench_10.err
done
for dbg in 1 2 3 4 12 20 8; do
  OF_PROFILE_LIB=1 OF_LF2_DBG=$dbg timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf24_dbg$dbg.json 2>> gpurun_out/r2_bench_10.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf24_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 300 python bench.py --workload kernels --steps 30 --no-cpu-baseline > gpurun_out/r2_kernels_10.json 2>> gpurun_out/r2_bench_10.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_10.json'))
    for k,v in d['roofline']['kernels'].items(): print("%-55s %8.1f us %7.1f GB/s %.3f"%(k,v['us'],v['achieved'],v['frac']))
except Exception as e: print('ERR',e)
PY
tail -3 gpurun_out/r2_bench_10.err
