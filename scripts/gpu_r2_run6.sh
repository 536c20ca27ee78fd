#!/bin/bash
# round 2, call 6: loss kernel v2.2 (64-px tiles, builder-released slots, slow polling) + BN train mode + warp bwd wide tiles
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q -k "loss or graph_step or golden or warp or bn_train" ) > gpurun_out/r2_pytest_6.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_6.log
tail -12 gpurun_out/r2_pytest_6.log
for v in "" g13 g21; do
  for fl in smooth iid; do
    OF_PROFILE_LIB=$v timeout 200 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_lf22_${v:-g17}_$fl.json 2>> gpurun_out/r2_bench_6.err
  done
done
for dbg in 2 3 10 11 20 22; do
  OF_PROFILE_LIB=1 OF_LF2_DBG=$dbg timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf22_dbg$dbg.json 2>> gpurun_out/r2_bench_6.err
done
timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_lf22_c5.json 2>> gpurun_out/r2_bench_6.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf22_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 300 python bench.py --workload kernels --steps 30 --no-cpu-baseline > gpurun_out/r2_kernels_6.json 2>> gpurun_out/r2_bench_6.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_6.json'))
    for k,v in d['roofline']['kernels'].items(): print("%-55s %8.1f us %7.1f GB/s %.3f"%(k,v['us'],v['achieved'],v['frac']))
except Exception as e: print('ERR',e)
PY
tail -3 gpurun_out/r2_bench_6.err
