#!/bin/bash
# round 2, call 11: GPU gate (driver form, -x) twice, loss kernel skeleton dissection (profile-hooks build, OF_LF2_DBG bits)
mkdir -p gpurun_out
for rep in a b; do
  ( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_11$rep.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/r2_pytest_11$rep.log
  tail -3 gpurun_out/r2_pytest_11$rep.log
done
timeout 100 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf25_default.json 2>> gpurun_out/r2_bench_11.err
for dbg in 3 11 67 2 10 66 1 4 12 20; do
  OF_PROFILE_LIB=1 OF_LF2_DBG=$dbg timeout 60 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf25_dbg$dbg.json 2>> gpurun_out/r2_bench_11.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf25_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/r2_bench_11.err
