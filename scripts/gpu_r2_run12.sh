#!/bin/bash
# round 2, call 12: GPU gate (driver form), loss kernel with the watcher relay (A/B against all-CTA polling), dissection,
# default bench line, C5-shaped train step (configs[4]: batch 4 at 1024x2048)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_12.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_12.log
tail -3 gpurun_out/r2_pytest_12.log
for v in "" norelay; do
  for fl in smooth iid; do
    OF_PROFILE_LIB=$v timeout 100 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_lf26_${v:-relay}_$fl.json 2>> gpurun_out/r2_bench_12.err
  done
done
for dbg in 3 2 4; do
  OF_PROFILE_LIB=1 OF_LF2_DBG=$dbg timeout 60 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf26_dbg$dbg.json 2>> gpurun_out/r2_bench_12.err
done
timeout 100 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_lf26_c5_smooth.json 2>> gpurun_out/r2_bench_12.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf26_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 600 python bench.py > gpurun_out/r2_bench_12.json 2>> gpurun_out/r2_bench_12.err
timeout 600 python bench.py --batch 4 --height 1024 --width 2048 --no-cpu-baseline > gpurun_out/r2_bench_12_c5.json 2>> gpurun_out/r2_bench_12.err
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_12.json','gpurun_out/r2_bench_12_c5.json'):
    try:
        d=json.load(open(f))
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline'].get('loss_fused',{}).get('flows'))
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_12.err
