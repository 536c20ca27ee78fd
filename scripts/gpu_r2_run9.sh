#!/bin/bash
# round 2, call 9: full GPU gate with the reworked loss kernel (packed build, mbarrier parking) + BN test rewrite + no-alloc entry;
# loss kernel A/B over builder / gather warp splits; default bench line
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_9.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_9.log
tail -4 gpurun_out/r2_pytest_9.log
for v in "" b6g19 b10g15 b12g13; do
  for fl in smooth iid; do
    OF_PROFILE_LIB=$v timeout 200 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_lf23_${v:-b8g17}_$fl.json 2>> gpurun_out/r2_bench_9.err
  done
done
timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --option OF_LF_V1=1 > gpurun_out/r2_lf23_v1_smooth.json 2>> gpurun_out/r2_bench_9.err
timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_lf23_c5_smooth.json 2>> gpurun_out/r2_bench_9.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf23_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 600 python bench.py > gpurun_out/r2_bench_9.json 2>> gpurun_out/r2_bench_9.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_9.json'))
    print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['loss_fused']['flows'])
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_9.err
