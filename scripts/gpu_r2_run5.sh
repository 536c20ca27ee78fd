#!/bin/bash
# round 2, call 5: loss kernel v2.1 parity, gather-warp A/B, dissection, ncu capture
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q -x -k "loss or graph_step or golden or warp" ) > gpurun_out/r2_pytest_5.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_5.log
tail -6 gpurun_out/r2_pytest_5.log
for v in "" g13 g21; do
  for fl in smooth iid; do
    OF_PROFILE_LIB=$v timeout 200 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_lf21_${v:-g17}_$fl.json 2>> gpurun_out/r2_bench_5.err
  done
done
for dbg in 2 3 10 11 20 22; do
  OF_PROFILE_LIB=1 OF_LF2_DBG=$dbg timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf21_dbg$dbg.json 2>> gpurun_out/r2_bench_5.err
done
timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_lf21_c5.json 2>> gpurun_out/r2_bench_5.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf21_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/r2_bench_5.err
timeout 300 python bench.py --workload loss --flows smooth --steps 8 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2_ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:loss_ws -s 4 -c 2 -o gpurun_out/r2_lf21_prof -f python bench.py --workload loss --flows smooth --steps 8 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2_ncu.log 2>&1
tail -3 gpurun_out/r2_ncu.log
