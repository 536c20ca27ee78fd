#!/bin/bash
# round 2, call 7 (fresh container): full GPU gate, smoke, default bench line (both arms), launch list of the step
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_7.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_7.log
tail -6 gpurun_out/r2_pytest_7.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" ) > gpurun_out/r2_smoke_7.log 2>&1
tail -2 gpurun_out/r2_smoke_7.log
timeout 600 python bench.py > gpurun_out/r2_bench_7.json 2> gpurun_out/r2_bench_7.err
echo "bench rc=$?"; head -c 3000 gpurun_out/r2_bench_7.json; tail -3 gpurun_out/r2_bench_7.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_7.json 2> gpurun_out/r2_bench_ref_7.err
echo "ref rc=$?"; head -c 800 gpurun_out/r2_bench_ref_7.json
for fl in smooth iid; do
  timeout 200 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_loss_7_$fl.json 2>> gpurun_out/r2_bench_7.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r2_loss_7_$fl.json')); print('loss $fl', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
except Exception as e: print('ERR', e)
PY
done
timeout 300 python bench.py --workload kernels --steps 30 --no-cpu-baseline > gpurun_out/r2_kernels_7.json 2>> gpurun_out/r2_bench_7.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_kernels_7.json'))
    for k,v in d['roofline']['kernels'].items(): print("%-55s %8.1f us %7.1f GB/s %.3f"%(k,v['us'],v['achieved'],v['frac']))
except Exception as e: print('ERR',e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_train_launches_7.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --graph off > gpurun_out/r2_ncu_7.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r2_train_launches_7.csv
