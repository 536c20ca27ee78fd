#!/bin/bash
# round 2 final evidence: GPU gate, smoke, default bench (both arms), launch list of the step, ncu --set full of the top kernels
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_25.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_25.log
tail -3 gpurun_out/r2_pytest_25.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" ) > gpurun_out/r2_smoke_25.log 2>&1
tail -1 gpurun_out/r2_smoke_25.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_25.json > gpurun_out/r2_bench_25.json 2> gpurun_out/r2_bench_25.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_25.json')); r=d['roofline']
    print(d['ms_per_step'], d['value'], d['e2e']['value'], 'frac', r['frac'], 'all', r['all_tc_launches']['frac'], 'loss', r['loss_fused']['frac'])
    for k,v in d.get('kernels',{}).items(): print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_25.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_train_launches_25.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --graph off > gpurun_out/r2_ncu_25.log 2>&1
echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_tc3_kernel|conv_tc_wgrad_kernel|cvtc_bwd_kernel" -c 12 -f -o gpurun_out/r2_top_kernels_spec \
  python scripts/prof_r2_kernels.py > gpurun_out/r2_ncu_top_25.log 2>&1
echo "ncu top rc=$?"; ls -la gpurun_out/r2_top_kernels_spec.ncu-rep
