#!/bin/bash
# round 2, call 8: full GPU gate (after the graph-step tolerance fix + the allocation-free entry point), default bench with the
# delay-queued timeline, ncu --set full of the fused loss kernel and of the warp backward
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_8.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_8.log
tail -4 gpurun_out/r2_pytest_8.log
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > gpurun_out/r2_pytest_8_noalloc_alone.log 2>&1
echo "no_alloc alone rc=$?" | tee -a gpurun_out/r2_pytest_8_noalloc_alone.log
tail -2 gpurun_out/r2_pytest_8_noalloc_alone.log
timeout 600 python bench.py --dump-timeline gpurun_out/r2_timeline_8.json > gpurun_out/r2_bench_8.json 2> gpurun_out/r2_bench_8.err
echo "bench rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_bench_8.json'))
    print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel'])
    for k,v in d.get('kernels',{}).items(): print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
except Exception as e: print('ERR', e)
PY
tail -3 gpurun_out/r2_bench_8.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:loss_ws_kernel -c 3 -f -o gpurun_out/r2_loss_ws \
  python bench.py --workload loss --steps 8 --no-graph --no-cpu-baseline > gpurun_out/r2_ncu_loss_8.log 2>&1
echo "ncu loss rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:sample_bwd_kernel -c 14 -f -o gpurun_out/r2_sample_bwd \
  python bench.py --workload kernels --steps 3 --no-cpu-baseline > gpurun_out/r2_ncu_warp_8.log 2>&1
echo "ncu warp rc=$?"
ls -la gpurun_out/*.ncu-rep
