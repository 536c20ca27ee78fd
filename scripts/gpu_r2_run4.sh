#!/bin/bash
# round 2, call 4: (a) loss + graph tests, (b) loss kernel dissection with the profile-hooks library,
# (c) train bench graph on/off + per-launch timeline dump
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q -x -k "loss or graph_step" ) > gpurun_out/r2_pytest_4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_4.log
tail -8 gpurun_out/r2_pytest_4.log
export OF_PROFILE_LIB=1
for dbg in 0 2 3 8 10 11 20 22; do
  OF_LF2_DBG=$dbg timeout 200 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline > gpurun_out/r2_lf2_dbg$dbg.json 2>> gpurun_out/r2_bench_4.err
done
unset OF_PROFILE_LIB
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf2_dbg*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us')
    except Exception as e: print(f, 'ERR', e)
PY
timeout 600 python bench.py --graph off --no-cpu-baseline --dump-timeline gpurun_out/r2_timeline.json > gpurun_out/r2_bench_4_eager.json 2>> gpurun_out/r2_bench_4.err
timeout 600 python bench.py --graph on --no-cpu-baseline --no-extras > gpurun_out/r2_bench_4_graph.json 2>> gpurun_out/r2_bench_4.err
timeout 600 python bench.py --graph off --no-cpu-baseline --no-extras --batch 4 --height 192 --width 640 --steps 50 > gpurun_out/r2_bench_4_small_eager.json 2>> gpurun_out/r2_bench_4.err
timeout 600 python bench.py --graph on --no-cpu-baseline --no-extras --batch 4 --height 192 --width 640 --steps 50 > gpurun_out/r2_bench_4_small_graph.json 2>> gpurun_out/r2_bench_4.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_4_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step'],3),'ms', d['e2e']['value'], d['config'].get('launch'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/r2_bench_4.err
