#!/bin/bash
# round 2, call 3: new warp-specialised loss kernel: parity, then A/B timing against the first-generation kernel
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q -k "loss or golden or smoke" ) > gpurun_out/r2_pytest_3.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_3.log
tail -15 gpurun_out/r2_pytest_3.log
for fl in smooth iid; do
  timeout 300 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline > gpurun_out/r2_lf2_$fl.json 2>> gpurun_out/r2_bench_3.err
  timeout 300 python bench.py --workload loss --flows $fl --steps 400 --no-cpu-baseline --option OF_LF_V1=1 > gpurun_out/r2_lf1_$fl.json 2>> gpurun_out/r2_bench_3.err
done
timeout 300 python bench.py --workload loss --flows smooth --steps 400 --no-cpu-baseline --batch 4 --height 1024 --width 2048 > gpurun_out/r2_lf2_c5.json 2>> gpurun_out/r2_bench_3.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_lf?_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/r2_bench_3.err
