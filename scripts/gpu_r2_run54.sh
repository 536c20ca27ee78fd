#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_b6_gpu_bf16x3.py tests/test_b4_bench_shapes.py -m gpu -x -q -k "bf16x3 or split" 2>&1 | tail -3
timeout 300 python scripts/timeline_mode.py bf16x3 2>&1 | tail -30 | head -12 | tee gpurun_out/r2_timeline_bf16x3.txt
timeout 600 python bench.py --conv-mode bf16x3 --no-cpu-baseline --no-extras --steps 10 --warmup 3 > gpurun_out/r2_bench_54_bf16x3.json 2> gpurun_out/r2_bench_54.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_54_bf16x3.json')); print('bf16x3', d['ms_per_step'], d['value'], d['e2e']['value'])
PY
tail -2 gpurun_out/r2_bench_54.err
