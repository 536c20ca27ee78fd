# quick iteration: a subset of the GPU tests, then the default bench (value / e2e only)
set -x
python -m pytest tests/test_b2_gpu_conv_tc.py -x -q -m gpu 2>&1 | tail -8
python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('STEP ms', d['ms_per_step'], 'pairs/s', d['value'], 'e2e', d['e2e']['value'], 'clk', d['clocks'])"
