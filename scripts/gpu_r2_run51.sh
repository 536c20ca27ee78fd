#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_b4_bench_shapes.py -m gpu -x -q -k "step_oracle" 2>&1 | tail -5
python - <<'PY'
import json, numpy as np
for m in ("fp32","bf16x3"):
    try:
        a=json.load(open('gpurun_out/parity_%s_step_384x1280.json'%m)); rl=a['rel_l2']; w=max(rl.items(), key=lambda kv: kv[1])
        print(m, 'loss %.9f ref %.9f worst %.3e (%s) median %.3e' % (a['loss'], a['ref_loss'], w[1], w[0], float(np.median(list(rl.values())))))
    except Exception as e: print(m, 'ERR', e)
PY
