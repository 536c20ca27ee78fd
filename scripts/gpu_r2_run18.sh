#!/bin/bash
# round 2: K-step skip (separate unrolled path, -DT3_KSKIP) A/B on the same box: conv parity tests on the variant, per-layer
# timings, whole step; loss kernel with the elect.sync producer
mkdir -p gpurun_out
( OF_PROFILE_LIB=kskip timeout 600 python -m pytest tests/test_b2_gpu_conv_tc.py tests/test_b4_bench_shapes.py -m gpu -x -q -k "tc_fprop or tc_dgrad or tc_conv_at_bench or sign_bit or post_add" ) > gpurun_out/r2_pytest_19_kskip.log 2>&1
tail -2 gpurun_out/r2_pytest_19_kskip.log
echo "== default"; timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -8
echo "== kskip"; OF_PROFILE_LIB=kskip timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -8
for v in "" kskip; do
  OF_PROFILE_LIB=$v timeout 300 python bench.py --no-cpu-baseline --steps 40 > gpurun_out/r2_bench_19_${v:-default}.json 2>> gpurun_out/r2_bench_19.err
done
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_19_default.json','gpurun_out/r2_bench_19_kskip.json'):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], 'frac', r['frac'], 'loss', r['loss_fused']['flows']['smooth']['us'])
    except Exception as e: print(f,'ERR', e)
PY
tail -3 gpurun_out/r2_bench_19.err
