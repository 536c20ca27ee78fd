#!/bin/bash
# round 2: conv_tc3 half-K variants (KS = 2) for single sources of <= 32 channels: tests, A/B on the step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_b2_gpu_conv_tc.py tests/test_b3_gpu_net.py -m gpu -x -q 2>&1 | tail -3
for v in default OF_T3_NO_HALFK=1; do
  opt=""; [ "$v" != default ] && opt="--option $v"
  timeout 300 python bench.py --no-cpu-baseline $opt --dump-timeline gpurun_out/r2_timeline_55_$v.json > gpurun_out/r2_bench_55_$v.json 2>> gpurun_out/r2_bench_55.err
done
python - <<'PY'
import json, glob
for v in ("default","OF_T3_NO_HALFK=1"):
    d=json.load(open('gpurun_out/r2_bench_55_%s.json'%v)); t=json.load(open('gpurun_out/r2_timeline_55_%s.json'%v)); n=len(t)//3
    heads=[round(r['us'],1) for r in t[:n] if r['kernel']=='conv_tc3_kernel<head>']
    c3=sum(r['us'] for r in t[:n] if r['kernel']=='conv_tc3_kernel')
    print(v, d['ms_per_step'], d['value'], 'frac', d['roofline']['frac'], 'heads', heads, 'conv_tc3', round(c3,1))
PY
tail -2 gpurun_out/r2_bench_55.err
