#!/bin/bash
# round 2: 16-byte bf16 paths of maxpool forward / add+relu / act backward / zero-insertion scatter: tests + step A/B is by timeline
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_b1_gpu_ops.py tests/test_b3_gpu_net.py -m gpu -x -q -k "streaming or maxpool or relu or act or pool" 2>&1 | tail -3
timeout 300 python bench.py --no-cpu-baseline --dump-timeline gpurun_out/r2_timeline_52.json > gpurun_out/r2_bench_52.json 2> gpurun_out/r2_bench_52.err
python - <<'PY'
import json, collections
d=json.load(open('gpurun_out/r2_bench_52.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'])
t=json.load(open('gpurun_out/r2_timeline_52.json')); n=len(t)//3
agg=collections.defaultdict(lambda:[0,0.0])
for r in t[:n]: agg[r['kernel']][0]+=1; agg[r['kernel']][1]+=r['us']
for k in ('maxpool2_fwd','add_relu','act_bwd','scatter_upsample2_kernel','maxpool2_bwd_add_act','colsum4_kernel','conv_tc_wgrad_kernel'):
    print(k, agg.get(k))
print('total', sum(v[1] for v in agg.values()), 'launches', n)
PY
tail -2 gpurun_out/r2_bench_52.err
