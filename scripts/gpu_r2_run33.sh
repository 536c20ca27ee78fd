#!/bin/bash
# round 2: CTA-pair (cta_group::2) wgrad - parity against the one-CTA kernel, then per-layer timings and the step
mkdir -p gpurun_out
timeout 120 python scripts/check_wgrad_pair.py > gpurun_out/r2_wgpair_33.txt 2>&1; echo "rc=$?" >> gpurun_out/r2_wgpair_33.txt
tail -30 gpurun_out/r2_wgpair_33.txt
if grep -q "PARITY OK" gpurun_out/r2_wgpair_33.txt; then
  timeout 200 python scripts/bench_wgrad_layer.py OF_WG_PAIR=1 2>&1 | grep "128\|96->64" | tee gpurun_out/r2_wgrad_layers_33.txt
  timeout 300 python bench.py --no-cpu-baseline --option OF_WG_PAIR=1 > gpurun_out/r2_bench_33_pair.json 2>> gpurun_out/r2_bench_33.err
  timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_33_default.json 2>> gpurun_out/r2_bench_33.err
  python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_bench_33_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, d['ms_per_step'], d['value'], 'frac', r['frac'])
        for k,v in list(d.get('kernels',{}).items())[:12]:
            if 'wgrad' in k: print("  %-40s %8.1f us n=%5.1f %.3f"%(k[:40],v['us'],v['launches'],v['share']))
    except Exception as e: print(f,'ERR', e)
PY
  tail -3 gpurun_out/r2_bench_33.err
fi
