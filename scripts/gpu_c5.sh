# BASELINE configs[4] shape on one GPU (batch 4 at 1024x2048) and configs[3] per-GPU batches 16 / 32
for args in "--batch 4 --height 1024 --width 2048" "--batch 16" "--batch 32"; do
  echo "== $args"; python bench.py $args --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('ms', round(d['ms_per_step'],2), 'pairs/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['config']['workload'][:90], 'step_frac', round(d['roofline']['step']['frac'],3))"
done
python bench.py --workload loss --batch 4 --height 1024 --width 2048 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-900
