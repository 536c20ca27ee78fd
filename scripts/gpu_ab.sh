# A/B of one env toggle on the same box: bash scripts/gpu_ab.sh ENVVAR
fmt='import sys, json; d = json.loads(sys.stdin.read()); print("STEP ms %.3f pairs/s %.1f e2e %.1f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]))'
for rep in 1 2; do
  echo "default:"; python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "$fmt"
  echo "$1=1:"; env $1=1 python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "$fmt"
done
