# usage: gpu_ab2.sh "ENV=VAL" ["ENV=VAL" ...]: bench with each setting on the same box
fmt='import sys, json; d = json.loads(sys.stdin.read()); print("STEP ms %.3f pairs/s %.1f e2e %.1f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]))'
for rep in 1 2; do
  for kv in "$@"; do
    echo "$kv:"; env $kv python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "$fmt"
  done
done
