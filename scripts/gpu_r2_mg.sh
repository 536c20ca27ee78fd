#!/bin/bash
# multi-GPU check: N = $1 GPUs (weak scaling, graph on / off) and the strong-scaling point (global batch 64)
N=${1:-2}
mkdir -p gpurun_out
for g in on off; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 --graph $g --no-cpu-baseline --no-extras > gpurun_out/r2_mg${N}_graph_$g.json 2> gpurun_out/r2_mg${N}_graph_$g.err
  tail -c 600 gpurun_out/r2_mg${N}_graph_$g.json; tail -2 gpurun_out/r2_mg${N}_graph_$g.err
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus $N --steps 10 --warmup 3 --scaling strong --global-batch 64 --no-cpu-baseline --no-extras > gpurun_out/r2_mg${N}_strong.json 2> gpurun_out/r2_mg${N}_strong.err
tail -c 600 gpurun_out/r2_mg${N}_strong.json; tail -2 gpurun_out/r2_mg${N}_strong.err
python - <<PY
import json
for f in ("graph_on","graph_off","strong"):
    try:
        d=json.load(open("gpurun_out/r2_mg${N}_%s.json"%f)); print(f, d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), round(d["e2e"]["value"],1))
    except Exception as e: print(f,"ERR",e)
PY
