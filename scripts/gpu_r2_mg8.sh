#!/bin/bash
# 8-GPU check: weak scaling (8 pairs per GPU, graph replay, full default line incl. e2e) and C4's literal reading (global batch 64)
N=${1:-8}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_mg${N}_weak.json 2> gpurun_out/r2_mg${N}_weak.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus $N --steps 20 --warmup 5 --scaling strong --global-batch 64 --no-cpu-baseline --no-extras > gpurun_out/r2_mg${N}_strong.json 2> gpurun_out/r2_mg${N}_strong.err
python - <<PY
import json
for f in ("weak","strong"):
    try:
        txt=open("gpurun_out/r2_mg${N}_%s.json"%f).read()
        d=json.loads(txt[txt.find('{"metric'):txt.rfind('}')+1]); print(f, d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), round(d["e2e"]["value"],1), d["config"]["global_batch"])
    except Exception as e: print(f,"ERR",e)
PY
tail -2 gpurun_out/r2_mg${N}_weak.err
