#!/bin/bash
# round 2 closing evidence: ncu launch list of the final tree's step (the bench command itself ran in the gate call before)
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_train_launches_final.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --graph off > gpurun_out/r2_ncu_final.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/r2_train_launches_final.csv
