#!/bin/bash
# the driver's GPU gate, twice in one call (flakiness probe on one more box)
mkdir -p gpurun_out
for k in 1 2; do
  ( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest_gate_$k.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/r2_pytest_gate_$k.log
  tail -3 gpurun_out/r2_pytest_gate_$k.log
done
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader
