#!/bin/bash
# round 2: SIMT bias gradient in wgrad - one polling lane per warp (default now), all threads polling, one bias job per range
mkdir -p gpurun_out
( timeout 300 python -m pytest tests -m gpu -x -q -k "wgrad or stem" ) > gpurun_out/r2_pytest_37.log 2>&1
tail -2 gpurun_out/r2_pytest_37.log
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_MMA_BIAS=1 OF_WG_BIAS_RANGE=1 NOBIAS 2>&1 | tee gpurun_out/r2_wgrad_layers_37.txt
