#!/bin/bash
# round 2: what do the fused bias-gradient MMAs cost in wgrad, one-CTA and CTA-pair kernels
mkdir -p gpurun_out
timeout 300 python scripts/bench_wgrad_layer.py NOBIAS OF_WG_PAIR=1 OF_WG_PAIR=1,NOBIAS 2>&1 | tee gpurun_out/r2_wgrad_layers_34.txt
