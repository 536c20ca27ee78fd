#!/bin/bash
# round 2: wgrad N-stacked main loop with the K slice outermost (consecutive MMAs into different accumulators)
mkdir -p gpurun_out
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_KOUTER=1 NOBIAS NOBIAS,OF_WG_KOUTER=1 OF_WG_KOUTER=1,OF_WG_BIAS_T=1 2>&1 | grep "96->64\|128" | tee gpurun_out/r2_wgrad_layers_41.txt
OF_PROFILE_LIB=1 timeout 300 python scripts/bench_wgrad_layer.py NOBIAS,OF_WG_DBG=2 NOBIAS,OF_WG_DBG=2,OF_WG_KOUTER=1 2>&1 | grep "DBG" | grep "96->64\|128" | tee -a gpurun_out/r2_wgrad_layers_41.txt
