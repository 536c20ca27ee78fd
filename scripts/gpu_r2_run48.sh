#!/bin/bash
# round 2: flow_head_bwd_kernel occupancy / prefetch-depth variants
mkdir -p gpurun_out
for v in "" m3 m4 p8 m3p8 p2m4; do OF_PROFILE_LIB=$v timeout 100 python scripts/bench_head_bwd.py 2>&1 | tail -4; done | tee gpurun_out/r2_head_bwd_48.txt
