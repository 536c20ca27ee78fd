#!/bin/bash
# round 2: wgrad main loop dissection (profile-hooks build): MMAs only (no TMA after the first ring fill) / loads only (no MMAs)
mkdir -p gpurun_out
OF_PROFILE_LIB=1 timeout 300 python scripts/bench_wgrad_layer.py NOBIAS NOBIAS,OF_WG_DBG=2 NOBIAS,OF_WG_DBG=4 NOBIAS,OF_WG_DBG=6 2>&1 | tee gpurun_out/r2_wgrad_dissect_39.txt
