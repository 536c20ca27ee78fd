#!/bin/bash
# round 2: is the N-stacked wgrad bound by L2 -> SM bytes?  Same MMA work, slab halo varied through the pixel-tile shape
# (TW x TH = 8x8: 36 KB per K tile, 16x4: 40 KB (default), 32x2: 48 KB at 128->128)
mkdir -p gpurun_out
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_STATIC=1 OF_WG_STATIC=1,OF_WG_TW=8 OF_WG_STATIC=1,OF_WG_TW=32 OF_WG_TW=8 2>&1 | tee gpurun_out/r2_wgrad_layers_32.txt
