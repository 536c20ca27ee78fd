#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/timeline_mode.py bf16x3 2>&1 | tail -30 | tee gpurun_out/r2_timeline_bf16x3.txt
