#!/bin/bash
# round 2: ncu --set full of the SPECIALISED conv_tc3 launches (fprop + sign bits, dgrad * bit mask) and of wgrad; launch list summary
mkdir -p gpurun_out
timeout 120 python scripts/prof_r2_kernels.py > gpurun_out/r2_prof_plain_26.log 2>&1; tail -1 gpurun_out/r2_prof_plain_26.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_tc3_kernel" -c 8 -f -o gpurun_out/r2_conv_tc3_spec \
  python scripts/prof_r2_kernels.py > gpurun_out/r2_ncu_top_26.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/r2_conv_tc3_spec.ncu-rep
timeout 200 python scripts/bench_conv_layer.py 2>&1 | tail -8
