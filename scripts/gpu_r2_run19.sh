#!/bin/bash
# round 2: conv_tc3 ring-depth sweep (bytes in flight vs TMA latency on the narrow layers)
mkdir -p gpurun_out
for r in "" "3,6" "4,7" "5,3" "2,12" "4,4"; do
  echo "== rings ${r:-default} (2CTA policy default)"; X_OF_T3_RINGS=$r timeout 120 python scripts/bench_conv_layer.py 2>&1 | grep -v "^option" | tail -8
done
echo "== rings 4,7 with 2CTA off"; X_OF_T3_RINGS=4,7 X_OF_T3_2CTA=0 timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -8
echo "== default rings with 2CTA off"; X_OF_T3_2CTA=0 timeout 120 python scripts/bench_conv_layer.py 2>&1 | tail -8
