#!/bin/bash
# compute-sanitizer over the inter-CTA protocols of the fused loss (ticket / per-image counters / progress words) and
# the scatter atomics of the warp backward: ONE tool per gpurun call (B200_PROFILING.md).  usage: gpu_sanitize.sh memcheck|racecheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SEL='test_loss_layer_fused or (test_loss_fast_path_equals_generic and not 384) or test_warp_features_bwd or test_warp_bwd_smooth or test_my_loss_fused or test_loss_layer_deterministic'
# the same selection must pass without the tool first
timeout 600 python -m pytest tests/test_b1_gpu_ops.py -m gpu -q -x -k "$SEL" > gpurun_out/r2_sanitize_plain.log 2>&1 || { tail -5 gpurun_out/r2_sanitize_plain.log; exit 1; }
tail -2 gpurun_out/r2_sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 --log-file gpurun_out/r2_sanitize_$TOOL.txt \
  python -m pytest tests/test_b1_gpu_ops.py -m gpu -q -x -k "$SEL" > gpurun_out/r2_sanitize_${TOOL}_pytest.log 2>&1
echo "compute-sanitizer $TOOL rc=$?" | tee -a gpurun_out/r2_sanitize_$TOOL.txt
tail -3 gpurun_out/r2_sanitize_${TOOL}_pytest.log
tail -5 gpurun_out/r2_sanitize_$TOOL.txt
