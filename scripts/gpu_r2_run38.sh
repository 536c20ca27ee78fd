#!/bin/bash
# round 2: bias-gradient ones-MMAs with M = 64 instead of 128 (only row 0 is read)
mkdir -p gpurun_out
( timeout 300 python -m pytest tests -m gpu -x -q -k "wgrad or stem" ) > gpurun_out/r2_pytest_38.log 2>&1
tail -2 gpurun_out/r2_pytest_38.log
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_BIAS_M64=1 NOBIAS 2>&1 | tee gpurun_out/r2_wgrad_layers_38.txt
timeout 100 python - <<'PY'
import sys, torch
sys.path.insert(0, '.')
from optical_flow_b200 import conv_tc, _lib as L
dev = torch.device('cuda', 0)
for (N, H, W, Cin, Cout) in [(2, 24, 32, 128, 128), (1, 16, 48, 64, 32), (1, 16, 16, 256, 256), (2, 20, 40, 32, 64), (1, 16, 32, 96, 64)]:
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16(); g = torch.randn((N, H, W, Cout), device=dev).bfloat16()
    dw0, db0 = conv_tc.conv2d_wgrad_tc(x, g, (3, 3), stride=1)
    with L.option("OF_WG_BIAS_M64"):
        dw1, db1 = conv_tc.conv2d_wgrad_tc(x, g, (3, 3), stride=1)
    ref = g.float().sum(dim=(0, 1, 2))
    print(Cin, Cout, 'db rel m64 vs m128 %.2e, vs torch %.2e; dw %.2e' % (float((db1 - db0).norm() / db0.norm()), float((db1 - ref).norm() / ref.norm()), float((dw1 - dw0).norm() / dw0.norm())))
PY
