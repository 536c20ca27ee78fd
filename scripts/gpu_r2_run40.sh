#!/bin/bash
# round 2: bias gradient as TRANSPOSED MMAs (A = dy tile, B = ones, N = 16) against the ones x dy MMAs (N = Cout)
mkdir -p gpurun_out
timeout 100 python - <<'PY' 2>&1 | tee gpurun_out/r2_wgrad_biast_40.txt
import sys, torch
sys.path.insert(0, '.')
from optical_flow_b200 import conv_tc, _lib as L
dev = torch.device('cuda', 0)
bad = 0
for (N, H, W, Cin, Cout, K, s) in [(2, 24, 32, 128, 128, 3, 1), (1, 16, 48, 64, 32, 3, 1), (1, 16, 16, 256, 256, 3, 1), (2, 20, 40, 32, 64, 3, 1),
                             (1, 16, 32, 96, 64, 3, 1), (1, 16, 48, 128, 96, 3, 1), (2, 24, 32, 64, 128, 3, 2), (1, 24, 32, 64, 128, 1, 2),
                             (1, 10, 24, 160, 48, 3, 1), (2, 12, 64, 64, 16, 3, 1), (1, 16, 16, 64, 192, 3, 1)]:
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16(); g = torch.randn((N, (H + s - 1) // s, (W + s - 1) // s, Cout), device=dev).bfloat16()
    dw0, db0 = conv_tc.conv2d_wgrad_tc(x, g, (K, K), stride=s)
    with L.option("OF_WG_BIAS_T"):
        dw1, db1 = conv_tc.conv2d_wgrad_tc(x, g, (K, K), stride=s)
    ref = g.float().sum(dim=(0, 1, 2))
    e = float((db1 - ref).norm() / ref.norm()); ew = float((dw1 - dw0).norm() / dw0.norm())
    bad += (e > 1e-5) or (ew > 1e-5)
    print(Cin, Cout, K, s, 'db T vs torch %.2e (old %.2e); dw %.2e' % (e, float((db0 - ref).norm() / ref.norm()), ew))
print("PARITY", "OK" if not bad else "FAIL")
PY
timeout 300 python scripts/bench_wgrad_layer.py OF_WG_BIAS_T=1 NOBIAS 2>&1 | tee gpurun_out/r2_wgrad_layers_40.txt
