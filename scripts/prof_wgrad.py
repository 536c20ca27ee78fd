"""ncu target: a few wgrad launches at the decoder level-3 shapes (see scripts/gpu_prof_wgrad.sh)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc
dev = torch.device("cuda", 0)
for (N, H, W, Cin, Cout) in [(8, 192, 640, 128, 128), (8, 192, 640, 96, 64), (8, 192, 640, 64, 32), (16, 96, 320, 64, 64)]:
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16()
    g = torch.randn((N, H, W, Cout), device=dev).bfloat16()
    for _ in range(2):
        conv_tc.conv2d_wgrad_tc(x, g, (3, 3), stride=1)
    torch.cuda.synchronize()
