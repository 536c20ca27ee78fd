"""One level-3 decoder layer (128->128, N8 192x640): fprop, then its dgrad with the LeakyReLU mask epilogue."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc, _lib as L
dev = torch.device("cuda", 0)
N, H, W, C = 8, 192, 640, 128
x = torch.randn((N, H, W, C), device=dev).bfloat16()
w = torch.randn((3, 3, C, C), device=dev) * 0.05
b = torch.randn((C,), device=dev)
dz = torch.randn((N, H, W, C), device=dev).bfloat16()
yprev = torch.randn((N, H, W, C), device=dev).bfloat16()
for _ in range(3):
    y = conv_tc.conv2d_fprop_tc(x, w, b, act=L.OF_ACT_LEAKY)
    dx = conv_tc.conv2d_dgrad_tc(dz, w, (H, W), 1, mask=yprev, mask_act=L.OF_ACT_LEAKY)
    dx2 = conv_tc.conv2d_dgrad_tc(dz, w, (H, W), 1)
torch.cuda.synchronize()
def t(f, n=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
print("fprop %.1f us" % t(lambda: conv_tc.conv2d_fprop_tc(x, w, b, act=L.OF_ACT_LEAKY)))
print("dgrad+mask %.1f us" % t(lambda: conv_tc.conv2d_dgrad_tc(dz, w, (H, W), 1, mask=yprev, mask_act=L.OF_ACT_LEAKY)))
print("dgrad nomask %.1f us" % t(lambda: conv_tc.conv2d_dgrad_tc(dz, w, (H, W), 1)))
