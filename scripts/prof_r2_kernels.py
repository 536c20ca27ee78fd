"""One launch each of the kernels profiled at the end of round 2 (run under ncu): conv_tc3 fprop 128->128 and 64->64
(N16 96x320), wgrad 128->128, cost-volume backward (level 3)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc, functional as F, _lib as L
dev = torch.device("cuda", 0)


def conv_case(N, H, W, Cin, Cout):
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16()
    w = torch.randn((3, 3, Cin, Cout), device=dev) * 0.05
    b = torch.randn((Cout,), device=dev)
    dz = torch.randn((N, H, W, Cout), device=dev).bfloat16()
    bits = torch.zeros((N, H, W, 16), dtype=torch.uint8, device=dev)
    for _ in range(2):
        conv_tc.conv2d_fprop_tc(x, w, b, act=L.OF_ACT_LEAKY, sign_out=bits)     # as the train step launches it
        conv_tc.conv2d_dgrad_tc(dz, w, (H, W), 1, mask_bits=bits, mask_act=L.OF_ACT_LEAKY)
        conv_tc.conv2d_wgrad_tc(x, dz, (3, 3))
    torch.cuda.synchronize()


conv_case(8, 192, 640, 128, 128)
conv_case(16, 96, 320, 64, 64)
B, h, w_, C = 8, 192, 640, 64
f1 = torch.randn((B, h, w_, C), device=dev).bfloat16()
f2 = torch.randn((B, h, w_, C), device=dev).bfloat16()
dcv = torch.randn((B, h, w_, 64), device=dev).bfloat16()
d1, d2 = torch.empty_like(f1), torch.empty_like(f1)
for _ in range(2):
    L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(f1), L.ptr(f2), L.ptr(dcv), L.ptr(d1), L.ptr(d2), None, B, h, w_, C, 3, 64,
                                             L.stream_ptr()), "of_cost_volume_bwd_bf16")
torch.cuda.synchronize()
print("done")
