"""Time the bf16 tensor-core cost-volume backward (both launches) at the decoder shapes."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import _lib as L
dev = torch.device("cuda", 0)
for (B, H, W, C) in [(8, 192, 640, 64), (8, 96, 320, 64), (8, 48, 160, 128)]:
    f1 = [torch.randn((B, H, W, C), device=dev).bfloat16() for _ in range(2)]
    f2 = [torch.randn((B, H, W, C), device=dev).bfloat16() for _ in range(2)]
    g = [torch.randn((B, H, W, 64), device=dev).bfloat16() for _ in range(2)]
    d1, d2 = torch.empty_like(f1[0]), torch.empty_like(f1[0])
    def f(i):
        L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(f1[i % 2]), L.ptr(f2[i % 2]), L.ptr(g[i % 2]), L.ptr(d1), L.ptr(d2), None,
                                                 B, H, W, C, 3, 64, L.stream_ptr()), "bwd")
    for i in range(3): f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10): f(i)
    e1.record(); torch.cuda.synchronize()
    print("cost volume bwd (d_f1 + d_f2) B%d %dx%d C%d: %.1f us" % (B, H, W, C, e0.elapsed_time(e1) / 10 * 1e3), flush=True)
