"""Time the flow-head backward kernel at the level-3 / level-2 shapes."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import _lib as L
dev = torch.device("cuda", 0)
for (N, H, W) in [(8, 192, 640), (8, 96, 320)]:
    xs = [torch.randn((N, H, W, 32), device=dev).bfloat16() for _ in range(4)]
    dy = torch.randn((N, H, W, 2), device=dev)
    w = torch.randn((3, 3, 32, 2), device=dev)
    dw, db, dz = torch.empty_like(w), torch.empty(2, device=dev), torch.empty_like(xs[0])
    def f(i):
        L.check(L.load().of_flow_head_bwd(L.ptr(xs[i % 4]), L.ptr(dy), L.ptr(w), L.ptr(dw), L.ptr(db), L.ptr(dz), N, H, W, 32,
                                          L.OF_ACT_LEAKY, L.OF_BF16, L.stream_ptr()), "head")
    for i in range(3): f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20): f(i)
    e1.record(); torch.cuda.synchronize()
    print("flow_head_bwd N%d %dx%d: %.1f us" % (N, H, W, e0.elapsed_time(e1) / 20 * 1e3), flush=True)
