"""Time flow_head_bwd_kernel at the step's four levels (bf16 activations), CUDA events, rotating inputs."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import _lib as L

dev = torch.device("cuda", 0)
for (N, H, W) in [(8, 192, 640), (8, 96, 320), (8, 48, 160), (8, 24, 80)]:
    xs = [torch.randn((N, H, W, 32), device=dev).bfloat16() for _ in range(3)]
    dys = [torch.randn((N, H, W, 2), device=dev) for _ in range(3)]
    w = torch.randn((3, 3, 32, 2), device=dev)
    dw = torch.empty_like(w); db = torch.empty(2, device=dev); dz = torch.empty_like(xs[0])
    def f(i):
        L.check(L.load().of_flow_head_bwd(L.ptr(xs[i % 3]), L.ptr(dys[i % 3]), L.ptr(w), L.ptr(dw), L.ptr(db), L.ptr(dz), N, H, W, 32,
                                          L.OF_ACT_LEAKY, L.dtype_code(xs[0]), L.stream_ptr()), "of_flow_head_bwd")
    for i in range(3): f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20): f(i)
    e1.record(); torch.cuda.synchronize()
    print("%s head_bwd N%d %dx%d: %.1f us  sum(dw) %.6e" % (os.environ.get("OF_PROFILE_LIB", "default"), N, H, W, e0.elapsed_time(e1) / 20 * 1e3, float(dw.double().sum())), flush=True)
