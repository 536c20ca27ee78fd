"""Time the tensor-core wgrad kernel at the step's layer shapes (CUDA events, rotating inputs), with mode toggles."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc


def run(N, H, W, Cin, Cout, iters=20, tag="", bias=True):
    dev = torch.device("cuda", 0)
    xs = [torch.randn((N, H, W, Cin), device=dev).bfloat16() for _ in range(3)]
    gs = [torch.randn((N, H, W, Cout), device=dev).bfloat16() for _ in range(3)]
    def f(i):
        conv_tc.conv2d_wgrad_tc(xs[i % 3], gs[i % 3], (3, 3), stride=1, need_bias_grad=bias)
    for i in range(3): f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters): f(i)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    fl = 2.0 * N * H * W * Cin * Cout * 9
    print("%-10s N%d %dx%d %d->%d: %.1f us  %.0f TFLOP/s" % (tag, N, H, W, Cin, Cout, us, fl / us / 1e6), flush=True)


if __name__ == "__main__":
    shapes = [(8, 192, 640, 64, 32), (16, 96, 320, 64, 64), (8, 96, 320, 64, 32), (8, 192, 640, 96, 64),
              (8, 192, 640, 128, 96), (8, 192, 640, 128, 128)]
    # variants from the command line ("OF_WG_STATIC=1" ...), default first
    variants = [""] + sys.argv[1:]
    from optical_flow_b200 import _lib as L
    for env in variants:
        opts = [o.partition("=") for o in env.split(",") if o and o != "NOBIAS"]
        for k, _, v in opts: L.check(L.load().of_set_option(k.encode(), (v or "1").encode()), "of_set_option")
        for s in shapes:
            run(*s, tag=env or "default", bias="NOBIAS" not in env)
        for k, _, v in opts: L.load().of_set_option(k.encode(), None)
