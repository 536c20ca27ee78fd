"""Time one tensor-core conv layer shape (fprop / dgrad with mask) with CUDA events."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc, _lib as L
def run(N, H, W, Cin, Cout, iters=20):
    dev = torch.device("cuda", 0)
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16()
    w = torch.randn((3, 3, Cin, Cout), device=dev) * 0.05
    b = torch.randn((Cout,), device=dev)
    segs = [(0, Cin, 0)]
    kb = (Cin + 63) // 64 * 64
    n_mma = (Cout + 15) // 16 * 16
    wp = conv_tc.pack_weights(w, segs, n_mma, kb, transpose=False)
    out = torch.empty((N, H, W, Cout), dtype=torch.bfloat16, device=dev)
    def f():
        conv_tc.tc_run([conv_tc.Source(x)], wp, n_mma, Cout, N, H, W, H, W, 3, 3, 1, 1, 1, +1, out, bias=b, act=L.OF_ACT_LEAKY)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    fl = 2.0 * N * H * W * Cin * Cout * 9
    print("N%d %dx%d %d->%d: %.1f us  %.0f TFLOP/s" % (N, H, W, Cin, Cout, us, fl / us / 1e6))
if __name__ == "__main__":
    for opt in ("OF_T3_RINGS", "OF_T3_2CTA"):          # ring depths "na,nb" / two CTAs per SM policy, through of_set_option
        if os.environ.get("X_" + opt):
            L.check(L.load().of_set_option(opt.encode(), os.environ["X_" + opt].encode()), "of_set_option")
            print("option", opt, os.environ["X_" + opt])
    shapes = ((128, 128), (128, 96), (96, 64), (64, 32), (32, 64), (64, 64))
    if os.environ.get("OF_CONV_SHAPES") == "pow2":      # the swapped-operand timing hook over-reads a 96-row weight tile
        shapes = ((128, 128), (128, 64), (64, 64), (64, 32), (32, 64))
    for cin, cout in shapes:
        run(8, 192, 640, cin, cout)
    run(16, 96, 320, 64, 64)
    if os.environ.get("OF_CONV_SHAPES") != "pow2":
        run(8, 96, 320, 96, 64)
