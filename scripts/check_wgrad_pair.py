"""CTA-pair wgrad (option OF_WG_PAIR): parity against the one-CTA kernel on the same inputs, then timings."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optical_flow_b200 import conv_tc, _lib as L

dev = torch.device("cuda", 0)


def one(N, H, W, Cin, Cout, bias=True):
    torch.manual_seed(1)
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16()
    g = torch.randn((N, H, W, Cout), device=dev).bfloat16()
    dw0, db0 = conv_tc.conv2d_wgrad_tc(x, g, (3, 3), stride=1, need_bias_grad=bias)
    with L.option("OF_WG_PAIR"), L.option("OF_WG_PAIR_MIN", "1"):
        dw1, db1 = conv_tc.conv2d_wgrad_tc(x, g, (3, 3), stride=1, need_bias_grad=bias)
    torch.cuda.synchronize()
    e = float((dw1 - dw0).norm() / dw0.norm())
    eb = float((db1 - db0).norm() / db0.norm()) if bias else 0.0
    print("N%d %dx%d %d->%d bias=%d: rel dw %.3e db %.3e" % (N, H, W, Cin, Cout, bias, e, eb), flush=True)
    if e > 1e-4:
        d = (dw1 - dw0).abs()
        for kh in range(3):
            print("   ", ["%.2e" % float(d[kh, kw].max()) for kw in range(3)],
                  "cols", [int(i) for i in torch.nonzero(d[kh].amax(dim=(0, 1)) > 1e-2 * float(dw0.abs().max()))[:6, 0]])
    return e


if __name__ == "__main__":
    shapes = [(1, 16, 32, 128, 128), (2, 24, 32, 128, 96), (1, 16, 48, 120, 128), (1, 12, 16, 160, 128), (1, 16, 32, 256, 128)]
    bad = 0
    for s in shapes:
        for b in (True, False):
            bad += one(*s, bias=b) > 1e-4
    print("PARITY", "OK" if not bad else "FAIL %d" % bad, flush=True)
