"""GPU parity tests (run with -m gpu on the B200 box): every hand-written CUDA op
called THROUGH THE C ABI against the CPU oracle on the same seeded inputs.

Bars (north_star): warp sample indices bit-exact; warped images, losses and
gradients within 1e-5 relative in fp32.  For scatter-add gradients (float
atomics reorder the sum) the 1e-5 is relative to the tensor's max-abs.
"""
import numpy as np
import pytest
import torch

from oracle import ref_np as R, ref_np_ops as O, ref_torch as T

pytestmark = pytest.mark.gpu

REL = 1e-5


def rng(seed):
    return np.random.default_rng(seed)


def dev(a, cuda):
    return torch.tensor(np.ascontiguousarray(a), device=cuda)


def assert_close_maxrel(got, exp, rel=REL, what=""):
    got = got.detach().float().cpu().numpy() if isinstance(got, torch.Tensor) else got
    scale = max(float(np.abs(exp).max()), 1e-30)
    err = float(np.abs(got - exp).max())
    assert err <= rel * scale, "%s: max|err| %.3e > %.1e * max|ref| %.3e" % (what, err, rel, scale)


def adversarial_flow(r, B, H, W, sigma):
    """N(0, sigma^2) flows plus values exactly at / one ulp around integers and far out of range."""
    fl = (r.standard_normal((B, H, W, 2)) * sigma).astype(np.float32)
    flat = fl.reshape(-1)
    n = flat.size
    k = max(n // 16, 1)
    pos = r.choice(n, size=4 * k, replace=False)
    ints = r.integers(-6, 7, size=k).astype(np.float32)
    flat[pos[:k]] = ints
    flat[pos[k:2 * k]] = np.nextafter(ints, np.float32(np.inf))
    flat[pos[2 * k:3 * k]] = np.nextafter(ints, np.float32(-np.inf))
    flat[pos[3 * k:]] = r.choice(np.array([-1000.5, 1000.5, -37.25, 41.75], np.float32), size=k)
    return fl


# ------------------------------------------------------------------ sampler --
@pytest.mark.parametrize("shape", [(2, 24, 40, 3), (2, 17, 29, 64), (1, 8, 26, 128), (1, 5, 7, 6), (3, 9, 4, 16)])
def test_bilinear_interpolation_fwd_indices_bit_exact(cuda, shape):
    from optical_flow_b200.transformations import bilinear_interpolation
    B, H, W, C = shape
    r = rng(1)
    inp = r.standard_normal(shape).astype(np.float32)
    base = np.stack(np.meshgrid(np.arange(W), np.arange(H), indexing="xy"), axis=-1).astype(np.float32)[None]
    pts = (base + adversarial_flow(r, B, H, W, 4.0)).astype(np.float32)
    exp, (x0, x1, y0, y1) = R.bilinear_interpolation(inp, pts, return_indices=True)
    out, idx = bilinear_interpolation(dev(inp, cuda), dev(pts, cuda), return_indices=True)
    idx = idx.cpu().numpy()
    assert np.array_equal(idx[..., 0], x0) and np.array_equal(idx[..., 1], x1)
    assert np.array_equal(idx[..., 2], y0) and np.array_equal(idx[..., 3], y1)
    # no FMA contraction + reference accumulate order -> fp32 values are bit-exact too
    assert np.array_equal(out.cpu().numpy(), exp)


@pytest.mark.parametrize("mode", ["ref_ij", "xy"])
@pytest.mark.parametrize("shape", [(2, 24, 80, 64), (1, 16, 52, 128), (2, 13, 21, 3), (1, 6, 9, 8)])
def test_warp_features_fwd_bit_exact(cuda, mode, shape):
    from optical_flow_b200.model import warp_features
    B, H, W, C = shape
    r = rng(2)
    feats = r.standard_normal(shape).astype(np.float32)
    flow = adversarial_flow(r, B, H, W, 4.0)
    fn = R.warp_features if mode == "ref_ij" else R.warp_xy
    exp, (x0, x1, y0, y1) = fn(flow, feats, return_indices=True)
    out, idx = warp_features(dev(flow, cuda), dev(feats, cuda), mode=mode, return_indices=True)
    idx = idx.cpu().numpy()
    assert np.array_equal(idx[..., 0], x0) and np.array_equal(idx[..., 1], x1)
    assert np.array_equal(idx[..., 2], y0) and np.array_equal(idx[..., 3], y1)
    assert np.array_equal(out.cpu().numpy(), exp)


def test_warp_zero_flow_known_answers(cuda):
    from optical_flow_b200.model import warp_features
    B, H, W, C = 2, 6, 9, 8
    f = rng(3).standard_normal((B, H, W, C)).astype(np.float32)
    z = torch.zeros((B, H, W, 2), device=cuda)
    out = warp_features(z, dev(f, cuda)).cpu().numpy()
    exp = np.stack([[[f[b, min(j, H - 1), i] for j in range(W)] for i in range(H)] for b in range(B)])
    assert np.array_equal(out[:, :, :H], exp[:, :, :H])      # transposed gather
    np.testing.assert_allclose(out, exp, atol=1e-6)
    assert np.array_equal(warp_features(z, dev(f, cuda), mode="xy").cpu().numpy(), f)   # identity


@pytest.mark.parametrize("mode", ["ref_ij", "xy"])
@pytest.mark.parametrize("shape", [(2, 24, 80, 64), (1, 16, 52, 128), (2, 13, 21, 3), (1, 10, 12, 256), (1, 6, 9, 20)])
def test_warp_features_bwd(cuda, mode, shape):
    from optical_flow_b200 import functional as F
    B, H, W, C = shape
    r = rng(4)
    feats = r.standard_normal(shape).astype(np.float32)
    flow = (r.standard_normal((B, H, W, 2)) * 3.0).astype(np.float32)
    g = r.standard_normal(shape).astype(np.float32)
    bw = R.warp_features_bwd if mode == "ref_ij" else R.warp_xy_bwd
    e_feats, e_flow, e_abs = bw(flow, feats, g, return_abs=True)
    d_feats, d_flow = F.warp_features_bwd(dev(flow, cuda), dev(feats, cuda), dev(g, cuda), mode=mode)
    # fp32 scatter-add: 1e-5 of max|ref|, or (where clipped taps with |w| >> 1 cancel: ref_ij with
    # W > H folds every column j >= H onto row H-1) 4 fp32 ulps of the summed |contributions| - the
    # rounding bound of ANY fp32 summation order, the reference's own scatter_nd included.
    err = np.abs(d_feats.cpu().numpy() - e_feats)
    bound = np.maximum(REL * np.abs(e_feats).max(), 4 * np.finfo(np.float32).eps * e_abs)
    assert (err <= bound).all(), "d_feats: max excess %.3e" % float((err - bound).max())
    assert_close_maxrel(d_flow, e_flow, what="d_flow")


def test_warp_bwd_smooth_flow_collisions(cuda):
    """Smooth flow -> neighbouring pixels scatter to the same taps (exercises the
    warp-shuffle-combined atomics of the small-C path)."""
    from optical_flow_b200 import functional as F
    B, H, W, C = 2, 32, 48, 3
    r = rng(5)
    im = r.standard_normal((B, H, W, C)).astype(np.float32)
    base = np.full((B, H, W, 2), 0.3, np.float32) + (r.standard_normal((B, 1, 1, 2)) * 0.1).astype(np.float32)
    g = r.standard_normal((B, H, W, C)).astype(np.float32)
    ii, jj = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    for mode, bw in (("ref_ij", R.warp_features_bwd), ("xy", R.warp_xy_bwd)):
        flow = base.copy()
        # a 16x16 block whose pixels all sample the same in-range point: 256-way collisions
        # (in range, so the reference's clipped-tap weights stay in [0,1]: no fp32 cancellation)
        c0 = (ii if mode == "ref_ij" else jj)[8:24, 8:24]
        c1 = (jj if mode == "ref_ij" else ii)[8:24, 8:24]
        flow[:, 8:24, 8:24, 0] = 5.25 - c0
        flow[:, 8:24, 8:24, 1] = 7.25 - c1
        e_im, e_flow = bw(flow, im, g)
        d_im, d_flow = F.warp_features_bwd(dev(flow, cuda), dev(im, cuda), dev(g, cuda), mode=mode)
        assert_close_maxrel(d_im, e_im, what="d_im " + mode)
        assert_close_maxrel(d_flow, e_flow, what="d_flow " + mode)


def test_bilinear_interpolation_bwd(cuda):
    from optical_flow_b200.transformations import bilinear_interpolation_backward
    B, H, W, C = 2, 11, 14, 12
    r = rng(6)
    inp = r.standard_normal((B, H, W, C)).astype(np.float32)
    pts = (r.random((B, H, W, 2)) * np.array([W + 4, H + 4]) - 2).astype(np.float32)
    g = r.standard_normal((B, H, W, C)).astype(np.float32)
    e_in, e_pts = R.bilinear_interpolation_bwd(inp, pts, g)
    d_in, d_pts = bilinear_interpolation_backward(dev(inp, cuda), dev(pts, cuda), dev(g, cuda))
    assert_close_maxrel(d_in, e_in, what="d_in")
    assert_close_maxrel(d_pts, e_pts, what="d_pts")


@pytest.mark.parametrize("shape", [(2, 24, 40, 3), (1, 14, 45, 32), (1, 7, 5, 5)])
def test_image_warp_fwd_bwd(cuda, shape):
    from optical_flow_b200.old.image_warp import image_warp, image_warp_tf, image_warp_backward
    B, H, W, C = shape
    r = rng(7)
    im = r.standard_normal(shape).astype(np.float32)
    flow = adversarial_flow(r, B, H, W, 3.0)
    exp, (x0, x1, y0, y1) = R.image_warp(im, flow, return_indices=True)
    out, idx = image_warp(dev(im, cuda), dev(flow, cuda), return_indices=True)
    idx = idx.cpu().numpy()
    assert np.array_equal(idx[..., 0], x0) and np.array_equal(idx[..., 1], x1)
    assert np.array_equal(idx[..., 2], y0) and np.array_equal(idx[..., 3], y1)
    assert np.array_equal(out.cpu().numpy(), exp)
    assert np.array_equal(image_warp_tf(dev(im, cuda), dev(flow, cuda)).cpu().numpy(), exp)
    g = r.standard_normal(shape).astype(np.float32)
    flow2 = (r.standard_normal((B, H, W, 2)) * 3.0).astype(np.float32)
    e_im, e_flow = R.image_warp_bwd(im, flow2, g)
    d_im, d_flow = image_warp_backward(dev(im, cuda), dev(flow2, cuda), dev(g, cuda))
    assert_close_maxrel(d_im, e_im, what="d_im")
    assert_close_maxrel(d_flow, e_flow, what="d_flow")


def test_bf16_feature_warp(cuda):
    """bf16 storage (conv fast path): same indices; values within bf16 rounding (2^-8 rel)."""
    from optical_flow_b200 import functional as F
    B, H, W, C = 1, 12, 20, 64
    r = rng(8)
    feats = torch.tensor(r.standard_normal((B, H, W, C)).astype(np.float32)).bfloat16()
    flow = (r.standard_normal((B, H, W, 2)) * 2).astype(np.float32)
    exp = R.warp_features(flow, feats.float().numpy())
    out = F.warp_features_fwd(dev(flow, cuda), feats.to(cuda))
    np.testing.assert_allclose(out.float().cpu().numpy(), exp, rtol=2 ** -7, atol=2 ** -7)
    g = r.standard_normal((B, H, W, C)).astype(np.float32)
    e_f, e_fl = R.warp_features_bwd(flow, feats.float().numpy(), g)
    d_f, d_fl = F.warp_features_bwd(dev(flow, cuda), feats.to(cuda), dev(g, cuda))
    assert_close_maxrel(d_f, e_f, what="d_feats(bf16 storage)")
    assert_close_maxrel(d_fl, e_fl, what="d_flow(bf16 storage)")


# ------------------------------------------------------------------- resize --
@pytest.mark.parametrize("hw,ohw", [((32, 48), (16, 24)), ((32, 48), (2, 3)), ((16, 24), (5, 7)), ((8, 26), (16, 52))])
def test_resize_bilinear(cuda, hw, ohw):
    from optical_flow_b200 import functional as F
    x = rng(9).standard_normal((2, hw[0], hw[1], 6)).astype(np.float32)
    out = F.resize_bilinear(dev(x, cuda), ohw[0], ohw[1])
    np.testing.assert_allclose(out.cpu().numpy(), O.resize_bilinear(x, ohw[0], ohw[1]), rtol=0, atol=1e-6)


@pytest.mark.parametrize("hw", [(24, 80), (3, 5), (1, 1), (8, 26)])
def test_upscale_flow_fwd_bwd(cuda, hw):
    from optical_flow_b200.model import upscale_flow
    from optical_flow_b200 import functional as F
    r = rng(10)
    flow = r.standard_normal((2, hw[0], hw[1], 2)).astype(np.float32)
    up = upscale_flow(dev(flow, cuda))
    np.testing.assert_allclose(up.cpu().numpy(), O.upscale_flow(flow), rtol=0, atol=2e-6)
    g = r.standard_normal((2, 2 * hw[0], 2 * hw[1], 2)).astype(np.float32)
    assert_close_maxrel(F.upscale_flow_bwd(dev(g, cuda)), O.upscale_flow_bwd(g), what="upscale bwd")


# -------------------------------------------------------------- cost volume --
@pytest.mark.parametrize("shape", [(2, 24, 80, 64), (1, 8, 26, 256), (1, 5, 6, 4), (2, 33, 17, 128)])
def test_cost_volume_fwd_bwd(cuda, shape):
    from optical_flow_b200.model import create_cost_volume
    from optical_flow_b200 import functional as F
    r = rng(11)
    f1 = r.standard_normal(shape).astype(np.float32)
    f2 = r.standard_normal(shape).astype(np.float32)
    cv = create_cost_volume(dev(f1, cuda), dev(f2, cuda), 3)
    exp = O.create_cost_volume(f1, f2, 3)
    assert tuple(cv.shape) == exp.shape
    assert_close_maxrel(cv, exp, what="cost volume")
    g = r.standard_normal(exp.shape).astype(np.float32)
    e1, e2 = O.create_cost_volume_bwd(f1, f2, 3, g)
    d1, d2 = F.cost_volume_bwd(dev(f1, cuda), dev(f2, cuda), dev(g, cuda), 3)
    assert_close_maxrel(d1, e1, what="d_f1")
    assert_close_maxrel(d2, e2, what="d_f2")


@pytest.mark.parametrize("tc", [True, False])
@pytest.mark.parametrize("shape", [(2, 24, 80, 64), (1, 8, 26, 256), (2, 33, 17, 128), (1, 6, 200, 64), (3, 5, 300, 128)])
def test_cost_volume_fwd_bf16(cuda, shape, tc, variant):
    """bf16 cost volume written into the 64-channel decoder input buffer: the tcgen05 banded-GEMM kernel
    (costvol_tc.cu) and the SIMT kernel (costvol7.cu, option OF_CV_NO_TC) against the oracle on bf16-rounded inputs."""
    from optical_flow_b200 import _lib as L
    if not tc:
        variant("OF_CV_NO_TC", "1")
    B, H, W, C = shape
    r = rng(14)
    bf = lambda a: torch.tensor(a).bfloat16()
    f1, f2 = bf(r.standard_normal(shape).astype(np.float32)), bf(r.standard_normal(shape).astype(np.float32))
    exp = O.create_cost_volume(f1.float().numpy(), f2.float().numpy(), 3)
    cv = torch.full((B, H, W, 64), 7.0, dtype=torch.bfloat16, device=cuda)
    f1d, f2d = f1.to(cuda), f2.to(cuda)
    L.check(L.load().of_cost_volume_fwd(L.ptr(f1d), L.ptr(f2d), L.ptr(cv), B, H, W, C, 3, 64, L.OF_BF16, L.stream_ptr()),
            "of_cost_volume_fwd")
    got = cv.float().cpu().numpy()
    assert (got[..., 49:] == 7.0).all()                      # channels 49..63 belong to the caller (flow, padding)
    err = np.abs(got[..., :49] - exp)
    assert (err <= 2.0 ** -8 * np.abs(exp) + 1e-5 * np.abs(exp).max()).all(), err.max()


@pytest.mark.parametrize("tc", [True, False])
@pytest.mark.parametrize("shape", [(2, 24, 80, 64), (1, 8, 26, 256), (2, 33, 17, 128), (1, 6, 200, 64), (3, 5, 300, 128)])
def test_cost_volume_bwd_bf16_fused(cuda, shape, tc, variant):
    """of_cost_volume_bwd_bf16 (bf16 tensors, d_f1 += the conv-path gradient) vs the oracle on the same
    bf16-rounded inputs; the only extra error is the final bf16 rounding (2^-8 relative to the value)."""
    from optical_flow_b200 import _lib as L
    if not tc:
        variant("OF_CV_NO_TC", "1")
    B, H, W, C = shape
    r = rng(12)
    bf = lambda a: torch.tensor(a).bfloat16()
    f1, f2 = bf(r.standard_normal(shape).astype(np.float32)), bf(r.standard_normal(shape).astype(np.float32))
    g = np.zeros((B, H, W, 64), np.float32)
    g[..., :49] = r.standard_normal((B, H, W, 49))
    g = bf(g)
    add = bf(r.standard_normal(shape).astype(np.float32) * 3)
    e1, e2 = O.create_cost_volume_bwd(f1.float().numpy(), f2.float().numpy(), 3, g.float().numpy()[..., :49])
    e1 = e1 + add.float().numpy()
    d1 = torch.empty(shape, dtype=torch.bfloat16, device=cuda)
    d2 = torch.empty(shape, dtype=torch.bfloat16, device=cuda)
    f1d, f2d, gd, addd = f1.to(cuda), f2.to(cuda), g.to(cuda), add.to(cuda)
    L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(f1d), L.ptr(f2d), L.ptr(gd), L.ptr(d1), L.ptr(d2), L.ptr(addd),
                                             B, H, W, C, 3, 64, L.stream_ptr()), "of_cost_volume_bwd_bf16")
    for got, exp in ((d1, e1), (d2, e2)):
        err = np.abs(got.float().cpu().numpy() - exp)
        assert (err <= 2.0 ** -8 * np.abs(exp) + 1e-5 * np.abs(exp).max()).all(), err.max()
    # without the add operand
    L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(f1d), L.ptr(f2d), L.ptr(gd), L.ptr(d1), L.ptr(d2), None,
                                             B, H, W, C, 3, 64, L.stream_ptr()), "of_cost_volume_bwd_bf16")
    exp = e1 - add.float().numpy()
    err = np.abs(d1.float().cpu().numpy() - exp)
    assert (err <= 2.0 ** -8 * np.abs(exp) + 1e-5 * np.abs(e1).max()).all(), err.max()


def test_warp_features_bwd_bf16_incoming_gradient(cuda):
    """of_warp_features_bwd_g with a bf16 d_out == the fp32-d_out call on the same (bf16-representable) values."""
    from optical_flow_b200 import functional as F
    r = rng(13)
    shape = (2, 12, 20, 64)
    feats = torch.tensor(r.standard_normal(shape).astype(np.float32)).bfloat16().to(cuda)
    flow = torch.tensor((r.standard_normal((2, 12, 20, 2)) * 3.0).astype(np.float32)).to(cuda)
    g16 = torch.tensor(r.standard_normal(shape).astype(np.float32)).bfloat16().to(cuda)
    for mode in ("ref_ij", "xy"):
        a1, b1 = F.warp_features_bwd(flow, feats, g16, mode=mode)
        a2, b2 = F.warp_features_bwd(flow, feats, g16.float(), mode=mode)
        # scatter order differs run to run (atomics): compare with an fp32 summation bound
        assert (a1 - a2).abs().max().item() <= 1e-5 * a2.abs().max().item()
        # the bf16-gradient call takes the 16-byte path (8 channels per thread): same products, another fp32
        # summation order over the channels
        assert (b1 - b2).abs().max().item() <= 1e-5 * b2.abs().max().item()


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("shape", [(2, 12, 20, 32), (1, 7, 150, 32), (1, 5, 9, 16)])
def test_flow_head_backward_fused(cuda, shape, dtype):
    """of_flow_head_bwd (model.py:114 head: dw, db, chained dz in one pass) vs torch-CPU autograd of
    leaky_relu -> conv3x3(Cout = 2)."""
    from optical_flow_b200 import _lib as L
    N, H, W, Cin = shape
    r = rng(15)
    y_prev = r.standard_normal(shape).astype(np.float32)             # previous conv's post-activation output
    if dtype == "bf16":
        y_prev = torch.tensor(y_prev).bfloat16().float().numpy()
    w = (r.standard_normal((3, 3, Cin, 2)) / 10).astype(np.float32)
    g = r.standard_normal((N, H, W, 2)).astype(np.float32)
    xt = torch.tensor(y_prev, requires_grad=True)
    wt = torch.tensor(w, requires_grad=True)
    bt = torch.zeros(2, requires_grad=True)
    T.conv2d_same(xt, wt, bt, 1).backward(torch.tensor(g))
    exp_dz = xt.grad.numpy() * np.where(y_prev > 0, 1.0, 0.3)
    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    xd = torch.tensor(y_prev, device=cuda).to(tdt)
    dz = torch.empty_like(xd)
    dw = torch.empty((3, 3, Cin, 2), device=cuda)
    db = torch.empty((2,), device=cuda)
    gd, wd = dev(g, cuda), dev(w, cuda)                               # keep the device buffers alive across the call
    L.check(L.load().of_flow_head_bwd(L.ptr(xd), L.ptr(gd), L.ptr(wd), L.ptr(dw), L.ptr(db), L.ptr(dz),
                                      N, H, W, Cin, L.OF_ACT_LEAKY, L.dtype_code(xd), L.stream_ptr()), "of_flow_head_bwd")
    assert_close_maxrel(dw, wt.grad.numpy(), what="head dw")
    assert_close_maxrel(db, bt.grad.numpy(), what="head db")
    if dtype == "f32":
        assert_close_maxrel(dz, exp_dz, what="head dz")
    else:
        err = np.abs(dz.float().cpu().numpy() - exp_dz)
        assert (err <= 2.0 ** -8 * np.abs(exp_dz) + 1e-6).all(), err.max()


def test_cost_volume_shape_assert(cuda):
    from optical_flow_b200.model import create_cost_volume
    a = torch.zeros((1, 4, 4, 8), device=cuda)
    b = torch.zeros((1, 4, 5, 8), device=cuda)
    with pytest.raises(AssertionError):
        create_cost_volume(a, b, 3)


# ------------------------------------------------------------- fused losses --
@pytest.mark.parametrize("B,H,W", [(2, 32, 64), (1, 128, 416), (2, 48, 80)])
def test_loss_layer_fused(cuda, B, H, W):
    from optical_flow_b200.loss import LossLayer
    r = rng(12)
    imgs = T.synthetic_pairs(B, H, W, seed=5)
    flows = [(r.standard_normal((B, int(H / 2 ** (s + 1)), int(W / 2 ** (s + 1)), 2)) * (4.0 / 2 ** s)).astype(np.float32)
             for s in range(4)]
    e_loss, e_d = O.loss_layer(imgs, flows, return_grads=True)
    layer = LossLayer()
    loss = layer(dev(imgs, cuda), [dev(f, cuda) for f in flows])
    assert abs(loss.item() - float(e_loss)) <= REL * abs(float(e_loss))
    for s in range(4):
        assert_close_maxrel(layer.d_flows[s], e_d[s], what="d_flow scale %d" % s)


@pytest.mark.parametrize("mode", ["ref_ij", "xy"])
@pytest.mark.parametrize("B,H,W", [(1, 16, 16), (3, 48, 80), (2, 64, 208), (8, 384, 1280)])
def test_loss_fast_path_equals_generic(cuda, variant, B, H, W, mode):
    """The L2-pyramid fast path (loss_fused.cu; H, W multiples of 16) against the direct-gather
    kernel (loss.cu) on the same inputs, up to BASELINE's full 8 x 384 x 1280, twice in a row
    (the kernel re-arms its own counters)."""
    from optical_flow_b200 import functional as F
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + H)
    imgs = (torch.rand((B, H, W, 6), generator=g) - 0.5).to(cuda)
    flows = [(torch.randn((B, H >> (s + 1), W >> (s + 1), 2), generator=g) * (4.0 / 2 ** s)).to(cuda) for s in range(4)]
    variant("OF_LOSS_GENERIC", "1")
    l_ref, d_ref = F.loss_l1_pyramid(imgs, flows, mode=mode)
    variant("OF_LOSS_GENERIC", None)
    for _ in range(2):
        l_fast, d_fast = F.loss_l1_pyramid(imgs, flows, mode=mode)
        assert abs(l_fast.item() - l_ref.item()) <= 1e-6 * abs(l_ref.item())
        for a, b in zip(d_fast, d_ref):
            assert_close_maxrel(a, b.cpu().numpy(), rel=1e-6, what="d_flow")


def test_loss_layer_shape_assert(cuda):
    from optical_flow_b200.loss import LossLayer
    imgs = torch.zeros((1, 32, 64, 6), device=cuda)
    flows = [torch.zeros((1, 16, 32, 2), device=cuda), torch.zeros((1, 8, 15, 2), device=cuda)]
    with pytest.raises(AssertionError):
        LossLayer()(imgs, flows)


def test_loss_layer_deterministic(cuda):
    from optical_flow_b200.loss import LossLayer
    r = rng(13)
    imgs = dev(T.synthetic_pairs(2, 64, 96, seed=6), cuda)
    flows = [dev((r.standard_normal((2, 64 >> (s + 1), 96 >> (s + 1), 2)) * 2).astype(np.float32), cuda) for s in range(4)]
    vals = {LossLayer()(imgs, flows).item() for _ in range(5)}
    assert len(vals) == 1


@pytest.mark.parametrize("B,H,W", [(2, 32, 64), (1, 48, 80)])
def test_my_loss_fused(cuda, B, H, W):
    from optical_flow_b200.old.loss import MyLoss
    r = rng(14)
    imgs = T.synthetic_pairs(B, H, W, seed=7)
    flows = [(r.standard_normal((B, H // d, W // d, 2)) * 2.0).astype(np.float32) for d in (16, 8, 4, 2, 1)]
    e_loss, e_d = O.my_loss(imgs, flows, return_grads=True)
    loss, d_flows = MyLoss(dev(imgs, cuda), [dev(f, cuda) for f in flows], return_grads=True)
    assert abs(loss.item() - float(e_loss)) <= REL * abs(float(e_loss))
    for k in range(5):
        assert_close_maxrel(d_flows[k], e_d[k], what="d_flow %d" % k)
    # y_pred in the reference's 4-list form, with the motion term
    motions = [(r.standard_normal((B, H // d, W // d, 3))).astype(np.float32) for d in (16, 8, 4, 2, 1)]
    e2 = O.my_loss(imgs, flows, motions=motions)
    l2 = MyLoss(dev(imgs, cuda), [[dev(f, cuda) for f in flows], [dev(m, cuda) for m in motions], None, None])
    assert abs(l2.item() - float(e2)) <= 1e-5 * abs(float(e2))


# ---------------------------------------------------- pooling / elementwise --
def test_maxpool_add_relu_act(cuda):
    from optical_flow_b200 import functional as F, _lib as L
    r = rng(15)
    x = np.maximum(r.standard_normal((2, 8, 12, 16)), 0).astype(np.float32)   # post-ReLU input (ties at 0)
    xt = torch.tensor(x, requires_grad=True)
    pt = torch.nn.functional.max_pool2d(xt.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
    out = F.maxpool2_fwd(dev(x, cuda))
    assert np.array_equal(out.cpu().numpy(), pt.detach().numpy())
    g = r.standard_normal(pt.shape).astype(np.float32)
    pt.backward(torch.tensor(g))
    d_in = F.maxpool2_bwd(dev(x, cuda), dev(g, cuda)).cpu().numpy()
    # only compare where the window max is unique (ties at 0 are killed by the ReLU upstream)
    mask = x > 0
    np.testing.assert_allclose(d_in[mask], xt.grad.numpy()[mask], atol=0)
    assert abs(d_in.sum() - g.sum()) < 1e-3
    a, b = r.standard_normal((4, 32)).astype(np.float32), r.standard_normal((4, 32)).astype(np.float32)
    assert np.array_equal(F.add_relu_fwd(dev(a, cuda), dev(b, cuda)).cpu().numpy(), np.maximum(a + b, 0))
    y = r.standard_normal((4, 32)).astype(np.float32)
    dz = F.act_bwd(dev(y, cuda), dev(a, cuda), L.OF_ACT_LEAKY).cpu().numpy()
    np.testing.assert_allclose(dz, a * np.where(y > 0, 1.0, 0.3).astype(np.float32), rtol=1e-7)
    dz = F.act_bwd(dev(y, cuda), dev(a, cuda), L.OF_ACT_RELU).cpu().numpy()
    assert np.array_equal(dz, a * (y > 0))


# -------------------------------------------------------------- convolution --
@pytest.mark.parametrize("cfg", [
    # N, H, W, Cin, K, Cout, stride
    (2, 16, 20, 3, 7, 64, 2),
    (1, 12, 16, 64, 3, 64, 1),
    (2, 12, 16, 64, 3, 128, 2),
    (1, 12, 16, 64, 1, 128, 2),
    (1, 9, 13, 115, 3, 32, 1),
    (1, 8, 10, 32, 3, 2, 1),
    (1, 7, 9, 5, 3, 6, 2),
])
def test_conv_simt_fprop_dgrad_wgrad(cuda, cfg):
    from optical_flow_b200 import functional as F, _lib as L
    N, H, W, Cin, K, Cout, s = cfg
    r = rng(16)
    x = r.standard_normal((N, H, W, Cin)).astype(np.float32)
    w = (r.standard_normal((K, K, Cin, Cout)) / np.sqrt(K * K * Cin)).astype(np.float32)
    b = r.standard_normal((Cout,)).astype(np.float32)
    xt, wt, bt = (torch.tensor(v, requires_grad=True) for v in (x, w, b))
    yt = torch.nn.functional.leaky_relu(T.conv2d_same(xt, wt, bt, s), 0.3)
    y = F.conv2d_fprop(dev(x, cuda), dev(w, cuda), dev(b, cuda), stride=s, act=L.OF_ACT_LEAKY)
    assert tuple(y.shape) == tuple(yt.shape)
    assert_close_maxrel(y, yt.detach().numpy(), what="fprop")
    # backward of the linear part
    zt = T.conv2d_same(xt, wt, bt, s)
    g = r.standard_normal(tuple(zt.shape)).astype(np.float32)
    zt.backward(torch.tensor(g))
    dx = F.conv2d_dgrad(dev(g, cuda), dev(w, cuda), (H, W), stride=s)
    assert_close_maxrel(dx, xt.grad.numpy(), what="dgrad")
    dw, db = F.conv2d_wgrad(dev(x, cuda), dev(g, cuda), (K, K), stride=s)
    assert_close_maxrel(dw, wt.grad.numpy(), rel=2e-5, what="wgrad")
    assert_close_maxrel(db, bt.grad.numpy(), rel=2e-5, what="bgrad")


# ---------------------------------------------------------------- optimiser --
def test_adam_steps(cuda):
    from optical_flow_b200 import functional as F
    r = rng(17)
    n = 10007
    p = r.standard_normal(n).astype(np.float32)
    m = np.zeros(n, np.float32)
    v = np.zeros(n, np.float32)
    pd, md, vd = dev(p, cuda), dev(m, cuda), dev(v, cuda)
    for step in range(1, 4):
        g = r.standard_normal(n).astype(np.float32) * 0.01
        p, m, v = O.adam_step(p, g, m, v, step)
        F.adam_step(pd, dev(g, cuda), md, vd, step)
    # p: three updates of size ~lr = 1e-4 each known to ~1e-5 relative (m cancels) -> 5e-9 absolute
    np.testing.assert_allclose(pd.cpu().numpy(), p, rtol=1e-6, atol=5e-9)
    # m, v: one fp32 rounding of O(1e-3) values (FMA contraction order) -> absolute floor 1e-9
    np.testing.assert_allclose(md.cpu().numpy(), m, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(vd.cpu().numpy(), v, rtol=1e-6, atol=1e-13)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_maxpool2_bwd_add_act_equals_three_passes(cuda, dtype):
    """of_maxpool2_bwd_add_act == maxpool2_bwd, then += add, then act_bwd (the three passes it replaces)."""
    from optical_flow_b200 import functional as F, _lib as L
    r = rng(16)
    shape = (2, 12, 20, 64)
    x = torch.tensor(r.standard_normal(shape).astype(np.float32), device=cuda).to(dtype)
    add = torch.tensor(r.standard_normal(shape).astype(np.float32), device=cuda).to(dtype)
    d = torch.tensor(r.standard_normal((2, 6, 10, 64)).astype(np.float32), device=cuda).to(dtype)
    fused = F.maxpool2_bwd_add_act(x, d, add, L.OF_ACT_RELU)
    ref = (F.maxpool2_bwd(x, d).float() + add.float()) * (x.float() > 0)
    assert torch.equal(fused.float(), ref.to(dtype).float())


# ------------------------------------------------------------------ bf16 streaming kernels, 16-byte paths --
@pytest.mark.parametrize("shape", [(2, 12, 20, 64), (1, 6, 10, 24), (1, 4, 6, 12)])    # C % 8 == 0 twice, C % 8 != 0 once
def test_bf16_streaming_kernels_16_byte_paths(cuda, shape):
    """MaxPool2D forward, add + ReLU, activation backward and the stride-2 zero-insertion scatter on bf16 tensors:
    the 16-byte (8-channel) kernels against torch on the same bf16 values (exact: max / select / one rounded add)."""
    from optical_flow_b200 import functional as F, _lib as L
    from optical_flow_b200.conv_tc import PackCache, zero_upsample2
    B, H, W, C = shape
    g = torch.Generator(device="cpu").manual_seed(3)
    x = torch.randn(shape, generator=g).to(cuda).bfloat16()
    y = torch.randn(shape, generator=g).to(cuda).bfloat16()
    mp = F.maxpool2_fwd(x)
    exp = torch.nn.functional.max_pool2d(x.float().permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1).bfloat16()
    assert torch.equal(mp, exp)
    ar = F.add_relu_fwd(x, y)
    assert torch.equal(ar, torch.relu(x.float() + y.float()).bfloat16())
    for act, slope in ((L.OF_ACT_RELU, 0.0), (L.OF_ACT_LEAKY, 0.3)):
        dz = F.act_bwd(x, y, act)
        exp = (y.float() * torch.where(x.float() > 0, torch.ones_like(x.float()), torch.full_like(x.float(), slope))).bfloat16()
        assert torch.equal(dz, exp)
    cache = PackCache()
    up = zero_upsample2(x, 2 * H, 2 * W, cache=cache, key=1)
    exp = torch.zeros((B, 2 * H, 2 * W, C), dtype=torch.bfloat16, device=cuda)
    exp[:, ::2, ::2] = x
    assert torch.equal(up, exp)
    up2 = zero_upsample2(y, 2 * H, 2 * W, cache=cache, key=1)          # the cached buffer: only the quarter is rewritten
    exp[:, ::2, ::2] = y
    assert up2.data_ptr() == up.data_ptr() and torch.equal(up2, exp)
