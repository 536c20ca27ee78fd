"""CPU tests of the oracle: the authored known-answer tests of SURVEY 8c and the
two-implementation rule (NumPy literal vs torch autograd).  The reference has no
tests/golden vectors of its own (parity unpinned), so these pin the oracle to the
cited reference lines' semantics."""
import numpy as np
import pytest
import torch

from oracle import ref_np as R, ref_np_ops as O, ref_torch as T

RNG = np.random.default_rng(7)


def rand(*shape, scale=1.0):
    return (RNG.standard_normal(shape) * scale).astype(np.float32)


# ---- (1) zero flow, REF_IJ: out[b,i,j] == in[b, min(j,H-1), i] ------------------
def test_kat_zero_flow_ref_ij_transposes():
    B, H, W, C = 2, 6, 9, 5
    f = rand(B, H, W, C)
    out = R.warp_features(np.zeros((B, H, W, 2), np.float32), f)
    exp = np.stack([[[f[b, min(j, H - 1), i] for j in range(W)] for i in range(H)] for b in range(B)])
    # exact where no clamping occurs; clamped taps blend two copies of the same row
    assert np.array_equal(out[:, :, :H], exp[:, :, :H])
    np.testing.assert_allclose(out, exp, rtol=0, atol=1e-6)


# ---- (2) zero flow, XY mode and UnFlow warp: identity, bit-exact ---------------
def test_kat_zero_flow_identity_xy_and_unflow():
    f = rand(2, 5, 7, 3)
    z = np.zeros((2, 5, 7, 2), np.float32)
    assert np.array_equal(R.warp_xy(z, f), f)
    assert np.array_equal(R.image_warp(f, z), f)


# ---- (3) integer flow: shifted copy with border replication --------------------
def test_kat_integer_flow_shift():
    f = rand(1, 6, 8, 2)
    flow = np.zeros((1, 6, 8, 2), np.float32)
    flow[..., 0] = 2.0   # x + 2
    flow[..., 1] = -1.0  # y - 1
    exp = np.zeros_like(f)
    for i in range(6):
        for j in range(8):
            exp[0, i, j] = f[0, min(max(i - 1, 0), 5), min(max(j + 2, 0), 7)]
    assert np.array_equal(R.image_warp(f, flow), exp)
    # the reference sampler blends two copies of the border pixel when a tap is
    # clamped (weights from the clipped x1), so it is exact only up to rounding
    np.testing.assert_allclose(R.warp_xy(flow, f), exp, rtol=0, atol=1e-6)


# ---- (4) half-integer flow on a linear ramp -------------------------------------
def test_kat_half_integer_ramp():
    H, W = 6, 10
    jj = np.arange(W, dtype=np.float32).reshape(1, 1, W, 1)
    ii = np.arange(H, dtype=np.float32).reshape(1, H, 1, 1)
    ramp = (3.0 * jj + 5.0 * ii).astype(np.float32) * np.ones((1, H, W, 1), np.float32)
    flow = np.full((1, H, W, 2), 0.5, np.float32)
    out = R.image_warp(ramp, flow)
    exp = ramp + 3.0 * 0.5 + 5.0 * 0.5
    np.testing.assert_allclose(out[:, :-1, :-1], exp[:, :-1, :-1], rtol=0, atol=1e-5)
    out2 = R.warp_xy(flow, ramp)
    np.testing.assert_allclose(out2[:, :-1, :-1], exp[:, :-1, :-1], rtol=0, atol=1e-5)


# ---- (5) far out-of-range: border value, weights sum to 1, d_flow = 0 ----------
def test_kat_out_of_range_border():
    f = rand(1, 4, 5, 3)
    # moderately out of range: the weights a*b etc. grow like distance^2 and cancel
    # (fp32), so the replicate-border value is exact only up to ~d^2 * eps
    pts = np.full((1, 4, 5, 2), 9.25, np.float32)
    out = R.bilinear_interpolation(f, pts)
    np.testing.assert_allclose(out, np.broadcast_to(f[0, 3, 4], out.shape), rtol=0, atol=2e-5)
    _, d_pts = R.bilinear_interpolation_bwd(f, pts, np.ones_like(f))
    assert np.all(d_pts == 0)


# ---- (6) index op order: (A) floor(pos + flow) vs (B) pos + floor(flow) ---------
def test_kat_index_op_order_differs_between_variants():
    # pos + flow rounds up to an integer in fp32, floor(flow) does not
    W = 4096
    flow = np.zeros((1, 1, W, 2), np.float32)
    flow[0, 0, :, 0] = np.float32(-1e-7)    # tiny negative x displacement
    (x0a, _, _, _) = R.sample_indices((np.ascontiguousarray(R.warp_grid(1, W)[..., ::-1]) + flow).astype(np.float32), 1, W)
    (x0b, _, _, _), _ = R.image_warp_indices(flow, 1, W)
    j = 3000
    assert x0a[0, 0, j] == j          # fp32: 3000 - 1e-7 == 3000
    assert x0b[0, 0, j] == j - 1      # floor(-1e-7) == -1


# ---- (7) resize: central 2x2 mean / {0.25,0.75} stencil -------------------------
@pytest.mark.parametrize("k", [2, 4, 8, 16])
def test_kat_resize_down_is_central_2x2_mean(k):
    x = rand(2, 32, 48, 6)
    out = O.resize_bilinear(x, 32 // k, 48 // k)
    c = k // 2
    exp = 0.25 * (x[:, c - 1::k, c - 1::k] + x[:, c - 1::k, c::k] + x[:, c::k, c - 1::k] + x[:, c::k, c::k])
    np.testing.assert_allclose(out, exp, rtol=0, atol=1e-6)


def test_kat_upscale_stencil_and_torch():
    x = rand(1, 4, 5, 2)
    up = O.upscale_flow(x)
    # interior even/odd output rows use {0.25, 0.75} weights; edges clamp
    np.testing.assert_allclose(up[0, 0, 0], 2 * x[0, 0, 0], atol=1e-6)
    np.testing.assert_allclose(up[0, 1, 0], 2 * (0.75 * x[0, 0, 0] + 0.25 * x[0, 1, 0]), atol=1e-6)
    np.testing.assert_allclose(up[0, 2, 0], 2 * (0.25 * x[0, 0, 0] + 0.75 * x[0, 1, 0]), atol=1e-6)
    np.testing.assert_allclose(up, T.upscale_flow(torch.tensor(x)).numpy(), atol=1e-6)
    g = rand(*up.shape)
    xt = torch.tensor(x, requires_grad=True)
    T.upscale_flow(xt).backward(torch.tensor(g))
    np.testing.assert_allclose(O.upscale_flow_bwd(g), xt.grad.numpy(), atol=1e-5)


@pytest.mark.parametrize("oh,ow", [(5, 7), (12, 9), (3, 3)])
def test_resize_general_matches_torch(oh, ow):
    x = rand(2, 16, 24, 3)
    np.testing.assert_allclose(O.resize_bilinear(x, oh, ow), T.resize_bilinear(torch.tensor(x), oh, ow).numpy(),
                               atol=1e-6)


# ---- (8) cost volume -------------------------------------------------------------
def test_kat_cost_volume_explicit_loop():
    f1, f2 = rand(1, 5, 6, 4), rand(1, 5, 6, 4)
    cv = O.create_cost_volume(f1, f2, 3)
    assert cv.shape == (1, 5, 6, 49)
    np.testing.assert_allclose(cv[..., 24], np.sum(f1 * f2, axis=-1), rtol=1e-6, atol=1e-6)
    for (y, x, i, j) in [(0, 0, 0, 0), (2, 3, 6, 1), (4, 5, 3, 6), (1, 1, 2, 5)]:
        yy, xx = y + i - 3, x + j - 3
        exp = 0.0 if not (0 <= yy < 5 and 0 <= xx < 6) else float(np.dot(f1[0, y, x], f2[0, yy, xx]))
        assert abs(cv[0, y, x, i * 7 + j] - exp) < 1e-5


# ---- (9) TF SAME padding table ---------------------------------------------------
def test_kat_same_padding_table():
    assert O.tf_same_pad(384, 7, 2) == (192, 2, 3)
    assert O.tf_same_pad(96, 3, 2) == (48, 0, 1)
    assert O.tf_same_pad(96, 1, 2) == (48, 0, 0)
    assert O.tf_same_pad(96, 3, 1) == (96, 1, 1)
    assert T.tf_same_pad(384, 7, 2) == (192, 2, 3)


# ---- (10) two-implementation rule: NumPy vs torch autograd + fp64 differences ----
def test_two_impl_warp_fwd_bwd():
    B, H, W, C = 2, 7, 9, 4
    f, fl, g = rand(B, H, W, C), rand(B, H, W, 2, scale=3), rand(B, H, W, C)
    for mode, fw, bw in (("ref_ij", R.warp_features, R.warp_features_bwd), ("xy", R.warp_xy, R.warp_xy_bwd)):
        ft, flt = torch.tensor(f, requires_grad=True), torch.tensor(fl, requires_grad=True)
        o = T.warp_features(flt, ft, mode)
        assert np.array_equal(fw(fl, f), o.detach().numpy())
        o.backward(torch.tensor(g))
        d_f, d_fl = bw(fl, f, g)
        np.testing.assert_allclose(d_f, ft.grad.numpy(), atol=2e-5)
        np.testing.assert_allclose(d_fl, flt.grad.numpy(), atol=2e-5)


def test_two_impl_image_warp_and_central_difference():
    B, H, W, C = 1, 6, 7, 3
    f, fl, g = rand(B, H, W, C), rand(B, H, W, 2, scale=2), rand(B, H, W, C)
    ft, flt = torch.tensor(f, requires_grad=True), torch.tensor(fl, requires_grad=True)
    o = T.image_warp(ft, flt)
    assert np.array_equal(R.image_warp(f, fl), o.detach().numpy())
    o.backward(torch.tensor(g))
    d_im, d_fl = R.image_warp_bwd(f, fl, g)
    np.testing.assert_allclose(d_im, ft.grad.numpy(), atol=1e-5)
    np.testing.assert_allclose(d_fl, flt.grad.numpy(), atol=1e-5)
    # fp64 central difference on one flow entry (away from an integer crossing)
    f64, fl64 = torch.tensor(f, dtype=torch.float64), torch.tensor(fl, dtype=torch.float64)
    g64 = torch.tensor(g, dtype=torch.float64)
    h = 1e-6
    for (i, j, c) in [(2, 3, 0), (4, 1, 1)]:
        if abs(fl[0, i, j, c] - round(float(fl[0, i, j, c]))) < 1e-3:
            continue
        p, m = fl64.clone(), fl64.clone()
        p[0, i, j, c] += h
        m[0, i, j, c] -= h
        num = ((T.image_warp(f64, p) * g64).sum() - (T.image_warp(f64, m) * g64).sum()) / (2 * h)
        assert abs(num.item() - d_fl[0, i, j, c]) < 1e-4


def test_two_impl_cost_volume_bwd():
    f1, f2, g = rand(2, 5, 7, 4), rand(2, 5, 7, 4), rand(2, 5, 7, 49)
    a, b = torch.tensor(f1, requires_grad=True), torch.tensor(f2, requires_grad=True)
    cv = T.create_cost_volume(a, b, 3)
    np.testing.assert_allclose(O.create_cost_volume(f1, f2, 3), cv.detach().numpy(), atol=1e-5)
    cv.backward(torch.tensor(g))
    d1, d2 = O.create_cost_volume_bwd(f1, f2, 3, g)
    np.testing.assert_allclose(d1, a.grad.numpy(), atol=1e-5)
    np.testing.assert_allclose(d2, b.grad.numpy(), atol=1e-5)
    # the tape-free large-shape form used by tests/test_b4_bench_shapes.py
    cv3, e1, e2 = T.create_cost_volume_fwd_bwd(torch.tensor(f1), torch.tensor(f2), 3, torch.tensor(g))
    np.testing.assert_allclose(cv3.numpy(), cv.detach().numpy(), atol=1e-6)
    np.testing.assert_allclose(e1.numpy(), a.grad.numpy(), atol=1e-5)
    np.testing.assert_allclose(e2.numpy(), b.grad.numpy(), atol=1e-5)


def test_two_impl_loss_layer():
    imgs = T.synthetic_pairs(2, 32, 64, seed=1)
    flows = [rand(2, 32 >> (s + 1), 64 >> (s + 1), 2, scale=2) for s in range(4)]
    loss, d_flows = O.loss_layer(imgs, flows, return_grads=True)
    ft = [torch.tensor(x, requires_grad=True) for x in flows]
    lt = T.loss_layer(torch.tensor(imgs), ft)
    lt.backward()
    assert abs(loss - lt.item()) < 1e-6 * max(1.0, abs(lt.item()))
    for d, t in zip(d_flows, ft):
        np.testing.assert_allclose(d, t.grad.numpy(), atol=1e-5 * np.abs(d).max() + 1e-12)


def test_two_impl_my_loss():
    imgs = T.synthetic_pairs(2, 32, 64, seed=2)
    flows = [rand(2, 32 // d, 64 // d, 2, scale=2) for d in (16, 8, 4, 2, 1)]
    loss, d_flows = O.my_loss(imgs, flows, return_grads=True)
    ft = [torch.tensor(x, requires_grad=True) for x in flows]
    lt = T.my_loss(torch.tensor(imgs), ft)
    lt.backward()
    assert abs(loss - lt.item()) < 1e-5 * abs(lt.item())
    for d, t in zip(d_flows, ft):
        np.testing.assert_allclose(d, t.grad.numpy(), atol=1e-5 * np.abs(d).max())


# ---- (11) Adam closed form ---------------------------------------------------------
def test_kat_adam_first_step_closed_form():
    p, g = rand(100), rand(100)
    p1, m1, v1 = O.adam_step(p, g, np.zeros(100, np.float32), np.zeros(100, np.float32), step=1)
    # t=1: m = 0.1 g, v = 0.001 g^2, lr_t = lr*sqrt(0.001)/0.1 -> p - lr * g/(|g| + eps*sqrt(1000)) approx
    lr_t = 1e-4 * np.sqrt(1 - 0.999) / (1 - 0.9)
    exp = p - lr_t * (0.1 * g) / (np.sqrt(0.001 * g * g) + 1e-7)
    np.testing.assert_allclose(p1, exp, rtol=1e-5, atol=1e-9)
    pt, mt, vt = T.adam_step(torch.tensor(p), torch.tensor(g), torch.zeros(100), torch.zeros(100), 1)
    np.testing.assert_allclose(p1, pt.numpy(), rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(m1, mt.numpy(), rtol=1e-6)
    np.testing.assert_allclose(v1, vt.numpy(), rtol=1e-6)


# ---- network: shapes, parameter count, TF-default details ---------------------------
def test_flow_net_ref_shapes_and_param_count():
    params, stats = T.init_params(seed=42)
    n = sum(int(np.prod(v.shape)) for v in params.values())
    assert n == 4937608
    net = T.FlowNetRef(params, stats)
    imgs = torch.tensor(T.synthetic_pairs(1, 32, 64, seed=3))
    with torch.no_grad():
        flows = net(imgs)
    assert [tuple(f.shape) for f in flows] == [(1, 16, 32, 2), (1, 8, 16, 2), (1, 4, 8, 2), (1, 2, 4, 2)]


def test_conv_same_asymmetric_padding_stride2():
    # 3x3/2 on even size pads (0,1): output (0,0) must not see a top/left halo
    x = torch.zeros(1, 4, 4, 1)
    x[0, 0, 0, 0] = 1.0
    w = torch.zeros(3, 3, 1, 1)
    w[0, 0, 0, 0] = 1.0       # top-left tap
    y = T.conv2d_same(x, w, None, 2)
    assert y.shape == (1, 2, 2, 1) and y[0, 0, 0, 0] == 1.0
