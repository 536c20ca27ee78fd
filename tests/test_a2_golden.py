"""Golden vectors produced by the reference's OWN Python (tests/golden/make_golden.py runs
/root/reference's model.py / loss.py / transformations.py / old/* over a torch-backed
TensorFlow stand-in).  CPU tests pin the oracle to them; GPU tests pin the CUDA path."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_np as R, ref_np_ops as O, ref_torch as T

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
OPS = np.load(os.path.join(GOLD, "ops.npz"))
NET = np.load(os.path.join(GOLD, "net_c1.npz"))


def close(got, exp, rel=1e-5, what=""):
    got = got.detach().float().cpu().numpy() if isinstance(got, torch.Tensor) else np.asarray(got)
    scale = max(float(np.abs(exp).max()), 1e-30)
    err = float(np.abs(got - exp).max())
    assert err <= rel * scale, "%s: max|err| %.3e > %.1e * %.3e" % (what, err, rel, scale)


# ------------------------------------------------------------- oracle vs golden (CPU) --
@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_warp_matches_reference_code(tag):
    feats, flow, g = OPS["warp_%s_feats" % tag], OPS["warp_%s_flow" % tag], OPS["warp_%s_g" % tag]
    assert np.array_equal(R.warp_features(flow, feats), OPS["warp_%s_out" % tag])      # bit-exact
    assert np.array_equal(R.bilinear_interpolation(feats, OPS["bil_%s_pts" % tag]), OPS["bil_%s_out" % tag])
    d_feats, d_flow, mag = R.warp_features_bwd(flow, feats, g, return_abs=True)
    bound = np.maximum(1e-5 * np.abs(d_feats).max(), 8 * np.finfo(np.float32).eps * mag)
    assert (np.abs(d_feats - OPS["warp_%s_dfeats" % tag]) <= bound).all()
    # the reference run is fp32: where a point lies far outside the image both taps collapse onto
    # one border pixel and weights of magnitude |w| >> 1 cancel; its own rounding
    # noise there is ~|w| * eps * |g||v|, so the bar scales with the per-pixel weight magnitude
    B, H, W, _ = feats.shape
    pts = (R.warp_grid(H, W) + flow).astype(np.float32)
    x0, x1, y0, y1 = R.sample_indices(pts, H, W)
    a, b = x1.astype(np.float32) - pts[..., 0], y1.astype(np.float32) - pts[..., 1]
    wmax = np.maximum.reduce([np.abs(a), np.abs(1 - a), np.abs(b), np.abs(1 - b), np.ones_like(a)])
    exp = OPS["warp_%s_dflow" % tag]
    assert (np.abs(d_flow - exp) <= 1e-5 * np.abs(exp).max() * wmax[..., None]).all()


def test_oracle_cost_volume_upscale_match_reference_code():
    close(O.create_cost_volume(OPS["cv_f1"], OPS["cv_f2"], 3), OPS["cv_out"], 1e-6, "cv")
    d1, d2 = O.create_cost_volume_bwd(OPS["cv_f1"], OPS["cv_f2"], 3, OPS["cv_g"])
    close(d1, OPS["cv_df1"], what="d_f1")
    close(d2, OPS["cv_df2"], what="d_f2")
    close(O.upscale_flow(OPS["up_flow"]), OPS["up_out"], 1e-6, "upscale")
    close(O.upscale_flow_bwd(OPS["up_g"]), OPS["up_dflow"], what="upscale bwd")


def test_oracle_loss_layer_matches_reference_code():
    flows = [OPS["loss_flow%d" % s] for s in range(4)]
    lv, d = O.loss_layer(OPS["loss_imgs"], flows, return_grads=True)
    assert abs(float(lv) - float(OPS["loss_value"])) <= 1e-6 * abs(float(OPS["loss_value"]))
    for s in range(4):
        close(d[s], OPS["loss_dflow%d" % s], what="d_flow%d" % s)


def test_oracle_image_warp_myloss_match_reference_code():
    assert np.array_equal(R.image_warp(OPS["iw_im"], OPS["iw_flow"]), OPS["iw_out"])   # bit-exact
    d_im, d_fl = R.image_warp_bwd(OPS["iw_im"], OPS["iw_flow"], OPS["iw_g"])
    close(d_im, OPS["iw_dim"], what="d_im")
    close(d_fl, OPS["iw_dflow"], what="d_flow")
    flows = [OPS["myloss_flow%d" % k] for k in range(5)]
    lv, d = O.my_loss(OPS["myloss_imgs"], flows, return_grads=True)
    assert abs(float(lv) - float(OPS["myloss_value"])) <= 2e-6 * abs(float(OPS["myloss_value"]))
    for k in range(5):
        close(d[k], OPS["myloss_dflow%d" % k], what="d_flow%d" % k)


def _digest(g):
    flat = np.asarray(g, dtype=np.float64).reshape(-1)
    step = max(flat.size // 256, 1)
    return flat.sum(), np.sqrt((flat * flat).sum()), flat[::step][:256].astype(np.float32)


def check_net_grads(grads, rel, what):
    """grads: name -> array.  Every tensor's norm and strided samples vs the reference run."""
    for key in NET.files:
        if not key.startswith("gnorm:"):
            continue
        name = key[6:]
        gs, gn, smp = _digest(grads[name])
        en, es = float(NET[key]), NET["gsmp:" + name]
        assert abs(gn - en) <= rel * max(en, 1e-12), "%s %s: |g| %.6e vs %.6e" % (what, name, gn, en)
        scale = max(float(np.abs(es).max()), 1e-30)
        assert float(np.abs(smp - es).max()) <= rel * scale * 4, "%s %s samples" % (what, name)


def test_oracle_network_matches_reference_code():
    """ref_torch.FlowNetRef (the oracle the GPU tests and bench use) vs model.py + loss.py +
    resnet/models.py executed as written: loss, four flows, every weight gradient."""
    B, H, W, seed, dseed = [int(v) for v in NET["meta"]]
    params, stats = T.init_params(seed=seed, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=dseed)
    loss, flows, grads = T.FlowNetRef(params, stats).loss_and_grads(torch.tensor(imgs))
    assert abs(loss.item() - float(NET["loss"])) <= 1e-5 * abs(float(NET["loss"]))
    for s in range(4):
        close(flows[s], NET["flow%d" % s], 1e-4, "flow%d" % s)
    check_net_grads({k: v.numpy() for k, v in grads.items()}, 2e-3, "oracle")


# --------------------------------------------------------- CUDA path vs golden (GPU) --
def _dev(a, cuda):
    return torch.tensor(np.ascontiguousarray(a), device=cuda)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_cuda_warp_matches_reference_code(cuda, tag):
    from optical_flow_b200 import functional as F
    from optical_flow_b200.model import warp_features
    from optical_flow_b200.transformations import bilinear_interpolation
    feats, flow, g = OPS["warp_%s_feats" % tag], OPS["warp_%s_flow" % tag], OPS["warp_%s_g" % tag]
    out = warp_features(_dev(flow, cuda), _dev(feats, cuda))
    assert np.array_equal(out.cpu().numpy(), OPS["warp_%s_out" % tag])                   # bit-exact
    out = bilinear_interpolation(_dev(feats, cuda), _dev(OPS["bil_%s_pts" % tag], cuda))
    assert np.array_equal(out.cpu().numpy(), OPS["bil_%s_out" % tag])
    d_feats, d_flow = F.warp_features_bwd(_dev(flow, cuda), _dev(feats, cuda), _dev(g, cuda))
    _, _, mag = R.warp_features_bwd(flow, feats, g, return_abs=True)
    exp = OPS["warp_%s_dfeats" % tag]
    bound = np.maximum(1e-5 * np.abs(exp).max(), 8 * np.finfo(np.float32).eps * mag)
    assert (np.abs(d_feats.cpu().numpy() - exp) <= bound).all()
    B, H, W, _ = feats.shape
    pts = (R.warp_grid(H, W) + flow).astype(np.float32)
    x0, x1, y0, y1 = R.sample_indices(pts, H, W)
    a, b = x1.astype(np.float32) - pts[..., 0], y1.astype(np.float32) - pts[..., 1]
    wmax = np.maximum.reduce([np.abs(a), np.abs(1 - a), np.abs(b), np.abs(1 - b), np.ones_like(a)])
    exp = OPS["warp_%s_dflow" % tag]
    assert (np.abs(d_flow.cpu().numpy() - exp) <= 1e-5 * np.abs(exp).max() * wmax[..., None]).all()


@pytest.mark.gpu
def test_cuda_cost_volume_upscale_match_reference_code(cuda):
    from optical_flow_b200 import functional as F
    f1, f2, g = _dev(OPS["cv_f1"], cuda), _dev(OPS["cv_f2"], cuda), _dev(OPS["cv_g"], cuda)
    close(F.cost_volume_fwd(f1, f2, 3), OPS["cv_out"], what="cv")
    d1, d2 = F.cost_volume_bwd(f1, f2, g, 3)
    close(d1, OPS["cv_df1"], what="d_f1")
    close(d2, OPS["cv_df2"], what="d_f2")
    close(F.upscale_flow_fwd(_dev(OPS["up_flow"], cuda)), OPS["up_out"], what="upscale")
    close(F.upscale_flow_bwd(_dev(OPS["up_g"], cuda)), OPS["up_dflow"], what="upscale bwd")


@pytest.mark.gpu
def test_cuda_losses_match_reference_code(cuda):
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.old.image_warp import image_warp, image_warp_backward
    from optical_flow_b200.old.loss import MyLoss
    layer = LossLayer()
    lv = layer(_dev(OPS["loss_imgs"], cuda), [_dev(OPS["loss_flow%d" % s], cuda) for s in range(4)])
    assert abs(lv.item() - float(OPS["loss_value"])) <= 1e-5 * abs(float(OPS["loss_value"]))
    for s in range(4):
        close(layer.d_flows[s], OPS["loss_dflow%d" % s], what="d_flow%d" % s)
    im, fl, g = _dev(OPS["iw_im"], cuda), _dev(OPS["iw_flow"], cuda), _dev(OPS["iw_g"], cuda)
    assert np.array_equal(image_warp(im, fl).cpu().numpy(), OPS["iw_out"])               # bit-exact
    d_im, d_fl = image_warp_backward(im, fl, g)
    close(d_im, OPS["iw_dim"], what="d_im")
    close(d_fl, OPS["iw_dflow"], what="iw d_flow")
    lv, d = MyLoss(_dev(OPS["myloss_imgs"], cuda), [_dev(OPS["myloss_flow%d" % k], cuda) for k in range(5)],
                   return_grads=True)
    assert abs(lv.item() - float(OPS["myloss_value"])) <= 1e-5 * abs(float(OPS["myloss_value"]))
    for k in range(5):
        close(d[k], OPS["myloss_dflow%d" % k], what="myloss d_flow%d" % k)


@pytest.mark.gpu
def test_cuda_network_fp32_matches_reference_code(cuda):
    """The whole train-step graph (exact-parity fp32 conv mode) against model.py + loss.py run
    as written, BASELINE configs[0] (batch 1, 128x416): loss, flows, every weight gradient."""
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.model import build_flow_net
    B, H, W, seed, dseed = [int(v) for v in NET["meta"]]
    params, stats = T.init_params(seed=seed, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=dseed)
    net = build_flow_net(H, W, None, conv_mode="fp32")
    net.store.load_numpy(params, stats)
    x = torch.tensor(imgs, device=cuda)
    flows = net(x)
    layer = LossLayer()
    loss = layer(x, flows)
    assert abs(loss.item() - float(NET["loss"])) <= 1e-4 * abs(float(NET["loss"]))
    for s in range(4):
        close(flows[s], NET["flow%d" % s], 1e-3, "flow%d" % s)
    net.backward(layer.d_flows)
    check_net_grads({k: v.detach().cpu().numpy() for k, v in net.store.g.items()}, 2e-2, "cuda fp32")
