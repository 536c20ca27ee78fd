"""conv_mode="bf16x3": the tensor-core parity mode (fp32 storage, every convolution as three bf16 tcgen05 passes over
hi / lo split operands) against the fp32 oracle - the bar the round-1 review set for a tensor-core parity mode is 1e-3;
the per-layer bars here are 2e-5 (the dropped lo x lo term is 2^-16 relative), the whole-step bars are the fp32 mode's."""
import numpy as np
import pytest
import torch

from oracle import ref_torch as T

pytestmark = pytest.mark.gpu


def rel_l2(got, exp):
    got = got.detach().double().cpu().numpy() if isinstance(got, torch.Tensor) else np.asarray(got, np.float64)
    exp = exp.detach().double().cpu().numpy() if isinstance(exp, torch.Tensor) else np.asarray(exp, np.float64)
    return float(np.linalg.norm(got - exp) / max(np.linalg.norm(exp), 1e-30))


def test_split_bf16_kat(cuda):
    from optical_flow_b200.conv_tc import split_bf16
    t = torch.randn((3, 5, 7, 13), device=cuda) * 3.0
    hi, lo = split_bf16(t)
    assert hi.dtype == torch.bfloat16 and tuple(hi.shape) == (3, 5, 7, 16)
    exp_hi = t.bfloat16()
    assert torch.equal(hi[..., :13], exp_hi) and torch.equal(lo[..., :13], (t - exp_hi.float()).bfloat16())
    assert float(hi[..., 13:].float().abs().max()) == 0.0 and float(lo[..., 13:].float().abs().max()) == 0.0
    rec = hi[..., :13].float() + lo[..., :13].float()
    assert float(((rec - t).abs() / t.abs().clamp_min(1e-20)).max()) < 2.0 ** -15
    fh, fl = split_bf16(t, pad_to=1, as_f32=True)
    assert fh.dtype == torch.float32 and torch.equal(fh, exp_hi.float()) and torch.equal(fl, (t - exp_hi.float()).bfloat16().float())


class _Layer:
    def __init__(self, cin, cout, k, stride, act):
        self.cin, self.cout, self.k, self.stride, self.act = cin, cout, k, stride, act


@pytest.mark.parametrize("cfg", [(2, 24, 32, 64, 3, 64, 1), (1, 16, 48, 115, 3, 128, 1), (2, 24, 32, 64, 3, 128, 2),
                                 (1, 24, 32, 64, 1, 128, 2), (1, 16, 32, 128, 3, 96, 1), (1, 12, 16, 256, 3, 256, 1),
                                 (1, 16, 16, 3, 7, 64, 2), (1, 16, 16, 32, 3, 2, 1), (1, 12, 16, 307, 3, 128, 1)])
def test_bf16x3_conv_matches_fp32(cuda, cfg):
    """fprop, dgrad, wgrad of one Conv2D('same') layer against torch fp32 (float64 reference), UNROUNDED fp32 inputs."""
    from optical_flow_b200 import _lib as L
    from optical_flow_b200.conv_tc import Bf16x3Backend
    N, H, W, Cin, K, Cout, s = cfg
    r = np.random.default_rng(5)
    x = r.standard_normal((N, H, W, Cin)).astype(np.float32)
    w = (r.standard_normal((K, K, Cin, Cout)) / np.sqrt(K * K * Cin)).astype(np.float32)
    b = r.standard_normal((Cout,)).astype(np.float32)
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    bt = torch.tensor(b, dtype=torch.float64, requires_grad=True)
    yt = torch.nn.functional.leaky_relu(T.conv2d_same(xt, wt, bt, s), 0.3)
    g = r.standard_normal(tuple(yt.shape)).astype(np.float32)
    yt.backward(torch.tensor(g, dtype=torch.float64))
    be = Bf16x3Backend(None)
    layer = _Layer(Cin, Cout, K, s, L.OF_ACT_LEAKY)
    xd, wd, bd, gd = (torch.tensor(a, device=cuda) for a in (x, w, b, g))
    y = be.fprop(layer, xd, wd, bd)
    assert y.dtype == torch.float32 and rel_l2(y, yt) < 2e-5, rel_l2(y, yt)
    dz = be.act_bwd(layer, y, gd)
    dw = torch.empty_like(wd); db = torch.empty_like(bd)
    be.wgrad(layer, xd, dz, dw, db)
    assert rel_l2(dw, wt.grad) < 2e-5, rel_l2(dw, wt.grad)
    assert rel_l2(db, bt.grad) < 2e-5
    dx = be.dgrad(layer, dz, wd, (H, W))
    assert tuple(dx.shape) == (N, H, W, Cin) and rel_l2(dx, xt.grad) < 2e-5, rel_l2(dx, xt.grad)


@pytest.fixture(scope="module")
def small_case():
    B, H, W = 2, 64, 96
    params, stats = T.init_params(seed=42, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=11)
    loss, flows, grads = T.FlowNetRef(params, stats).loss_and_grads(torch.tensor(imgs))
    return dict(B=B, H=H, W=W, params=params, stats=stats, imgs=imgs, loss=loss, flows=flows, grads=grads)


def test_bf16x3_forward_loss_backward_parity(cuda, small_case):
    """The whole network in the tensor-core parity mode against the fp32 oracle, at the fp32 mode's own bars
    (tests/test_b3_gpu_net.py::test_fp32_forward_loss_backward_parity)."""
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.model import build_flow_net
    net = build_flow_net(small_case["H"], small_case["W"], None, conv_mode="bf16x3")
    net.store.load_numpy(small_case["params"], small_case["stats"])
    imgs = torch.tensor(small_case["imgs"], device=cuda)
    flows = net(imgs)
    for k, (f, e) in enumerate(zip(flows, small_case["flows"])):
        assert rel_l2(f, e) < 1e-4, "flow %d: %g" % (k, rel_l2(f, e))
    layer = LossLayer()
    loss = layer(imgs, flows)
    assert abs(loss.item() - small_case["loss"].item()) < 1e-4 * abs(small_case["loss"].item())
    net.backward(layer.d_flows)
    worst = max(((name, rel_l2(g, small_case["grads"][name])) for name, g in net.store.g.items()
                 if float(small_case["grads"][name].abs().max()) > 0), key=lambda kv: kv[1])
    assert worst[1] < 1e-3, "worst gradient tensor %s: %g" % worst


def test_bf16x3_train_step_runs_in_a_graph(cuda, small_case):
    """Trainer(conv_mode="bf16x3"): three steps, loss equal to the fp32 mode's within 1e-4."""
    from optical_flow_b200.train import Trainer
    losses = {}
    for mode in ("fp32", "bf16x3"):
        tr = Trainer(small_case["H"], small_case["W"], None, conv_mode=mode)
        tr.flow_net.store.load_numpy(small_case["params"], small_case["stats"])
        imgs = torch.tensor(small_case["imgs"], device=cuda)
        losses[mode] = [float(tr.train_step(imgs, i)[0]) for i in range(3)]
    for a, b in zip(losses["fp32"], losses["bf16x3"]):
        assert abs(a - b) < 1e-4 * abs(a), (losses["fp32"], losses["bf16x3"])
