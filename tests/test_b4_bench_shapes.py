"""GPU parity AT THE BENCHMARKED CONFIGURATIONS (BASELINE.json configs[1], [2], [4]).

Everything else in tests/ runs at shapes the oracle finishes instantly; this file is the evidence that the
tile / TMA-box / grid arithmetic is also right where bench.py measures:

  (i)   LossLayer fast path (loss_fused.cu) vs oracle.ref_np_ops.loss_layer at 8 x 384 x 1280 and 1 x 1024 x 2048
        (loss.py:5-32): loss 1e-5 relative, every d_flow within 1e-5 of max-abs plus the cancellation-aware term of
        dflow_bound() (clipped taps carry weights of magnitude w - h in the reference's transposed sampling).
  (ii)  a whole fp32-mode step (forward, loss, all gradients) vs oracle.ref_torch.FlowNetRef at 1 x 384 x 1280.
  (iii) bf16 tensor-core mode vs the repo's own fp32 mode at 2 x 384 x 1280 with the warp coordinates pinned to the
        fp32 run's and the SAME d_flows fed to both backward passes: per-tensor relative L2 of every weight gradient
        (stated bound below), plus a 30-step loss-curve agreement of the two modes.
  (iv)  conv_tc3 / conv_tc_wgrad / cvtc kernels at the exact layer shapes of the C2 step (decoder level 3 at
        N = 8, 192 x 640: 115 -> 128 two-source, 128 -> 128, 128 -> 96, 96 -> 64, 64 -> 32; level 0 at 24 x 80:
        305 -> 128 two-source) vs torch-CPU fp32 convolution on the same bf16-rounded operands (model.py:104-114).
  (v)   the 1024 x 2048 (configs[4]) network shape: fp32-mode forward + loss vs the oracle at batch 1.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ref_np as R
from oracle import ref_np_ops as O
from oracle import ref_torch as T

pytestmark = pytest.mark.gpu

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")


def record(name, payload):
    """Measured error figures go to gpurun_out/parity_<name>.json (scratch; summarised in profiles/ by hand)."""
    try:
        os.makedirs(OUT, exist_ok=True)
        with open(os.path.join(OUT, "parity_%s.json" % name), "w") as f:
            json.dump(payload, f, indent=1, sort_keys=True)
    except OSError:
        pass


def bf16_round(a):
    return torch.tensor(a).bfloat16().float().numpy()


def rel_err(got, exp):
    got = got.detach().float().cpu().numpy() if isinstance(got, torch.Tensor) else np.asarray(got)
    exp = exp.detach().cpu().numpy() if isinstance(exp, torch.Tensor) else np.asarray(exp)
    return float(np.abs(got - exp).max()) / max(float(np.abs(exp).max()), 1e-30)


def rel_l2(got, exp):
    got = got.detach().float().cpu().numpy().ravel().astype(np.float64)
    exp = (exp.detach().float().cpu().numpy() if isinstance(exp, torch.Tensor) else np.asarray(exp)).ravel().astype(np.float64)
    return float(np.linalg.norm(got - exp) / max(np.linalg.norm(exp), 1e-30))


def dflow_bound(flow, e_d, num_scales=4):
    """Cancellation-aware fp32 bound for d loss / d flow of LossLayer: 1e-5 of max|ref| plus the rounding of the two
    products that cancel when a tap pair is clipped.  The reference forms its weights from the CLIPPED x1 / y1
    (transformations.py:98-112), and its transposed sampling (model.py:65-71: y = column + flow, clipped to h - 1)
    puts every column j >= h out of range, so |weight| reaches w - h (hundreds): d = a (v10 - v00) + (1 - a)(v11 - v01)
    with v10 - v00 == v11 - v01 is the difference of two products ~|a| times larger than the result, and an fp32
    kernel legitimately carries |a| ulps there (the float64 oracle does not)."""
    B, h, w, _ = flow.shape
    pts = (R.warp_grid(h, w) + flow).astype(np.float32)
    x0, x1, y0, y1 = R.sample_indices(pts, h, w)
    ax = np.abs(x1.astype(np.float32) - pts[..., 0])
    ay = np.abs(y1.astype(np.float32) - pts[..., 1])
    wx, wy = ax + np.abs(1 - ax), ay + np.abs(1 - ay)
    wmag = np.stack([wy, wx], axis=-1)                                        # d0 uses (ay, by); d1 uses (ax, bx)
    inv_n = 1.0 / (B * h * w * 3.0 * num_scales)
    eps = 2.0 ** -24
    bound = 1e-5 * np.abs(e_d).max() + 4 * eps * wmag * (3 * 1.2) * inv_n     # 3 channels, |pixel difference| <= 1.2
    # the forward value sum_k w_k v_k rounds against sum |w_k| |v_k| <= wx wy max|v|: below that, sign(diff) is a tie
    tie = 2e-6 + 8 * eps * wx * wy * 0.6
    return bound, tie


# ------------------------------------------------------------------ (i) fused loss vs the oracle --
@pytest.mark.parametrize("B,H,W", [(8, 384, 1280), (1, 1024, 2048), (4, 1024, 2048)])
@pytest.mark.parametrize("flows_kind", ["iid", "smooth"])
def test_loss_layer_oracle_at_bench_shape(cuda, B, H, W, flows_kind):
    from optical_flow_b200.loss import LossLayer
    if B == 4 and flows_kind == "smooth":
        pytest.skip("one flow statistic is enough at the largest shape")
    g = torch.Generator(device="cpu").manual_seed(B * 7 + H)
    imgs = T.synthetic_pairs(B, H, W, seed=5)
    flows = []
    for s in range(4):
        h, w = H >> (s + 1), W >> (s + 1)
        if flows_kind == "iid":
            f = torch.randn((B, h, w, 2), generator=g) * (4.0 / 2 ** s)
        else:
            c = torch.randn((B, 2, max(h // 8, 2), max(w // 8, 2)), generator=g)
            f = torch.nn.functional.interpolate(c, size=(h, w), mode="bilinear", align_corners=False)
            f = (f * (4.0 / 2 ** s) / f.std()).permute(0, 2, 3, 1).contiguous()
        flows.append(f.numpy().astype(np.float32))
    e_loss, e_d, e_diff = O.loss_layer(imgs, flows, return_grads=True, return_diffs=True)
    # d loss / d flow carries sign(im1 - warped) (loss.py:28): where a channel's |diff| is within a few fp32 ulps of 0
    # the sign is a rounding tie-break and that pixel's whole gradient may legitimately flip; with millions of pixels
    # such ties exist (a few 1e-5 of the in-range pixels; more where clipped taps carry weights in the hundreds and the
    # forward value itself only holds ~1e-4), so they are excluded from the max-abs bound - and counted
    bounds = [dflow_bound(flows[s], e_d[s]) for s in range(4)]
    firm = [np.abs(e_diff[s]).min(axis=-1) >= bounds[s][1] for s in range(4)]
    for s in range(4):
        assert firm[s].mean() > 0.99, (s, firm[s].mean())
    layer = LossLayer()
    x = torch.tensor(imgs, device=cuda)
    fd = [torch.tensor(f, device=cuda) for f in flows]
    for rep in range(2):                                   # twice: the kernel re-arms its own counters
        loss = layer(x, fd)
        assert abs(loss.item() - float(e_loss)) <= 1e-5 * abs(float(e_loss)), (loss.item(), float(e_loss))
        for s in range(4):
            got = layer.d_flows[s].cpu().numpy()
            excess = (np.abs(got - e_d[s]) - bounds[s][0]).max(axis=-1)
            assert excess[firm[s]].max() <= 0, (s, excess[firm[s]].max() / np.abs(e_d[s]).max())


# ------------------------------------------------------------------ (ii) fp32 step vs the oracle --
@pytest.fixture(scope="module")
def oracle_step_384x1280():
    B, H, W = 1, 384, 1280
    params, stats = T.init_params(seed=42, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=21)
    return (params, stats, imgs) + tuple(T.FlowNetRef(params, stats).loss_and_grads(torch.tensor(imgs)))


@pytest.mark.parametrize("mode", ["fp32", "bf16x3"])
def test_fp32_step_oracle_at_384x1280(cuda, oracle_step_384x1280, mode):
    """The fp32-accurate modes (SIMT fp32; tensor-core parity mode bf16x3) against the oracle at the benchmark resolution."""
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.model import build_flow_net
    B, H, W = 1, 384, 1280
    params, stats, imgs, r_loss, r_flows, r_grads = oracle_step_384x1280
    net = build_flow_net(H, W, None, conv_mode=mode)
    net.store.load_numpy(params, stats)
    x = torch.tensor(imgs, device=cuda)
    flows = net(x)
    for k, (f, e) in enumerate(zip(flows, r_flows)):
        assert rel_err(f, e) < (1e-4 if mode == "fp32" else 5e-4), "flow %d: %g" % (k, rel_err(f, e))
    layer = LossLayer()
    loss = layer(x, flows)
    assert abs(loss.item() - r_loss.item()) < 1e-4 * abs(r_loss.item())
    net.backward(layer.d_flows)
    rows = {}
    for name, gt in net.store.g.items():
        e = r_grads[name]
        if float(e.norm()) == 0:
            continue
        rows[name] = rel_l2(gt, e)
    record("%s_step_384x1280" % mode, {"loss": loss.item(), "ref_loss": r_loss.item(), "rel_l2": rows})
    # a warp index flips where an fp32-noisy flow crosses an integer (isolated pixels); relative L2 per tensor
    worst = max(rows.items(), key=lambda kv: kv[1])
    med = float(np.median(list(rows.values())))
    if mode == "fp32":
        assert worst[1] < 1e-3, worst      # measured on B200: worst 1.1e-4 (flow1_conv3/kernel), median 1.8e-5
    else:
        # hi + lo bf16 operands carry 16-17 mantissa bits (fp32: 24, tf32: 11): ~5e-6 per convolution, amplified through
        # 40 layers and the sign() / floor() decisions of loss.py:28 and the warp.  Measured on B200: loss 2e-6 off,
        # worst tensor 2.0e-3 (flow0_conv2/kernel), median 4.5e-4 - against 2.4e-2 / 1.1e-2 for the bf16 mode
        assert worst[1] < 5e-3 and med < 1e-3, (worst, med)


# ------------------------------------------------------------------ (iii) bf16 vs fp32 mode --
BF16_GRAD_REL_L2 = 0.05          # stated bf16 tolerance per weight tensor with pinned warp coordinates (measured on
BF16_GRAD_REL_L2_MEDIAN = 0.02   # B200: worst 0.0205 conv1/kernel, median 0.0108; profiles/r02_parity_summary.md)


def test_bf16_vs_fp32_mode_frozen_warp_384x1280(cuda):
    """Same weights, same images, warp coordinates of the bf16 run pinned to the fp32 run's up-scaled flows, the
    SAME d_flows (the fp32 run's) pushed through both backward passes: what remains is conv arithmetic only."""
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.model import build_flow_net
    B, H, W = 2, 384, 1280
    params, stats = T.init_params(seed=42, randomize_bn=True)
    x = torch.tensor(T.synthetic_pairs(B, H, W, seed=31), device=cuda)
    ref = build_flow_net(H, W, None, conv_mode="fp32")
    ref.store.load_numpy(params, stats)
    f_ref = ref(x)
    layer = LossLayer()
    l_ref = layer(x, f_ref).item()
    d_flows = [d.clone() for d in layer.d_flows]
    ref.backward(d_flows)
    g_ref = {n: g.detach().clone() for n, g in ref.store.g.items()}
    frozen = {fm.level: fm.flow_up.clone() for fm in ref.flow_modules if fm.flow_up is not None}
    f_ref = [f.clone() for f in f_ref]
    del ref
    net = build_flow_net(H, W, None, conv_mode="bf16")
    net.store.load_numpy(params, stats)
    net.frozen_warp_flows = frozen
    flows = net(x)
    flow_err = [rel_l2(f, e) for f, e in zip(flows, f_ref)]
    l_bf = layer(x, flows).item()
    net.backward(d_flows)
    rows = {n: rel_l2(g, g_ref[n]) for n, g in net.store.g.items() if float(g_ref[n].norm()) > 0}
    med = float(np.median(list(rows.values())))
    worst = max(rows.items(), key=lambda kv: kv[1])
    record("bf16_vs_fp32_frozen_384x1280", {"loss_fp32": l_ref, "loss_bf16": l_bf, "flow_rel_l2": flow_err,
                                            "grad_rel_l2": rows, "median": med, "worst": list(worst)})
    assert max(flow_err) < 0.05, flow_err
    assert abs(l_bf - l_ref) < 2e-2 * abs(l_ref), (l_bf, l_ref)
    assert worst[1] < BF16_GRAD_REL_L2, worst
    assert med < BF16_GRAD_REL_L2_MEDIAN, med


def test_bf16_and_fp32_modes_train_alike(cuda):
    """Adam steps from the same weights on the same 2 batches in both conv modes.  30 steps are run and recorded; the
    ASSERTED window is the first 8.  Why: the warp backward's scatter atomics make even the fp32 mode non-deterministic
    in the last bit, and this optimisation problem (random init, L1 photometric loss, floor() in the warp) amplifies
    that - on one driver box the fp32 run itself left its own earlier trajectory after ~15 steps (loss on batch A
    rising to 0.159 where another fp32 run of the same build reached 0.145), so a 30-step bound compares two chaotic
    trajectories, not two arithmetic modes.  Within the first 8 steps the modes agree to 1.6 % on every box seen
    (bound: 3 % of the largest loss), and the second visit of each batch is below its first."""
    from optical_flow_b200.train import Trainer
    B, H, W = 2, 192, 640
    params, stats = T.init_params(seed=42)
    batches = [torch.tensor(T.synthetic_pairs(B, H, W, seed=40 + k), device=cuda) for k in range(2)]
    curves = {}
    for mode in ("fp32", "bf16"):
        tr = Trainer(H, W, None, conv_mode=mode, lr=1e-4)
        tr.flow_net.store.load_numpy(params, stats)
        losses = []
        for i in range(30):
            lv, _ = tr.train_step(batches[i % 2], i)
            losses.append(lv.item())
        curves[mode] = losses
        del tr
    a, b = np.array(curves["fp32"]), np.array(curves["bf16"])
    record("loss_curves_30_steps", {"fp32": curves["fp32"], "bf16": curves["bf16"]})
    K = 8
    assert np.abs(a[:K] - b[:K]).max() <= 3e-2 * np.abs(a[:K]).max(), float(np.abs(a[:K] - b[:K]).max() / np.abs(a[:K]).max())
    assert a[2] < a[0] and b[2] < b[0] and a[3] < a[1] and b[3] < b[1]
    assert np.isfinite(a).all() and np.isfinite(b).all()


# ------------------------------------------------------------------ (iv) kernels at the bench layer shapes --
# (N, H, W, [source channels], Cout)
BENCH_LAYERS = [
    (8, 192, 640, (64, 51), 128),     # flow3_conv1: concat [f1 | cost volume + flow], K = 115 * 9
    (8, 192, 640, (128,), 128),       # flow3_conv2
    (8, 192, 640, (128,), 96),        # flow3_conv3
    (8, 192, 640, (96,), 64),         # flow3_conv4
    (8, 192, 640, (64,), 32),         # flow3_conv5
    (8, 24, 80, (256, 49), 128),      # flow0_conv1: K = 305 * 9
]


def _layer_operands(cfg, seed):
    N, H, W, srcs, Cout = cfg
    r = np.random.default_rng(seed)
    cin = sum(srcs)
    x = torch.tensor(r.standard_normal((N, H, W, cin)).astype(np.float32)).bfloat16().float()
    w = torch.tensor((r.standard_normal((3, 3, cin, Cout)) / np.sqrt(9 * cin)).astype(np.float32)).bfloat16().float()
    b = torch.tensor(r.standard_normal((Cout,)).astype(np.float32))
    g = torch.tensor(r.standard_normal((N, H, W, Cout)).astype(np.float32)).bfloat16().float()
    return x, w, b, g


def _sources(cfg, x, cuda):
    """The activation as the kernels see it: one tensor, or two K-sources whose second one is padded to 64 channels."""
    from optical_flow_b200 import conv_tc
    srcs = cfg[3]
    if len(srcs) == 1:
        return x.to(cuda).bfloat16()
    c0, c1 = srcs
    second = torch.zeros(tuple(x.shape[:3]) + (64,), dtype=torch.bfloat16, device=cuda)
    second[..., :c1] = x[..., c0:].to(cuda).bfloat16()
    return conv_tc.MultiSource([conv_tc.Source(x[..., :c0].contiguous().to(cuda).bfloat16(), c0, 0),
                                conv_tc.Source(second, c1, c0)])


@pytest.mark.parametrize("cfg", BENCH_LAYERS, ids=lambda c: "N%d_%dx%d_%s_%d" % (c[0], c[1], c[2], "+".join(map(str, c[3])), c[4]))
def test_tc_conv_at_bench_layer_shape(cuda, cfg):
    from optical_flow_b200 import conv_tc, _lib as L
    torch.set_num_threads(os.cpu_count() or 1)
    N, H, W, srcs, Cout = cfg
    x, w, b, g = _layer_operands(cfg, 17)
    xt, wt, bt = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    z = T.conv2d_same(xt, wt, bt, 1)
    z.backward(g)
    ref_y = torch.nn.functional.leaky_relu(z.detach(), 0.3)
    xs = _sources(cfg, x, cuda)
    wd, bd, gd = w.to(cuda), b.to(cuda), g.to(cuda).bfloat16()
    # fprop (bf16 out, fused bias + LeakyReLU)
    y = conv_tc.conv2d_fprop_tc(xs, wd, bd, act=L.OF_ACT_LEAKY)
    assert rel_err(y, ref_y) <= 2.0 ** -7, rel_err(y, ref_y)
    # dgrad (fp32 out isolates the MMA path)
    if len(srcs) == 1:
        dx = conv_tc.conv2d_dgrad_tc(gd, wd, (H, W), 1, out_dtype=torch.float32)
        assert rel_err(dx, xt.grad) <= 2e-5, rel_err(dx, xt.grad)
    else:
        c0, c1 = srcs
        d0 = conv_tc.conv2d_dgrad_tc(gd, wd, (H, W), 1, seg=(0, c0), out_dtype=torch.float32)
        d1 = conv_tc.conv2d_dgrad_tc(gd, wd, (H, W), 1, seg=(c0, c1), out_dtype=torch.float32, out_channels=64)
        assert rel_err(d0, xt.grad[..., :c0]) <= 2e-5
        assert rel_err(d1[..., :c1], xt.grad[..., c0:]) <= 2e-5
    # wgrad + fused bias gradient (split-K with fp32 atomics over up to 983 040 pixels)
    dw, db = conv_tc.conv2d_wgrad_tc(xs, gd, (3, 3))
    assert rel_err(dw, wt.grad) <= 1e-4, rel_err(dw, wt.grad)
    assert rel_err(db, bt.grad) <= 1e-4, rel_err(db, bt.grad)


@pytest.mark.parametrize("shape", [(8, 192, 640, 64), (8, 24, 80, 256)])
def test_cost_volume_tc_at_bench_shape(cuda, shape):
    """cvtc_fwd / cvtc_bwd (costvol_tc.cu) at decoder level 3 / level 0 of the C2 step vs the oracle
    (model.py:29-42) on bf16-rounded inputs."""
    from optical_flow_b200 import _lib as L
    B, H, W, C = shape
    r = np.random.default_rng(14)
    bf = lambda a: torch.tensor(a).bfloat16()
    f1, f2 = bf(r.standard_normal(shape).astype(np.float32)), bf(r.standard_normal(shape).astype(np.float32))
    g = np.zeros((B, H, W, 64), np.float32)
    g[..., :49] = r.standard_normal((B, H, W, 49))
    g = bf(g)
    torch.set_num_threads(os.cpu_count() or 1)
    e, e1, e2 = (t.numpy() for t in T.create_cost_volume_fwd_bwd(f1.float(), f2.float(), 3, g.float()[..., :49].contiguous()))
    cv = torch.zeros((B, H, W, 64), dtype=torch.bfloat16, device=cuda)
    f1d, f2d, gd = f1.to(cuda), f2.to(cuda), g.to(cuda)
    L.check(L.load().of_cost_volume_fwd(L.ptr(f1d), L.ptr(f2d), L.ptr(cv), B, H, W, C, 3, 64, L.OF_BF16, L.stream_ptr()),
            "of_cost_volume_fwd")
    err = np.abs(cv.float().cpu().numpy()[..., :49] - e)
    assert (err <= 2.0 ** -8 * np.abs(e) + 1e-5 * np.abs(e).max()).all(), err.max()
    d1 = torch.empty(shape, dtype=torch.bfloat16, device=cuda)
    d2 = torch.empty(shape, dtype=torch.bfloat16, device=cuda)
    L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(f1d), L.ptr(f2d), L.ptr(gd), L.ptr(d1), L.ptr(d2), None,
                                             B, H, W, C, 3, 64, L.stream_ptr()), "of_cost_volume_bwd_bf16")
    for got, ex in ((d1, e1), (d2, e2)):
        err = np.abs(got.float().cpu().numpy() - ex)
        assert (err <= 2.0 ** -8 * np.abs(ex) + 1e-5 * np.abs(ex).max()).all(), err.max()


# ------------------------------------------------------------------ (v) configs[4] shape --
def test_fp32_forward_loss_oracle_at_1024x2048(cuda):
    from optical_flow_b200.loss import LossLayer
    from optical_flow_b200.model import build_flow_net
    B, H, W = 1, 1024, 2048
    params, stats = T.init_params(seed=42, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=22)
    ref = T.FlowNetRef(params, stats)
    with torch.no_grad():
        r_flows = ref(torch.tensor(imgs))
        r_loss = T.loss_layer(torch.tensor(imgs), r_flows)
    net = build_flow_net(H, W, None, conv_mode="fp32")
    net.store.load_numpy(params, stats)
    x = torch.tensor(imgs, device=cuda)
    flows = net(x)
    for k, (f, e) in enumerate(zip(flows, r_flows)):
        assert rel_err(f, e) < 1e-4, "flow %d: %g" % (k, rel_err(f, e))
    loss = LossLayer()(x, flows)
    assert abs(loss.item() - float(r_loss)) < 1e-4 * abs(float(r_loss))


def test_bf16_step_runs_at_1024x2048_batch4(cuda):
    """configs[4] (batch 4 per GPU at 1024 x 2048) through the tensor-core path: the step runs, the loss is finite and
    within 2 % of the fp32 mode's on the same batch-1 slice, every gradient tensor is finite and non-zero."""
    from optical_flow_b200.train import Trainer
    H, W = 1024, 2048
    params, stats = T.init_params(seed=42)
    x = torch.tensor(T.synthetic_pairs(4, H, W, seed=23), device=cuda)
    tr = Trainer(H, W, None, conv_mode="bf16")
    tr.flow_net.store.load_numpy(params, stats)
    loss4, flows = tr.forward_backward(x)
    assert np.isfinite(loss4.item())
    for n, g in tr.flow_net.store.g.items():
        assert torch.isfinite(g).all(), n
    assert float(tr.flow_net.store.grad.abs().sum()) > 0
    l_bf = tr.forward_backward(x[:1].contiguous())[0].item()
    del tr
    tr = Trainer(H, W, None, conv_mode="fp32")
    tr.flow_net.store.load_numpy(params, stats)
    l_32 = tr.forward_backward(x[:1].contiguous())[0].item()
    assert abs(l_bf - l_32) < 2e-2 * abs(l_32), (l_bf, l_32)
