"""Training-mode BatchNormalization (opt-in): the four CUDA passes against torch-CPU
F.batch_norm(training=True) + autograd (the semantics of Keras' BatchNormalization(training=True): biased batch
variance in the normalisation, momentum 0.99, epsilon 1e-3), and the cross-replica form against the full batch."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def reference(x, gamma, beta, g, act, eps=1e-3, y_kernel=None):
    """fp64 torch batch_norm(training=True) + autograd.  ``y_kernel``: take the activation's derivative mask from the
    CUDA output instead of the fp64 one - an element whose pre-activation is within fp32 rounding of 0 may land on
    either side, and one flipped ReLU mask moves dx by |g gamma / sigma| ~ 1, so over 31 M elements a max-abs bound
    with the fp64 mask is a coin toss (round-1 driver failure); with the kernel's own mask the comparison is of the
    arithmetic, not of a tie-break."""
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    gt = torch.tensor(gamma, dtype=torch.float64, requires_grad=True)
    bt = torch.tensor(beta, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.batch_norm(xt.permute(0, 3, 1, 2), None, None, gt, bt, True, 0.0, eps).permute(0, 2, 3, 1)
    if act != 0:
        neg = 0.0 if act == 1 else 0.3
        pos = (torch.tensor(np.asarray(y_kernel)) > 0) if y_kernel is not None else (y.detach() > 0)
        y = y * torch.where(pos, 1.0, neg).to(torch.float64)
    y.backward(torch.tensor(g, dtype=torch.float64))
    return y.detach().numpy(), xt.grad.numpy(), gt.grad.numpy(), bt.grad.numpy()


def close(a, b, tol):
    a = a.detach().float().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    err = np.abs(a.astype(np.float64) - b).max()
    assert err <= tol * max(np.abs(b).max(), 1e-30), (err, np.abs(b).max())


@pytest.mark.parametrize("shape,act", [((2, 6, 10, 64), 1), ((3, 17, 9, 128), 0), ((1, 5, 7, 256), 2), ((16, 96, 320, 64), 1)])
def test_bn_train_fwd_bwd_fp32(cuda, shape, act):
    from optical_flow_b200.resnet.batchnorm import BatchNormalizationTraining
    r = np.random.default_rng(sum(shape))
    x = (r.standard_normal(shape) * 1.7 + 0.4).astype(np.float32)
    g = r.standard_normal(shape).astype(np.float32)
    C = shape[-1]
    gamma, beta = (1 + 0.2 * r.standard_normal(C)).astype(np.float32), (0.1 * r.standard_normal(C)).astype(np.float32)
    bn = BatchNormalizationTraining(C, act=act, device=cuda)
    bn.gamma.copy_(torch.tensor(gamma)); bn.beta.copy_(torch.tensor(beta))
    y = bn(torch.tensor(x, device=cuda))
    dx = bn.backward(torch.tensor(g, device=cuda))
    ey, edx, eg, eb = reference(x, gamma, beta, g, act, y_kernel=y.float().cpu().numpy())
    close(y, ey, 2e-5)
    close(dx, edx, 1e-4)
    close(bn.d_gamma, eg, 1e-4)
    close(bn.d_beta, eb, 1e-4)
    # moving statistics: one Keras update from (0, 1) with momentum 0.99, unbiased batch variance (fused kernel)
    n = x.size // C
    m, v = x.reshape(-1, C).astype(np.float64).mean(0), x.reshape(-1, C).astype(np.float64).var(0)
    close(bn.moving_mean, 0.01 * m, 1e-4)
    close(bn.moving_variance, 0.99 + 0.01 * v * n / (n - 1), 1e-4)


def test_bn_train_bf16(cuda):
    from optical_flow_b200.resnet.batchnorm import BatchNormalizationTraining
    r = np.random.default_rng(5)
    shape = (2, 12, 20, 64)
    x = torch.tensor(r.standard_normal(shape).astype(np.float32)).bfloat16()
    g = torch.tensor(r.standard_normal(shape).astype(np.float32)).bfloat16()
    bn = BatchNormalizationTraining(64, act=1, device=cuda)
    y = bn(x.to(cuda))
    dx = bn.backward(g.to(cuda))
    ey, edx, eg, eb = reference(x.float().numpy(), np.ones(64, np.float32), np.zeros(64, np.float32), g.float().numpy(), 1,
                                y_kernel=y.float().cpu().numpy())
    close(y, ey, 2 ** -7)
    close(dx, edx, 2 ** -6)
    close(bn.d_gamma, eg, 2e-2)


class _TwoReplicaSum:
    """Stand-in communicator for two replicas living in one process: the 'all-reduce' adds the other replica's
    buffer (captured from a first pass)."""
    world_size = 2

    def __init__(self):
        self.other = []      # buffers of the other replica, in call order
        self.k = 0

    def allreduce_sum_(self, t):
        t.add_(self.other[self.k])
        self.k += 1


def test_bn_train_cross_replica_equals_full_batch(cuda):
    """SURVEY 8c (12) for sync BN: two replicas with half the batch each, statistics and gradient sums exchanged,
    give the full-batch output and gradients."""
    from optical_flow_b200.resnet.batchnorm import BatchNormalizationTraining
    r = np.random.default_rng(9)
    shape = (4, 10, 12, 64)
    x = (r.standard_normal(shape) + 0.3).astype(np.float32)
    g = r.standard_normal(shape).astype(np.float32)
    ey, edx, eg, eb = reference(x, np.ones(64, np.float32), np.zeros(64, np.float32), g, 1)
    halves = [(torch.tensor(x[:2], device=cuda), torch.tensor(g[:2], device=cuda)),
              (torch.tensor(x[2:], device=cuda), torch.tensor(g[2:], device=cuda))]
    # pass 1: each replica alone, to capture the buffers its peer would contribute
    local = []
    for xh, gh in halves:
        bn = BatchNormalizationTraining(64, act=1, device=cuda)
        bn(xh)
        local.append(bn._stats.clone())
    outs = []
    for k, (xh, gh) in enumerate(halves):
        comm = _TwoReplicaSum()
        bn = BatchNormalizationTraining(64, act=1, device=cuda, comm=comm)
        comm.other = [local[1 - k]]
        y = bn(xh)
        outs.append((bn, y, gh, comm))
    # gradient sums of each replica (computed with the GLOBAL statistics), then exchanged
    sums = []
    for bn, y, gh, comm in outs:
        probe = BatchNormalizationTraining(64, act=1, device=cuda)
        probe._stats.copy_(bn._stats); probe._saved = bn._saved
        saved_comm, probe.comm = probe.comm, None
        probe.backward(gh)
        sums.append((probe.d_beta.clone(), probe.d_gamma.clone()))
    for k, (bn, y, gh, comm) in enumerate(outs):
        comm.other = [local[1 - k], sums[1 - k][0], sums[1 - k][1]]
        comm.k = 1
        dx = bn.backward(gh)
        close(y, ey[2 * k:2 * k + 2], 2e-5)
        close(dx, edx[2 * k:2 * k + 2], 1e-4)
        close(bn.d_gamma, eg, 1e-4)
        close(bn.d_beta, eb, 1e-4)


def test_bn_train_is_deterministic(cuda):
    """Two-stage fixed-order reductions (no floating-point atomics): repeated runs are bit-identical."""
    from optical_flow_b200.resnet.batchnorm import BatchNormalizationTraining
    r = np.random.default_rng(21)
    shape = (8, 96, 160, 64)
    x = torch.tensor((r.standard_normal(shape) * 1.3 + 0.2).astype(np.float32), device=cuda)
    g = torch.tensor(r.standard_normal(shape).astype(np.float32), device=cuda)
    runs = []
    for _ in range(3):
        bn = BatchNormalizationTraining(64, act=1, device=cuda)
        y = bn(x)
        dx = bn.backward(g)
        runs.append((bn._stats.clone(), y.clone(), dx.clone(), bn.d_gamma.clone(), bn.d_beta.clone()))
    for other in runs[1:]:
        for a, b in zip(runs[0], other):
            assert torch.equal(a, b)


def test_bn_train_shifted_variance(cuda):
    """Large mean / small spread (mean^2 / var = 4e4): with the sums taken about the moving mean the variance keeps
    fp32 accuracy; the plain E[x^2] - E[x]^2 form (shift 0) loses ~4 digits here."""
    from optical_flow_b200.resnet.batchnorm import BatchNormalizationTraining
    r = np.random.default_rng(33)
    shape = (4, 40, 64, 64)
    x = (r.standard_normal(shape) * 0.5 + 100.0).astype(np.float32)
    g = r.standard_normal(shape).astype(np.float32)
    bn = BatchNormalizationTraining(64, act=0, device=cuda)
    bn.moving_mean.fill_(99.7)                       # a moving mean that tracks the batch mean
    y = bn(torch.tensor(x, device=cuda), update_moving=False)
    dx = bn.backward(torch.tensor(g, device=cuda))
    ey, edx, eg, eb = reference(x, np.ones(64, np.float32), np.zeros(64, np.float32), g, 0)
    close(y, ey, 1e-4)         # x itself carries 100 * 2^-24 = 6e-6 of rounding on a sigma of 0.5
    close(dx, edx, 2e-4)


def _train_bn_ref():
    from oracle import ref_torch as T

    class TrainBnRef(T.FlowNetRef):
        def _bn(self, x, name):
            y = torch.nn.functional.batch_norm(x.permute(0, 3, 1, 2), None, None, self.p[name + "/gamma"],
                                               self.p[name + "/beta"], True, 0.0, 1e-3)
            return y.permute(0, 2, 3, 1)

        def __call__(self, batch_imgs):         # FlowNet runs the shared encoder ONCE on the 2B stacked frames
            n = batch_imgs.shape[0]
            e = self.encoder(torch.cat([batch_imgs[..., :3], batch_imgs[..., 3:]], dim=0))
            e1, e2 = [t[:n] for t in e], [t[n:] for t in e]
            flow0 = self.flow_module(e1[-1], e2[-1], None, 0)
            flow1 = self.flow_module(e1[-2], e2[-2], flow0, 1)
            flow2 = self.flow_module(e1[-3], e2[-3], flow1, 2)
            flow3 = self.flow_module(e1[-4], e2[-4], flow2, 3)
            return [flow3, flow2, flow1, flow0]

    return T, TrainBnRef


@pytest.mark.parametrize("mode,tol_y,tol_g", [("fp32", 1e-4, 2e-3), ("bf16", 0.03, 0.25)])
def test_flownet_bn_train_mode_matches_torch(cuda, mode, tol_y, tol_g):
    """FlowNet(bn_mode="train"): every BatchNormalization of the encoder (model.py:14; resnet/models.py:19,24) on batch
    statistics, against the torch oracle with F.batch_norm(training=True): the four encoder outputs, and - for fixed
    upstream gradients - the gradient of every encoder parameter (relative L2 per tensor; bf16 tolerance stated).
    The comparison stops at the encoder on purpose: behind it the warp's floor() and the L1 loss's sign() turn a
    last-bit difference of the batch statistics (a different summation order is enough - the CPU oracle itself moves
    by 3e-4 in the loss and 60 % in some decoder gradients between a one-pass and a two-pass variance at this size)
    into index / sign flips, so an end-to-end bound measures the flips, not the BN arithmetic."""
    T, TrainBnRef = _train_bn_ref()
    from optical_flow_b200.model import FlowNet
    B, H, W = 2, 64, 96
    params, stats = T.init_params(seed=42, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=11)
    r = np.random.default_rng(5)
    ref = TrainBnRef(params, stats)
    for t in ref.p.values():
        t.grad = None
    x2 = torch.tensor(np.concatenate([imgs[..., :3], imgs[..., 3:]], axis=0))
    r_outs = ref.encoder(x2)
    d_np = [r.standard_normal(tuple(o.shape)).astype(np.float32) for o in r_outs]
    sum((o * torch.tensor(d)).sum() for o, d in zip(r_outs, d_np)).backward()

    net = FlowNet(H, W, None, conv_mode=mode, bn_mode="train")
    net.store.load_numpy(params, stats)
    mm0 = net.store.stats["layer1_bn/moving_mean"].clone()
    net.B = B
    net.prepare_weights()
    outs = net.encoder(net.backend_split(torch.tensor(imgs, device=cuda)))
    for k, (o, e) in enumerate(zip(outs, r_outs)):
        e = e.detach().numpy().astype(np.float64)
        rel = float(np.linalg.norm(o.float().cpu().numpy() - e) / np.linalg.norm(e))
        assert rel < tol_y, (k, rel)
    net.store.grad.zero_()
    net.encoder.backward([torch.tensor(d, device=cuda).to(net.backend.act_dtype) for d in d_np])
    worst = ("", 0.0)
    enc_names = [n for n in net.store.specs if not n.startswith("flow")]
    biggest = max(float(ref.p[n].grad.norm()) for n in enc_names)
    for name in enc_names:
        e = ref.p[name].grad.numpy().astype(np.float64).ravel()
        got = net.store.g[name].detach().cpu().numpy().astype(np.float64).ravel()
        if np.linalg.norm(e) < 1e-5 * biggest:
            # a conv bias in front of a batch-statistics BN has an exactly-zero gradient (the mean is subtracted):
            # both sides hold rounding noise there (bf16: the summed dz carries 2^-9 relative rounding per element)
            assert np.linalg.norm(got) < max(1e-3, 0.1 * tol_g) * biggest, name
            continue
        rel = float(np.linalg.norm(got - e) / np.linalg.norm(e))
        if rel > worst[1]:
            worst = (name, rel)
    assert worst[1] < tol_g, worst
    assert not torch.equal(net.store.stats["layer1_bn/moving_mean"], mm0)      # moving statistics were updated


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_trainer_bn_train_mode_step_runs(cuda, mode):
    """Trainer(bn_mode="train") end to end (forward, loss, backward through the decoder, warp and BN layers): the loss
    agrees with the torch oracle within 1 % (see the note above on why not tighter) and every gradient is finite."""
    T, TrainBnRef = _train_bn_ref()
    from optical_flow_b200.train import Trainer
    B, H, W = 2, 64, 96
    params, stats = T.init_params(seed=42, randomize_bn=True)
    imgs = T.synthetic_pairs(B, H, W, seed=11)
    r_loss, _, _ = TrainBnRef(params, stats).loss_and_grads(torch.tensor(imgs))
    tr = Trainer(H, W, None, conv_mode=mode, bn_mode="train")
    tr.flow_net.store.load_numpy(params, stats)
    loss, flows = tr.forward_backward(torch.tensor(imgs, device=cuda))
    assert abs(loss.item() - r_loss.item()) < (1e-2 if mode == "fp32" else 5e-2) * abs(r_loss.item())
    assert bool(torch.isfinite(tr.flow_net.store.grad).all())
