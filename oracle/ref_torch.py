"""Independent torch-CPU autograd restatement of the reference hot path.

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py (pinned to tests/golden/*.npz).

Second implementation for the two-implementation rule: every op
is written with differentiable torch primitives so that autograd derives the
same gradients TF's GradientTape would (floor / int casts / clips carry no
gradient; gather -> scatter-add).  It is also the CPU baseline timed by
bench.py (all host threads).  Layout NHWC like the reference; convolutions
permute to NCHW internally and use explicit TF-'SAME' padding.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F


# ---------------------------------------------------------------- sampler --

def bilinear_interpolation(inp, pts):
    """transformations.py:85-129."""
    B, H, W, C = inp.shape
    x = pts[..., 0]
    y = pts[..., 1]
    x0 = torch.floor(x).to(torch.int64)
    y0 = torch.floor(y).to(torch.int64)
    x1 = x0 + 1
    y1 = y0 + 1
    x0 = x0.clamp(0, W - 1)
    x1 = x1.clamp(0, W - 1)
    y0 = y0.clamp(0, H - 1)
    y1 = y1.clamp(0, H - 1)
    flat = inp.reshape(B, H * W, C)

    def gather(xx, yy):
        idx = (yy * W + xx).reshape(B, H * W, 1).expand(B, H * W, C)
        return torch.gather(flat, 1, idx).reshape(B, H, W, C)

    v00, v01, v10, v11 = gather(x0, y0), gather(x0, y1), gather(x1, y0), gather(x1, y1)
    a = (x1.to(inp.dtype) - x)
    b = (y1.to(inp.dtype) - y)
    w00 = (a * b).unsqueeze(-1)
    w01 = (a * (1.0 - b)).unsqueeze(-1)
    w10 = ((1.0 - a) * b).unsqueeze(-1)
    w11 = ((1.0 - a) * (1.0 - b)).unsqueeze(-1)
    return w00 * v00 + w01 * v01 + w10 * v10 + w11 * v11


def warp_features(flow, feats, mode="ref_ij"):
    """model.py:55-73 (mode 'ref_ij', the reference's behaviour) or the
    corrected x/y order (mode 'xy')."""
    B, H, W, C = feats.shape
    ii, jj = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    if mode == "ref_ij":
        grid = torch.stack([ii, jj], dim=-1)
    else:
        grid = torch.stack([jj, ii], dim=-1)
    pts = grid.to(flow.dtype).unsqueeze(0) + flow
    return bilinear_interpolation(feats, pts)


def image_warp(im, flow):
    """old/image_warp.py:7-78."""
    B, H, W, C = im.shape
    ff = torch.floor(flow)
    fl = ff.to(torch.int64)
    wts = flow - ff.detach()
    pos_x = torch.arange(W).reshape(1, 1, W)
    pos_y = torch.arange(H).reshape(1, H, 1)
    x0 = pos_x + fl[..., 0]
    y0 = pos_y + fl[..., 1]
    x1 = x0 + 1
    y1 = y0 + 1
    x0, x1 = x0.clamp(0, W - 1), x1.clamp(0, W - 1)
    y0, y1 = y0.clamp(0, H - 1), y1.clamp(0, H - 1)
    flat = im.reshape(B, H * W, C)

    def gather(xx, yy):
        idx = (yy * W + xx).reshape(B, H * W, 1).expand(B, H * W, C)
        return torch.gather(flat, 1, idx).reshape(B, H, W, C)

    xw, yw = wts[..., 0], wts[..., 1]
    wa = ((1 - xw) * (1 - yw)).unsqueeze(-1)
    wb = ((1 - xw) * yw).unsqueeze(-1)
    wc = (xw * (1 - yw)).unsqueeze(-1)
    wd = (xw * yw).unsqueeze(-1)
    return wa * gather(x0, y0) + wb * gather(x0, y1) + wc * gather(x1, y0) + wd * gather(x1, y1)


# ----------------------------------------------------------------- resize --

def resize_bilinear(x, oh, ow):
    """tf.image.resize default == F.interpolate(bilinear, align_corners=False,
    antialias=False) (checked in SURVEY F7)."""
    if x.shape[1] == oh and x.shape[2] == ow:
        return x
    y = F.interpolate(x.permute(0, 3, 1, 2), size=(oh, ow), mode="bilinear",
                      align_corners=False, antialias=False)
    return y.permute(0, 2, 3, 1)


def upscale_flow(flow):
    """model.py:76-77."""
    return resize_bilinear(flow, flow.shape[1] * 2, flow.shape[2] * 2) * 2.0


def create_cost_volume(f1, f2, max_disp):
    """model.py:29-42."""
    B, H, W, C = f1.shape
    m = max_disp
    f2p = F.pad(f2, (0, 0, m, m, m, m))
    corr = []
    for i in range(2 * m + 1):
        for j in range(2 * m + 1):
            corr.append((f1 * f2p[:, i:i + H, j:j + W, :]).sum(dim=-1))
    return torch.stack(corr, dim=-1)


def create_cost_volume_fwd_bwd(f1, f2, max_disp, d_cv):
    """model.py:29-42 forward and its hand-derived gradient in one multi-threaded torch loop that keeps no
    autograd tape (49 saved products of the full feature map are 12 GB at 8 x 192 x 640 x 64): the large-shape
    oracle of tests/test_b4_bench_shapes.py.  Agrees with create_cost_volume + autograd and with
    ref_np_ops.create_cost_volume_bwd (tests/test_a0_oracle.py)."""
    B, H, W, C = f1.shape
    m = max_disp
    f2p = F.pad(f2, (0, 0, m, m, m, m))
    cv = torch.empty((B, H, W, (2 * m + 1) ** 2), dtype=f1.dtype)
    d_f1 = torch.zeros_like(f1)
    d_f2p = torch.zeros_like(f2p)
    k = 0
    with torch.no_grad():
        for i in range(2 * m + 1):
            for j in range(2 * m + 1):
                win = f2p[:, i:i + H, j:j + W, :]
                cv[..., k] = (f1 * win).sum(dim=-1)
                gk = d_cv[..., k:k + 1]
                d_f1.addcmul_(gk, win)
                d_f2p[:, i:i + H, j:j + W, :].addcmul_(gk, f1)
                k += 1
    return cv, d_f1, d_f2p[:, m:m + H, m:m + W, :].contiguous()


# ----------------------------------------------------------------- losses --

def loss_layer(batch_imgs, flows):
    """loss.py:5-32."""
    num_scales = len(flows)
    loss = torch.zeros((), dtype=batch_imgs.dtype)
    for s in range(num_scales):
        sh = int(batch_imgs.shape[1] / (2.0 ** (s + 1)))
        sw = int(batch_imgs.shape[2] / (2.0 ** (s + 1)))
        flow = flows[s]
        assert flow.shape[1] == sh and flow.shape[2] == sw
        resized = resize_bilinear(batch_imgs, sh, sw)
        warped = warp_features(flow, resized[..., 3:])
        loss = loss + torch.mean(torch.abs(resized[..., :3] - warped))
    return loss / float(num_scales)


def my_loss(y_true, flows_coarse_to_fine, motions=None):
    """old/loss.py:7-105."""
    im1 = y_true[..., :3]
    im2 = y_true[..., 3:]
    B, H, W, _ = im1.shape
    photo = 0.0
    smooth = 0.0
    for k, flow in enumerate(flows_coarse_to_fine):
        div = 16 >> k
        h, w = H // div, W // div
        a = resize_bilinear(im1, h, w)
        b = resize_bilinear(im2, h, w)
        diff = image_warp(a, flow) - b
        photo = photo + torch.sqrt((diff ** 2).sum(dim=-1) + 1e-4).sum()
        di = flow[:, :-1] - flow[:, 1:]
        dj = flow[:, :, :-1] - flow[:, :, 1:]
        smooth = smooth + torch.sqrt(di ** 2 + 1e-4).sum() + torch.sqrt(dj ** 2 + 1e-4).sum()
    motion = 0.0
    if motions is not None:
        for mo in motions:
            motion = motion + torch.sqrt((mo ** 2).sum(dim=-1)).sum()
    return (photo + 0.2 * smooth + motion) / B


# ------------------------------------------------------------ convolution --

def tf_same_pad(in_size, k, s):
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    return out, total // 2, total - total // 2


def conv2d_same(x, w_hwio, bias, stride):
    """Keras Conv2D(padding='same'): NHWC input, HWIO kernel, asymmetric pad."""
    kh, kw = w_hwio.shape[0], w_hwio.shape[1]
    _, pt, pb = tf_same_pad(x.shape[1], kh, stride)
    _, pl, pr = tf_same_pad(x.shape[2], kw, stride)
    xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    y = F.conv2d(xn, w_hwio.permute(3, 2, 0, 1), bias, stride=stride)
    return y.permute(0, 2, 3, 1)


BN_EPS = 1e-3  # Keras BatchNormalization default epsilon


def bn_inference(x, gamma, beta, mean, var):
    """BatchNormalization called with training=None -> moving statistics (F5)."""
    return gamma * (x - mean) / torch.sqrt(var + BN_EPS) + beta


# ---------------------------------------------------- parameter inventory --

def encoder_param_specs():
    """Names/shapes in the reference's layer naming (model.py:12-24,
    resnet/models.py:15-30).  BN moving stats are non-trainable."""
    specs = OrderedDict()
    specs["conv1/kernel"] = (7, 7, 3, 64)
    specs["conv1/bias"] = (64,)
    specs["layer1_bn/gamma"] = (64,)
    specs["layer1_bn/beta"] = (64,)
    cin = 64
    for layer_idx, down in ((2, False), (3, True), (4, True)):
        for blk in (1, 2):
            ds = down and blk == 1
            cout = 2 * cin if ds else cin
            base = "layer%d_block_%d" % (layer_idx, blk)
            specs[base + "_conv1/kernel"] = (3, 3, cin, cout)
            specs[base + "_conv1/bias"] = (cout,)
            specs[base + "_bn1/gamma"] = (cout,)
            specs[base + "_bn1/beta"] = (cout,)
            specs[base + "_conv2/kernel"] = (3, 3, cout, cout)
            specs[base + "_conv2/bias"] = (cout,)
            specs[base + "_bn2/gamma"] = (cout,)
            specs[base + "_bn2/beta"] = (cout,)
            if ds:
                specs[base + "_shortcut/kernel"] = (1, 1, cin, cout)
            cin = cout
    return specs


def decoder_param_specs(max_disp=3):
    """flow_module convs in creation order (model.py:104-114); level 0 is the
    coarsest (H/16, 256 ch, no previous flow)."""
    specs = OrderedDict()
    ncv = (2 * max_disp + 1) ** 2
    feat_ch = [256, 128, 64, 64]
    for lvl in range(4):
        cin = feat_ch[lvl] + ncv + (2 if lvl > 0 else 0)
        for ci, cout in enumerate((128, 128, 96, 64, 32, 2)):
            specs["flow%d_conv%d/kernel" % (lvl, ci + 1)] = (3, 3, cin, cout)
            specs["flow%d_conv%d/bias" % (lvl, ci + 1)] = (cout,)
            cin = cout
    return specs


def bn_stat_names():
    names = ["layer1_bn"]
    for layer_idx in (2, 3, 4):
        for blk in (1, 2):
            names += ["layer%d_block_%d_bn1" % (layer_idx, blk), "layer%d_block_%d_bn2" % (layer_idx, blk)]
    return names


def init_params(seed=42, max_disp=3, randomize_bn=False):
    """Glorot-uniform kernels (Keras default), zero bias, gamma=1, beta=0,
    moving mean 0 / var 1.  randomize_bn=True perturbs BN params/statistics
    and biases so tests exercise every term."""
    rng = np.random.default_rng(seed)
    params = OrderedDict()
    specs = OrderedDict()
    specs.update(encoder_param_specs())
    specs.update(decoder_param_specs(max_disp))
    for name, shape in specs.items():
        if name.endswith("/kernel"):
            fan_in = shape[0] * shape[1] * shape[2]
            fan_out = shape[0] * shape[1] * shape[3]
            lim = math.sqrt(6.0 / (fan_in + fan_out))
            params[name] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
        elif name.endswith("/gamma"):
            params[name] = (np.ones(shape) + (0.2 * rng.standard_normal(shape) if randomize_bn else 0.0)).astype(np.float32)
        else:  # bias / beta
            params[name] = (0.05 * rng.standard_normal(shape) if randomize_bn else np.zeros(shape)).astype(np.float32)
    stats = OrderedDict()
    for bn in bn_stat_names():
        c = params[bn + "/gamma"].shape[0]
        stats[bn + "/moving_mean"] = (0.1 * rng.standard_normal(c) if randomize_bn else np.zeros(c)).astype(np.float32)
        stats[bn + "/moving_variance"] = (1.0 + 0.3 * rng.random(c) if randomize_bn else np.ones(c)).astype(np.float32)
    return params, stats


# ---------------------------------------------------------------- network --

class FlowNetRef:
    """model.py:119-143 build_flow_net, with torch tensors as parameters."""

    def __init__(self, params, stats, max_disp=3, dtype=torch.float32):
        self.p = OrderedDict((k, torch.tensor(v, dtype=dtype, requires_grad=True)) for k, v in params.items())
        self.s = OrderedDict((k, torch.tensor(v, dtype=dtype)) for k, v in stats.items())
        self.max_disp = max_disp

    def _bn(self, x, name):
        return bn_inference(x, self.p[name + "/gamma"], self.p[name + "/beta"],
                            self.s[name + "/moving_mean"], self.s[name + "/moving_variance"])

    def _block(self, x, down, layer_idx, blk):
        """resnet/models.py:8-36 (ReLU before the residual add)."""
        base = "layer%d_block_%d" % (layer_idx, blk)
        s = 2 if down else 1
        y = conv2d_same(x, self.p[base + "_conv1/kernel"], self.p[base + "_conv1/bias"], s)
        y = F.relu(self._bn(y, base + "_bn1"))
        y = conv2d_same(y, self.p[base + "_conv2/kernel"], self.p[base + "_conv2/bias"], 1)
        y = F.relu(self._bn(y, base + "_bn2"))
        if down:
            sc = conv2d_same(x, self.p[base + "_shortcut/kernel"], None, 2)
            y = y + sc
        else:
            y = y + x
        return F.relu(y)

    def encoder(self, img):
        """model.py:10-26."""
        x = conv2d_same(img, self.p["conv1/kernel"], self.p["conv1/bias"], 2)
        x = F.relu(self._bn(x, "layer1_bn"))
        outs = [x]
        x = F.max_pool2d(x.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)  # MaxPool2D(): 2x2/2 valid
        for layer_idx, down in ((2, False), (3, True), (4, True)):
            for blk in (1, 2):
                x = self._block(x, down and blk == 1, layer_idx, blk)
            outs.append(x)
        return outs

    def flow_module(self, f1, f2, prev_flow, lvl):
        """model.py:80-116."""
        if prev_flow is not None:
            flow_up = upscale_flow(prev_flow)
            f2w = warp_features(flow_up, f2)
        else:
            f2w = f2
        cv = create_cost_volume(f1, f2w, self.max_disp)
        x = torch.cat([f1, cv, flow_up], dim=-1) if prev_flow is not None else torch.cat([f1, cv], dim=-1)
        for ci in range(1, 7):
            x = conv2d_same(x, self.p["flow%d_conv%d/kernel" % (lvl, ci)], self.p["flow%d_conv%d/bias" % (lvl, ci)], 1)
            if ci < 6:
                x = F.leaky_relu(x, 0.3)   # Keras LeakyReLU() default alpha
        return x

    def __call__(self, batch_imgs):
        e1 = self.encoder(batch_imgs[..., :3])
        e2 = self.encoder(batch_imgs[..., 3:])
        flow0 = self.flow_module(e1[-1], e2[-1], None, 0)
        flow1 = self.flow_module(e1[-2], e2[-2], flow0, 1)
        flow2 = self.flow_module(e1[-3], e2[-3], flow1, 2)
        flow3 = self.flow_module(e1[-4], e2[-4], flow2, 3)
        return [flow3, flow2, flow1, flow0]

    def loss_and_grads(self, batch_imgs):
        """train.py:50-55: loss + gradients w.r.t. trainable weights."""
        for t in self.p.values():
            t.grad = None
        flows = self(batch_imgs)
        loss = loss_layer(batch_imgs, flows)
        loss.backward()
        grads = OrderedDict((k, (v.grad if v.grad is not None else torch.zeros_like(v)).detach().clone())
                            for k, v in self.p.items())
        return loss.detach(), [f.detach() for f in flows], grads


def adam_step(param, grad, m, v, step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-7):
    lr_t = lr * math.sqrt(1.0 - beta2 ** step) / (1.0 - beta1 ** step)
    # TF's ApplyAdam forms (1 - beta) in the variable dtype (fp32): T(1) - T(beta)
    omb1 = float(np.float32(1.0) - np.float32(beta1))
    omb2 = float(np.float32(1.0) - np.float32(beta2))
    m = m + (grad - m) * omb1
    v = v + (grad * grad - v) * omb2
    param = param - lr_t * m / (torch.sqrt(v) + eps)
    return param, m, v


# ----------------------------------------------------------- synthetic data --

def synthetic_pairs(batch, height, width, seed=1234):
    """SURVEY 8d synthetic inputs: smooth random texture (8 random 2-D
    sinusoids + noise) clipped to [0,1] minus the BGR mean/255
    (data_reader.py:58-63); frame 2 = frame 1 translated by U(-8,8) px + noise.
    Returns (B,H,W,6) float32."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(height, dtype=np.float32), np.arange(width, dtype=np.float32), indexing="ij")
    means = np.array([104.0, 117.0, 123.0], np.float32) / 255.0
    out = np.zeros((batch, height, width, 6), np.float32)
    for b in range(batch):
        dx, dy = rng.uniform(-8, 8, size=2)
        for c in range(3):
            im1 = np.zeros((height, width), np.float32)
            im2 = np.zeros((height, width), np.float32)
            for _ in range(8):
                fx, fy = rng.uniform(0.005, 0.15, size=2)
                ph = rng.uniform(0, 2 * np.pi)
                amp = rng.uniform(0.03, 0.12)
                im1 += amp * np.sin(2 * np.pi * (fx * xx + fy * yy) + ph)
                im2 += amp * np.sin(2 * np.pi * (fx * (xx + dx) + fy * (yy + dy)) + ph)
            im1 = im1 + 0.5 + rng.normal(0, 0.02, size=im1.shape)
            im2 = im2 + 0.5 + rng.normal(0, 0.01, size=im2.shape)
            out[b, :, :, c] = np.clip(im1, 0, 1) - means[c]
            out[b, :, :, 3 + c] = np.clip(im2, 0, 1) - means[c]
    return out
