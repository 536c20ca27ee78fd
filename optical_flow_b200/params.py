"""Flat fp32 parameter / gradient / Adam-state buffers with a name -> slice table
that uses the reference's layer names (model.py:12-24, resnet/models.py:15-30,
model.py:104-114 in creation order).  One contiguous buffer is what the Adam
kernel and the NCCL gradient all-reduce operate on.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch

BN_EPS = 1e-3   # Keras BatchNormalization default


def encoder_param_specs():
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


DECODER_WIDTHS = (128, 128, 96, 64, 32, 2)   # model.py:104-114
FEATURE_CHANNELS = (256, 128, 64, 64)        # encoder outputs, coarsest first (model.py:134-137)


def decoder_param_specs(max_disp=3):
    specs = OrderedDict()
    ncv = (2 * max_disp + 1) ** 2
    for lvl in range(4):
        cin = FEATURE_CHANNELS[lvl] + ncv + (2 if lvl > 0 else 0)
        for ci, cout in enumerate(DECODER_WIDTHS):
            specs["flow%d_conv%d/kernel" % (lvl, ci + 1)] = (3, 3, cin, cout)
            specs["flow%d_conv%d/bias" % (lvl, ci + 1)] = (cout,)
            cin = cout
    return specs


def bn_names():
    names = ["layer1_bn"]
    for layer_idx in (2, 3, 4):
        for blk in (1, 2):
            names += ["layer%d_block_%d_bn1" % (layer_idx, blk), "layer%d_block_%d_bn2" % (layer_idx, blk)]
    return names


class ParamStore:
    """params / grads / adam m,v as single flat fp32 CUDA tensors + named views."""

    def __init__(self, device, max_disp=3):
        self.device = device
        self.specs = OrderedDict()
        self.specs.update(encoder_param_specs())
        self.specs.update(decoder_param_specs(max_disp))
        self.offsets = OrderedDict()
        off = 0
        for name, shape in self.specs.items():
            n = int(np.prod(shape))
            self.offsets[name] = (off, n)
            off += (n + 3) // 4 * 4          # keep every slice 16-byte aligned
        self.numel = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.grad = torch.zeros(off, dtype=torch.float32, device=device)
        self.adam_m = torch.zeros(off, dtype=torch.float32, device=device)
        self.adam_v = torch.zeros(off, dtype=torch.float32, device=device)
        self.p = OrderedDict()
        self.g = OrderedDict()
        for name, shape in self.specs.items():
            o, n = self.offsets[name]
            self.p[name] = self.flat[o:o + n].view(shape)
            self.g[name] = self.grad[o:o + n].view(shape)
        # non-trainable BN moving statistics
        self.stats = OrderedDict()
        for bn in bn_names():
            c = self.specs[bn + "/gamma"][0]
            self.stats[bn + "/moving_mean"] = torch.zeros(c, dtype=torch.float32, device=device)
            self.stats[bn + "/moving_variance"] = torch.ones(c, dtype=torch.float32, device=device)

    def num_trainable(self):
        return sum(n for _, n in self.offsets.values())

    def init_glorot(self, seed=42):
        """Keras defaults: Glorot-uniform kernels, zero bias, gamma 1, beta 0, mean 0, var 1.
        Generated on the host with numpy so every rank (and the oracle) gets identical values."""
        rng = np.random.default_rng(seed)
        for name, shape in self.specs.items():
            if name.endswith("/kernel"):
                fan_in = shape[0] * shape[1] * shape[2]
                fan_out = shape[0] * shape[1] * shape[3]
                lim = math.sqrt(6.0 / (fan_in + fan_out))
                val = rng.uniform(-lim, lim, size=shape).astype(np.float32)
            elif name.endswith("/gamma"):
                val = np.ones(shape, np.float32)
            else:
                val = np.zeros(shape, np.float32)
            self.p[name].copy_(torch.from_numpy(val))

    def load_numpy(self, params, stats=None):
        for name in self.specs:
            self.p[name].copy_(torch.from_numpy(np.ascontiguousarray(params[name], dtype=np.float32)))
        if stats is not None:
            for name in self.stats:
                self.stats[name].copy_(torch.from_numpy(np.ascontiguousarray(stats[name], dtype=np.float32)))

    def to_numpy(self):
        out = OrderedDict((k, v.detach().cpu().numpy().copy()) for k, v in self.p.items())
        out.update((k, v.detach().cpu().numpy().copy()) for k, v in self.stats.items())
        return out

    def grads_numpy(self):
        return OrderedDict((k, v.detach().cpu().numpy().copy()) for k, v in self.g.items())
