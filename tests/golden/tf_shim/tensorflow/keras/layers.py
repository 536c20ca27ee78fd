"""Keras layers used on the hot path, TF defaults restated."""
import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

from .. import _wrap

BATCH = 1                    # static batch size of Input placeholders
REGISTRY = OrderedDict()     # layer name -> layer (creation order)
_counters = {}


def reset(batch=1):
    global BATCH
    BATCH = batch
    REGISTRY.clear()
    _counters.clear()


def _auto_name(base):
    k = _counters.get(base, 0)
    _counters[base] = k + 1
    return base if k == 0 else "%s_%d" % (base, k)


def all_weights(trainable=True):
    out = []
    for name, layer in REGISTRY.items():
        for wname, w, tr in layer.weights():
            if tr == trainable:
                out.append((name + "/" + wname, w))
    return out


def _same_pad(in_size, k, s):
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    return total // 2, total - total // 2


class _Layer:
    base = "layer"

    def __init__(self, name=None):
        self.name = name or _auto_name(self.base)
        assert self.name not in REGISTRY, "duplicate layer name " + self.name
        REGISTRY[self.name] = self

    def weights(self):
        return []


class Conv2D(_Layer):
    base = "conv2d"

    def __init__(self, filters, kernel_size, strides=1, padding="valid", activation=None, use_bias=True, name=None,
                 kernel_regularizer=None, bias_regularizer=None):
        super().__init__(name)
        assert padding == "same" and activation is None
        self.filters, self.k, self.s, self.use_bias = int(filters), int(kernel_size), int(strides), use_bias
        self.kernel = self.bias = None

    def build(self, cin):
        fan_in, fan_out = self.k * self.k * cin, self.k * self.k * self.filters
        lim = math.sqrt(6.0 / (fan_in + fan_out))          # glorot_uniform (Keras default)
        self.kernel = _wrap(torch.empty(self.k, self.k, cin, self.filters).uniform_(-lim, lim)).requires_grad_(True)
        if self.use_bias:
            self.bias = _wrap(torch.zeros(self.filters)).requires_grad_(True)

    def weights(self):
        w = [("kernel", self.kernel, True)]
        if self.use_bias:
            w.append(("bias", self.bias, True))
        return w

    def __call__(self, x):
        if self.kernel is None:
            self.build(int(x.shape[3]))
        pt, pb = _same_pad(int(x.shape[1]), self.k, self.s)
        pl, pr = _same_pad(int(x.shape[2]), self.k, self.s)
        xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
        y = F.conv2d(xn, self.kernel.permute(3, 2, 0, 1), self.bias, stride=self.s)
        return y.permute(0, 2, 3, 1)


class BatchNormalization(_Layer):
    """Called without training=True (train.py:51) -> moving statistics, eps = 1e-3."""
    base = "batch_normalization"
    eps = 1e-3

    def __init__(self, name=None):
        super().__init__(name)
        self.gamma = None

    def weights(self):
        return [("gamma", self.gamma, True), ("beta", self.beta, True),
                ("moving_mean", self.moving_mean, False), ("moving_variance", self.moving_variance, False)]

    def __call__(self, x, training=None):
        assert not training
        if self.gamma is None:
            c = int(x.shape[3])
            self.gamma = _wrap(torch.ones(c)).requires_grad_(True)
            self.beta = _wrap(torch.zeros(c)).requires_grad_(True)
            self.moving_mean = _wrap(torch.zeros(c))
            self.moving_variance = _wrap(torch.ones(c))
        # Keras (fused=False formula): (x - mean) * (gamma * rsqrt(var + eps)) + beta
        inv = self.gamma * torch.rsqrt(self.moving_variance + self.eps)
        return x * inv + (self.beta - self.moving_mean * inv)


class ReLU(_Layer):
    base = "re_lu"

    def __call__(self, x):
        return F.relu(x)


class LeakyReLU(_Layer):
    base = "leaky_re_lu"

    def __init__(self, alpha=0.3, name=None):
        super().__init__(name)
        self.alpha = alpha

    def __call__(self, x):
        return F.leaky_relu(x, self.alpha)


class MaxPool2D(_Layer):
    base = "max_pooling2d"

    def __init__(self, pool_size=2, strides=None, padding="valid", name=None):
        super().__init__(name)
        assert pool_size == 2 and strides is None and padding == "valid"

    def __call__(self, x):
        return F.max_pool2d(x.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
