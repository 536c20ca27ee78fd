"""Training-mode BatchNormalization (opt-in building block).

The reference's train step calls its Keras model as ``flow_net(batch_imgs)`` (train.py:51), so every
``layers.BatchNormalization`` (model.py:14; resnet/models.py:19,24) runs in inference mode - that is the
path ``FlowNet`` fuses into the convolutions.  This class is the batch-statistics variant with Keras' defaults
(momentum 0.99, epsilon 1e-3), an optional fused activation, explicit ``backward`` and optional cross-replica
statistics (``comm``: anything with ``world_size`` and ``allreduce_sum_(tensor)``)."""
from __future__ import annotations

import torch

from .. import _lib as L


class BatchNormalizationTraining:
    def __init__(self, channels, momentum=0.99, epsilon=1e-3, act=L.OF_ACT_NONE, device=None, comm=None, tensors=None):
        """``tensors``: optional dict of existing f32 (C,) device tensors to use in place (FlowNet passes views of its
        flat parameter / gradient / statistics buffers): gamma, beta, moving_mean, moving_variance, d_gamma, d_beta."""
        dev = device or torch.device("cuda", torch.cuda.current_device())
        self.C, self.momentum, self.epsilon, self.act, self.comm = int(channels), float(momentum), float(epsilon), act, comm
        t = tensors or {}
        self.gamma = t.get("gamma", None) if "gamma" in t else torch.ones(channels, device=dev)
        self.beta = t["beta"] if "beta" in t else torch.zeros(channels, device=dev)
        self.moving_mean = t["moving_mean"] if "moving_mean" in t else torch.zeros(channels, device=dev)
        self.moving_variance = t["moving_variance"] if "moving_variance" in t else torch.ones(channels, device=dev)
        self.d_gamma = t["d_gamma"] if "d_gamma" in t else torch.zeros(channels, device=dev)
        self.d_beta = t["d_beta"] if "d_beta" in t else torch.zeros(channels, device=dev)
        for v in (self.gamma, self.beta, self.moving_mean, self.moving_variance, self.d_gamma, self.d_beta):
            assert v.dtype == torch.float32 and v.numel() == channels and v.is_contiguous()
        self._stats = torch.zeros((2, channels), device=dev)      # sum, sum of squares (about _shift)
        self._shift = torch.zeros(channels, device=dev)           # snapshot of the moving mean taken by __call__
        self._ws = torch.empty(L.load().of_bn_train_workspace_bytes(int(channels)), dtype=torch.uint8, device=dev)
        self._saved = None

    def _count(self, npix):
        return float(npix * (self.comm.world_size if self.comm is not None else 1))

    def __call__(self, x, update_moving=True):
        """x (N,H,W,C) float32 / bfloat16 -> y of the same type; keeps what ``backward`` needs."""
        assert x.shape[-1] == self.C and x.is_contiguous()
        npix = x.numel() // self.C
        lib = L.load()
        # sums are taken about the moving mean (identical on every replica -> the buffers stay all-reducible);
        # a snapshot, because apply updates the moving mean and backward needs the value the sums were taken about
        self._shift.copy_(self.moving_mean)
        L.check(lib.of_bn_train_stats(L.ptr(x), L.ptr(self._shift), L.ptr(self._stats[0]), L.ptr(self._stats[1]), npix, self.C,
                                      L.dtype_code(x), L.ptr(self._ws), L.stream_ptr()), "of_bn_train_stats")
        if self.comm is not None:
            self.comm.allreduce_sum_(self._stats)
        count = self._count(npix)
        y = torch.empty_like(x)
        L.check(lib.of_bn_train_apply(L.ptr(x), L.ptr(self._stats[0]), L.ptr(self._stats[1]), L.ptr(self._shift), count, L.ptr(self.gamma),
                                      L.ptr(self.beta), self.epsilon, self.act, L.ptr(y),
                                      L.ptr(self.moving_mean) if update_moving else None,
                                      L.ptr(self.moving_variance) if update_moving else None, self.momentum, npix, self.C,
                                      L.dtype_code(x), L.stream_ptr()), "of_bn_train_apply")
        self._saved = (x, y, count)
        return y

    def backward(self, dy, act_applied=False):
        """dy: gradient w.r.t. the (activated) output.  Returns dx; fills ``d_gamma`` / ``d_beta``.
        ``act_applied``: dy was already multiplied by the activation's derivative (a fused upstream epilogue did it)."""
        x, y, count = self._saved
        act = L.OF_ACT_NONE if act_applied else self.act
        assert dy.shape == x.shape and dy.dtype == x.dtype and dy.is_contiguous()
        npix = x.numel() // self.C
        lib = L.load()
        L.check(lib.of_bn_train_bwd_reduce(L.ptr(x), L.ptr(y), L.ptr(dy), L.ptr(self._stats[0]), L.ptr(self._stats[1]),
                                           L.ptr(self._shift), count, self.epsilon, act, L.ptr(self.d_beta),
                                           L.ptr(self.d_gamma), npix, self.C, L.dtype_code(x), L.ptr(self._ws),
                                           L.stream_ptr()), "of_bn_train_bwd_reduce")
        if self.comm is not None:
            self.comm.allreduce_sum_(self.d_beta)
            self.comm.allreduce_sum_(self.d_gamma)
        dx = torch.empty_like(x)
        L.check(lib.of_bn_train_bwd_apply(L.ptr(x), L.ptr(y), L.ptr(dy), L.ptr(self._stats[0]), L.ptr(self._stats[1]),
                                          L.ptr(self._shift), count, L.ptr(self.gamma), self.epsilon, act, L.ptr(self.d_beta), L.ptr(self.d_gamma),
                                          L.ptr(dx), npix, self.C, L.dtype_code(x), L.stream_ptr()), "of_bn_train_bwd_apply")
        return dx
