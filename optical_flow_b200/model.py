"""Mirror of the reference's model.py: same entry points, CUDA underneath.

  warp_features(flow_1_to_2, features2)                    model.py:55-73
  create_cost_volume(features1, features2, max_disp)       model.py:29-42
  upscale_flow(flow)                                       model.py:76-77
  flow_module(features1, features2, previous_flow, max_disp[, net, level])   model.py:80-116
  reset18_encoder(height, width, name[, net])              model.py:10-26
  build_flow_net(height, width, pretrained_weights_path, max_disp=3)         model.py:119-143

Tensors are NHWC CUDA torch tensors used as buffers.  There is no autograd:
``FlowNet.__call__`` records what the reverse pass needs and
``FlowNet.backward(d_flows)`` runs the hand-written backward kernels in reverse
order, leaving d loss / d weight in ``net.store.grad`` (one flat fp32 buffer).

Two convolution back-ends share this graph:
  conv_mode="fp32"  SIMT fp32 implicit GEMM - exact-parity mode (<= 1e-5 vs oracle)
  conv_mode="bf16"  tcgen05/TMEM implicit GEMM, bf16 operands, fp32 accumulate
  conv_mode="bf16x3" tensor-core parity mode: fp32 storage, every conv as three bf16 tcgen05 passes (hi / lo split operands)
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import _lib as L
from . import functional as F
from .params import BN_EPS, DECODER_WIDTHS, FEATURE_CHANNELS, ParamStore


# --------------------------------------------------------------- functional --

def warp_features(flow_1_to_2, features2, mode="ref_ij", return_indices=False):
    """model.py:55-73.  mode='ref_ij' reproduces the reference exactly (its grid is
    [row, col] but the sampler reads (x, y): SURVEY F4); mode='xy' is the corrected order."""
    return F.warp_features_fwd(flow_1_to_2, features2, mode=mode, return_indices=return_indices)


def create_cost_volume(features1, features2, max_disp):
    """model.py:29-42."""
    return F.cost_volume_fwd(features1, features2, max_disp)


def upscale_flow(flow):
    """model.py:76-77."""
    return F.upscale_flow_fwd(flow)


def _copy_channels(src, src_off, dst, dst_off, nch, accumulate=False):
    """dst[..., dst_off:dst_off+nch] (=|+=) src[..., src_off:src_off+nch] (NHWC, any dtype mix)."""
    npix = src.numel() // src.shape[-1]
    assert dst.numel() // dst.shape[-1] == npix
    sp = src.data_ptr() + src_off * src.element_size()
    dp = dst.data_ptr() + dst_off * dst.element_size()
    L.check(L.load().of_copy_channels(sp, src.shape[-1], dp, dst.shape[-1], nch, npix, L.dtype_code(src),
                                      L.dtype_code(dst), 1 if accumulate else 0, L.stream_ptr()),
            "of_copy_channels")


def _accumulate(dst, src):
    """dst += src (flat)."""
    n = dst.numel()
    assert src.numel() == n
    g = 4 if n % 4 == 0 else 1
    L.check(L.load().of_copy_channels(L.ptr(src), g, L.ptr(dst), g, g, n // g, L.dtype_code(src),
                                      L.dtype_code(dst), 1, L.stream_ptr()), "of_copy_channels(acc)")


# -------------------------------------------------------------------- layers --

class ConvLayer:
    """One Keras Conv2D(padding='same') [+ inference-mode BatchNormalization] [+ activation]."""

    def __init__(self, net, name, ksize, cin, cout, stride=1, act=L.OF_ACT_NONE, bn=None, use_bias=True):
        self.net, self.name = net, name
        self.k, self.cin, self.cout, self.stride, self.act, self.bn = ksize, cin, cout, stride, act, bn
        self.use_bias = use_bias
        self.x = self.y = None
        self.prev_chain = None         # ConvLayer whose output is exactly this layer's input (sole consumer)
        self.incoming_is_dz = False    # set by a fused dgrad epilogue: dy already multiplied by act'(y)
        self.weights_ready = False     # FlowNet.prepare_weights() already produced w_eff / b_eff for this step
        if not hasattr(net, "conv_layers"):
            net.conv_layers = []
        net.conv_layers.append(self)

    # parameters (views into the flat buffers)
    def _p(self, suffix):
        return self.net.store.p[self.name + suffix]

    def _g(self, suffix):
        return self.net.store.g[self.name + suffix]

    def effective_weights(self):
        """(w, b) as used by the conv kernel: BN folded in when the layer has one."""
        net = self.net
        w = self._p("/kernel")
        b = self._p("/bias") if self.use_bias else None
        if self.bn is None or net.bn_mode == "train":     # batch-statistics BN is a layer of its own, nothing to fold
            return w, b
        K = self.k * self.k * self.cin
        wf = net.buf(self.name + ":wf", w.shape, torch.float32)
        bf = net.buf(self.name + ":bf", (self.cout,), torch.float32)
        st = net.store
        L.check(L.load().of_bn_fold_fwd(L.ptr(w), L.ptr(b), L.ptr(st.p[self.bn + "/gamma"]), L.ptr(st.p[self.bn + "/beta"]),
                                        L.ptr(st.stats[self.bn + "/moving_mean"]),
                                        L.ptr(st.stats[self.bn + "/moving_variance"]), BN_EPS, L.ptr(wf), L.ptr(bf),
                                        K, self.cout, L.stream_ptr()), "of_bn_fold_fwd")
        return wf, bf

    def forward(self, x, post_add=None):
        """``post_add`` (back-ends with ``can_post_add``): returns relu(act(conv(x)) + post_add) from the conv's own
        epilogue (resnet/models.py:24-35); the layer's activation then exists only as its sign-bit mask."""
        self.x = x
        if not self.weights_ready:
            self.w_eff, self.b_eff = self.effective_weights()
        self.weights_ready = False
        if post_add is not None:
            assert not (self.bn is not None and self.net.bn_mode == "train")
            self.y = None
            return self.net.backend.fprop(self, x, self.w_eff, self.b_eff, post_add=post_add)
        if self.bn is not None and self.net.bn_mode == "train":
            # conv (no activation) -> BatchNormalization(training=True) + activation (resnet/batchnorm.py)
            lin = self._linear_view()
            lin.x = x
            z = self.net.backend.fprop(lin, x, self.w_eff, self.b_eff)
            self.y = self._bn_layer()(z)
            return self.y
        self.y = self.net.backend.fprop(self, x, self.w_eff, self.b_eff)
        return self.y

    # -- opt-in training-mode BatchNormalization (FlowNet(bn_mode="train")) --------------
    def _linear_view(self):
        """This layer as the back-end sees it in bn_mode='train': the bare convolution (no folded BN, no fused
        activation, no chained epilogues), same weights and gradient buffers."""
        if getattr(self, "_lin", None) is None:
            lin = ConvLayer.__new__(ConvLayer)
            lin.__dict__.update(self.__dict__)
            lin.act, lin.bn, lin.prev_chain, lin.incoming_is_dz = L.OF_ACT_NONE, None, None, False
            self._lin = lin
        return self._lin

    def _bn_layer(self):
        if getattr(self, "_bn_obj", None) is None:
            from .resnet.batchnorm import BatchNormalizationTraining
            st = self.net.store
            tensors = dict(gamma=st.p[self.bn + "/gamma"], beta=st.p[self.bn + "/beta"],
                           moving_mean=st.stats[self.bn + "/moving_mean"],
                           moving_variance=st.stats[self.bn + "/moving_variance"],
                           d_gamma=st.g[self.bn + "/gamma"], d_beta=st.g[self.bn + "/beta"])
            self._bn_obj = BatchNormalizationTraining(self.cout, epsilon=BN_EPS, act=self.act, device=self.net.device,
                                                      comm=self.net.bn_comm, tensors=tensors)
        return self._bn_obj

    def _backward_bn_train(self, dy, need_dx, add):
        net = self.net
        bn = self._bn_layer()
        dz = bn.backward(dy, act_applied=self.incoming_is_dz)     # d loss / d conv output; fills d_gamma, d_beta
        self.incoming_is_dz = False
        if net.bn_comm is not None and net.bn_comm.world_size > 1:
            # the layer all-reduced its gradient sums; the trainer's exchange will sum the N identical copies again
            bn.d_gamma.mul_(1.0 / net.bn_comm.world_size)
            bn.d_beta.mul_(1.0 / net.bn_comm.world_size)
        lin = self._linear_view()
        net.backend.wgrad(lin, lin.x, dz, self._g("/kernel"), self._g("/bias") if self.use_bias else None)
        if not need_dx:
            return None
        in_hw = (lin.x.shape[1], lin.x.shape[2])
        if add is not None:
            return net.backend.dgrad(lin, dz, self.w_eff, in_hw, chain=None, add=add)
        return net.backend.dgrad(lin, dz, self.w_eff, in_hw, chain=None)

    def backward(self, dy, need_dx=True, add=None, out_mask=None):
        """dy: gradient w.r.t. the layer output (after activation).  Returns dx or None.
        ``add`` (back-ends with ``dgrad_can_add``): a tensor summed into dx by the dgrad epilogue.
        ``out_mask``: a post-ReLU tensor shaped like the layer input whose derivative mask the dgrad epilogue applies to
        dx + add (the producer of the input is not a ConvLayer this layer chains to, e.g. the previous residual block)."""
        net = self.net
        assert out_mask is None or (self.prev_chain is None and not (self.bn is not None and net.bn_mode == "train"))
        if (need_dx and self.bn is None and self.act == L.OF_ACT_NONE and self.k == 3 and self.stride == 1
                and self.cout == 2 and self.cin <= 32 and self.use_bias and hasattr(net.backend, "head_bwd")):
            return net.backend.head_bwd(self, dy)     # flow head: dw, db and the chained dz in one pass
        if self.bn is not None and net.bn_mode == "train":
            return self._backward_bn_train(dy, need_dx, add)
        if self.incoming_is_dz:
            dz, self.incoming_is_dz = dy, False
        else:
            dz = net.backend.act_bwd(self, self.y, dy)
        if self.bn is None:
            dw = self._g("/kernel")
            db = self._g("/bias") if self.use_bias else None
            net.backend.wgrad(self, self.x, dz, dw, db)
        else:
            K = self.k * self.k * self.cin
            dwf = net.buf(self.name + ":dwf", self._p("/kernel").shape, torch.float32)
            dbf = net.buf(self.name + ":dbf", (self.cout,), torch.float32)
            net.backend.wgrad(self, self.x, dz, dwf, dbf)
            st = net.store
            b = self._p("/bias") if self.use_bias else None
            if not getattr(net, "_bn_bwd_deferred", False):     # FlowNet.backward un-folds every layer in one launch
                L.check(L.load().of_bn_fold_bwd(
                    L.ptr(self._p("/kernel")), L.ptr(b), L.ptr(st.p[self.bn + "/gamma"]),
                    L.ptr(st.stats[self.bn + "/moving_mean"]), L.ptr(st.stats[self.bn + "/moving_variance"]), BN_EPS,
                    L.ptr(dwf), L.ptr(dbf), L.ptr(self._g("/kernel")), L.ptr(self._g("/bias")) if self.use_bias else None,
                    L.ptr(st.g[self.bn + "/gamma"]), L.ptr(st.g[self.bn + "/beta"]), K, self.cout, L.stream_ptr()),
                    "of_bn_fold_bwd")
        if not need_dx:
            return None
        chain = self._chain() if out_mask is None else _OutMask(out_mask)
        if add is not None:
            return net.backend.dgrad(self, dz, self.w_eff, (self.x.shape[1], self.x.shape[2]), chain=chain, add=add)
        return net.backend.dgrad(self, dz, self.w_eff, (self.x.shape[1], self.x.shape[2]), chain=chain)

    def _chain(self):
        """The producer whose activation derivative this layer's dgrad epilogue may apply - none when the producer's
        activation sits behind a training-mode BatchNormalization layer."""
        c = self.prev_chain
        return None if (c is not None and c.bn is not None and self.net.bn_mode == "train") else c


class _OutMask:
    """Stand-in for a chained producer: a post-ReLU tensor whose derivative mask a dgrad epilogue applies."""

    def __init__(self, y):
        self.y, self.act, self.ybits, self.incoming_is_dz = y, L.OF_ACT_RELU, None, False


class Fp32Backend:
    """SIMT fp32 implicit-GEMM convolutions (exact-parity mode)."""
    act_dtype = torch.float32

    def __init__(self, net):
        self.net = net

    def fprop(self, layer, x, w, b):
        return F.conv2d_fprop(x, w, b, stride=layer.stride, act=layer.act, out_dtype=torch.float32)

    def act_bwd(self, layer, y, dy):
        if layer.act == L.OF_ACT_NONE:
            return dy
        return F.act_bwd(y, dy, layer.act)

    def wgrad(self, layer, x, dz, dw, db):
        F.conv2d_wgrad(x, dz, (layer.k, layer.k), stride=layer.stride, dw=dw, db=db, need_bias_grad=db is not None)

    def dgrad(self, layer, dz, w, in_hw, chain=None):
        return F.conv2d_dgrad(dz, w, in_hw, stride=layer.stride, out_dtype=torch.float32)


class ResnetBlock:
    """resnet/models.py:8-36 resnet_block_simple (ReLU BEFORE the residual add)."""

    def __init__(self, net, cin, downsample, layer_idx, block_idx):
        base = "layer%d_block_%d" % (layer_idx, block_idx)
        cout = 2 * cin if downsample else cin
        s = 2 if downsample else 1
        self.net, self.down, self.cout = net, downsample, cout
        self.conv1 = ConvLayer(net, base + "_conv1", 3, cin, cout, s, L.OF_ACT_RELU, bn=base + "_bn1")
        self.conv2 = ConvLayer(net, base + "_conv2", 3, cout, cout, 1, L.OF_ACT_RELU, bn=base + "_bn2")
        self.shortcut = ConvLayer(net, base + "_shortcut", 1, cin, cout, 2, L.OF_ACT_NONE, use_bias=False) if downsample else None
        self.conv2.prev_chain = self.conv1

    def forward(self, x):
        self.x = x
        h1 = self.conv1.forward(x)
        sc = self.shortcut.forward(x) if self.down else x
        be = self.net.backend
        if self.net.bn_mode != "train" and hasattr(be, "can_post_add") and be.can_post_add(self.conv2):
            self.out = self.conv2.forward(h1, post_add=sc)        # relu(relu(bn2(conv2)) + shortcut) in conv2's epilogue
        else:
            self.out = F.add_relu_fwd(self.conv2.forward(h1), sc)
        return self.out

    def backward(self, d_out, d_is_pre=False, dx_mask=None, skip_add=None):
        """``d_is_pre``: d_out already carries this block's last ReLU derivative (the next block's dgrad epilogue applied
        it).  ``dx_mask``: the previous block's output - its ReLU derivative is applied to the returned gradient.
        ``skip_add`` (downsample blocks): a gradient that reaches the block input from elsewhere (the decoder), summed
        in by the shortcut's dgrad epilogue."""
        d_pre = d_out if d_is_pre else F.act_bwd(self.out, d_out, L.OF_ACT_RELU)        # relu3
        d_y1 = self.conv2.backward(d_pre)
        assert skip_add is None or self.down
        skip = self.shortcut.backward(d_pre, add=skip_add) if self.down else d_pre
        if getattr(self.net.backend, "dgrad_can_add", False):
            return self.conv1.backward(d_y1, add=skip, out_mask=dx_mask)   # residual sum (+ mask) in the dgrad epilogue
        assert dx_mask is None and skip_add is None
        d_x = self.conv1.backward(d_y1)
        _accumulate(d_x, skip)
        return d_x


class Encoder:
    """model.py:10-26 reset18_encoder.  Runs once on the 2B stacked frames (shared weights,
    model.py:131-132)."""

    def __init__(self, net):
        self.net = net
        self.conv1 = ConvLayer(net, "conv1", 7, 3, 64, 2, L.OF_ACT_RELU, bn="layer1_bn")
        self.blocks = []
        cin = 64
        for layer_idx, down in ((2, False), (3, True), (4, True)):
            for blk in (1, 2):
                b = ResnetBlock(net, cin, down and blk == 1, layer_idx, blk)
                self.blocks.append(b)
                cin = b.cout

    def __call__(self, x):
        return self.forward(x)

    def forward(self, x):
        out0 = self.conv1.forward(x)                 # (N, h/2, w/2, 64)
        self.out0 = out0
        y = F.maxpool2_fwd(out0)                     # model.py:17 MaxPool2D(): 2x2/2
        outs = [out0]
        for k, b in enumerate(self.blocks):
            y = b.forward(y)
            if k % 2 == 1:
                outs.append(y)
        return outs                                  # [h/2 x64, h/4 x64, h/8 x128, h/16 x256]

    def backward(self, d_outs):
        """d_outs: gradients w.r.t. the 4 outputs (same dtype as the activations)."""
        d = d_outs[3]
        # inside a stage the second block's dgrad epilogue also applies the first block's last ReLU derivative
        fuse = (self.net.fuse_residual_epilogues and self.net.bn_mode != "train"
                and getattr(self.net.backend, "dgrad_can_add", False))
        # ... and at a stage boundary the decoder's gradient of the previous stage's output rides the shortcut's dgrad
        # epilogue, so every block but the last receives its gradient with its own ReLU derivative already applied
        pre = False
        for k in range(5, -1, -1):
            boundary = k % 2 == 0 and k > 0
            dx_mask = self.blocks[k - 1].out if (fuse and k > 0) else None
            skip_add = d_outs[k // 2] if (fuse and boundary) else None
            d = self.blocks[k].backward(d, d_is_pre=pre, dx_mask=dx_mask, skip_add=skip_add)
            pre = dx_mask is not None
            if boundary and skip_add is None:
                _accumulate(d, d_outs[k // 2])       # output of the previous stage also feeds the decoder
        # pool backward + the finest decoder level's gradient + the stem's ReLU derivative, one pass
        dz0 = F.maxpool2_bwd_add_act(self.out0, d, d_outs[0], self.conv1.act)
        self.conv1.incoming_is_dz = True
        self.conv1.backward(dz0, need_dx=False)      # images are constants


class FlowModule:
    """model.py:80-116 flow_module for one pyramid level (0 = coarsest)."""

    def __init__(self, net, level):
        self.net, self.level = net, level
        self.C = FEATURE_CHANNELS[level]
        self.ncv = (2 * net.max_disp + 1) ** 2
        self.cin = self.C + self.ncv + (2 if level > 0 else 0)
        self.convs = []
        cin = self.cin
        for ci, cout in enumerate(DECODER_WIDTHS):
            act = L.OF_ACT_LEAKY if ci < 5 else L.OF_ACT_NONE      # LeakyReLU() alpha 0.3
            self.convs.append(ConvLayer(net, "flow%d_conv%d" % (level, ci + 1), 3, cin, cout, 1, act))
            if ci > 0:
                self.convs[ci].prev_chain = self.convs[ci - 1]
            cin = cout

    def forward(self, features1, features2, previous_flow):
        net = self.net
        assert features1.shape[1] == features2.shape[1]          # model.py:83-85
        assert features1.shape[2] == features2.shape[2]
        assert features1.shape[3] == features2.shape[3]
        self.f1, self.f2, self.has_prev = features1, features2, previous_flow is not None
        B, h, w, C = features1.shape
        if previous_flow is not None:
            assert previous_flow.shape[0] == features1.shape[0]  # model.py:88-91
            assert previous_flow.shape[1] * 2 == features1.shape[1]
            assert previous_flow.shape[2] * 2 == features1.shape[2]
            assert previous_flow.shape[3] == 2
            self.flow_up = F.upscale_flow_fwd(previous_flow)
            # sampling coordinates: this level's own up-scaled flow, unless the caller pinned them
            # (net.frozen_warp_flows: parity tests that separate conv rounding from warp-index flips)
            frozen = net.frozen_warp_flows
            self.warp_flow = frozen[self.level] if frozen is not None else self.flow_up
            self.f2w = F.warp_features_fwd(self.warp_flow, features2, mode=net.warp_mode)
        else:
            self.flow_up, self.f2w = None, features2
        x = net.backend_concat_fwd(self, features1, self.f2w, self.flow_up)
        for conv in self.convs:
            x = conv.forward(x)
        return x                                                  # (B, h, w, 2) float32

    def backward(self, d_flow, out_f1=None, out_f2=None):
        """d_flow: f32 gradient w.r.t. this level's flow.  Returns (d_f1, d_f2, d_prev_flow).
        ``out_f1`` / ``out_f2`` (optional, activation dtype): where the caller wants the feature
        gradients; a back-end that can write them in place returns those very tensors."""
        net = self.net
        d = d_flow                      # the flow head has no activation: its dz is the fp32 d_flow
        for conv in reversed(self.convs):
            d = conv.backward(d)
        d_f1, d_f2w, d_flow_up = net.backend_concat_bwd(self, d, out_f1, out_f2)
        if self.has_prev:
            d_f2, d_flow_up_b = F.warp_features_bwd(self.warp_flow, self.f2, d_f2w, mode=net.warp_mode)
            _accumulate(d_flow_up, d_flow_up_b)
            d_prev = F.upscale_flow_bwd(d_flow_up)
        else:
            d_f2, d_prev = d_f2w, None
        return d_f1, d_f2, d_prev


# ----------------------------------------------------------------------- net --

class FlowNet:
    """model.py:119-143 build_flow_net(...) result: ``flow_net(batch_imgs)`` returns
    [flow3 (H/2), flow2 (H/4), flow1 (H/8), flow0 (H/16)], finest first."""

    def __init__(self, height, width, pretrained_weights_path=None, max_disp=3, conv_mode="fp32",
                 warp_mode="ref_ij", device=None, seed=42, bn_mode="inference", bn_comm=None,
                 fuse_residual_epilogues=True):
        """bn_mode: "inference" (default = the reference: train.py:51 calls the model without training=True, SURVEY F5;
        BN is folded into the convolutions) or "train" (opt-in, SURVEY 8f rank 4: batch statistics, moving-statistics
        update, cross-replica sums through ``bn_comm`` - anything with world_size and allreduce_sum_)."""
        assert bn_mode in ("inference", "train")
        self.bn_mode, self.bn_comm = bn_mode, bn_comm
        # bf16 back-end: the residual blocks' element-wise passes move into conv epilogues - conv2's epilogue computes
        # relu(relu(.) + shortcut), the dgrad epilogues apply the previous block's ReLU mask and take the decoder's
        # gradient at stage boundaries.  With the run-time-flagged epilogue this was a wash (call 20: step 10.27 -> 10.19 ms,
        # conv_tc3 +0.27 ms); with its own compile-time specialisations (call 32, same box) 9.17 -> 8.98 ms, the conv
        # kernels paying 0.14 ms for 0.32 ms of removed element-wise launches - on by default
        self.fuse_residual_epilogues = bool(fuse_residual_epilogues)
        assert torch.cuda.is_available(), "optical_flow_b200 needs a CUDA device (no CPU fallback)"
        assert height % 16 == 0 and width % 16 == 0, "height/width must be multiples of 16 (4 stride-2 stages)"
        L.load()
        self.height, self.width, self.max_disp = height, width, max_disp
        self.warp_mode = warp_mode
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.store = ParamStore(self.device, max_disp)
        self.store.init_glorot(seed)
        self._bufs = {}
        self.frozen_warp_flows = None      # {level: (B,h,w,2) f32} overrides the warp coordinates (tests only)
        self.conv_mode = conv_mode
        if conv_mode == "fp32":
            self.backend = Fp32Backend(self)
        elif conv_mode == "bf16":
            from .conv_tc import Bf16Backend
            self.backend = Bf16Backend(self)
        elif conv_mode == "bf16x3":
            from .conv_tc import Bf16x3Backend
            self.backend = Bf16x3Backend(self)
        else:
            raise AssertionError("conv_mode must be 'fp32', 'bf16' or 'bf16x3'")
        self.encoder = Encoder(self)
        self.flow_modules = [FlowModule(self, lvl) for lvl in range(4)]
        if pretrained_weights_path is not None:
            self.load_weights(pretrained_weights_path, encoder_only=True)    # model.py:127-129

    # -- buffers ---------------------------------------------------------------
    def buf(self, key, shape, dtype):
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(tuple(shape), dtype=dtype, device=self.device)
            self._bufs[key] = t
        return t

    def zero_buf(self, key, shape, dtype):
        """Persistent buffer zero-filled once at allocation (padding channels stay zero)."""
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.zeros(tuple(shape), dtype=dtype, device=self.device)
            self._bufs[key] = t
        return t

    def to_act_dtype(self, t):
        if t.dtype == self.backend.act_dtype:
            return t
        out = torch.empty(t.shape, dtype=self.backend.act_dtype, device=t.device)
        _copy_channels(t, 0, out, 0, t.shape[-1])
        return out

    def to_f32(self, t):
        if t.dtype == torch.float32:
            return t
        out = torch.empty(t.shape, dtype=torch.float32, device=t.device)
        _copy_channels(t, 0, out, 0, t.shape[-1])
        return out

    # concat [features1, cost_volume, flow_up] (model.py:97-102) and its gradient split
    def backend_concat_fwd(self, fm, f1, f2w, flow_up):
        if hasattr(self.backend, "concat_fwd"):
            return self.backend.concat_fwd(fm, f1, f2w, flow_up)
        B, h, w, C = f1.shape
        x = torch.empty((B, h, w, fm.cin), dtype=f1.dtype, device=f1.device)
        _copy_channels(f1, 0, x, 0, C)
        cvp = x.data_ptr() + C * x.element_size()
        L.check(L.load().of_cost_volume_fwd(L.ptr(f1), L.ptr(f2w), cvp, B, h, w, C, self.max_disp, fm.cin,
                                            L.dtype_code(f1), L.stream_ptr()), "of_cost_volume_fwd")
        if flow_up is not None:
            _copy_channels(flow_up, 0, x, C + fm.ncv, 2)
        return x

    def backend_concat_bwd(self, fm, dx, out_f1=None, out_f2=None):
        if hasattr(self.backend, "concat_bwd"):
            return self.backend.concat_bwd(fm, dx, out_f1, out_f2)
        B, h, w, _ = dx.shape
        C = fm.C
        d_f1 = torch.empty((B, h, w, C), dtype=torch.float32, device=dx.device)
        d_f2w = torch.empty((B, h, w, C), dtype=torch.float32, device=dx.device)
        dcvp = dx.data_ptr() + C * dx.element_size()
        L.check(L.load().of_cost_volume_bwd(L.ptr(fm.f1), L.ptr(fm.f2w), dcvp, L.ptr(d_f1), L.ptr(d_f2w), B, h, w, C,
                                            self.max_disp, fm.cin, L.dtype_code(fm.f1), L.stream_ptr()),
                "of_cost_volume_bwd")
        _copy_channels(dx, 0, d_f1, 0, C, accumulate=True)
        d_flow_up = None
        if fm.has_prev:
            d_flow_up = torch.empty((B, h, w, 2), dtype=torch.float32, device=dx.device)
            _copy_channels(dx, C + fm.ncv, d_flow_up, 0, 2)
        return d_f1, d_f2w, d_flow_up

    # -- forward / backward ------------------------------------------------------
    @property
    def trainable_weights(self):
        return list(self.store.p.values())

    def __call__(self, batch_imgs):
        return self.forward(batch_imgs)

    def forward(self, batch_imgs):
        B, H, W, c6 = batch_imgs.shape
        assert (H, W, c6) == (self.height, self.width, 6) and batch_imgs.dtype == torch.float32
        self.B = B
        self.prepare_weights()
        x2 = self.backend_split(batch_imgs)                       # (2B, H, W, 3): frame 1s then frame 2s
        enc = self.encoder(x2)
        self.enc_outs = enc
        flows = []
        prev = None
        for lvl in range(4):
            e = enc[3 - lvl]
            prev = self.flow_modules[lvl].forward(e[:B], e[B:], prev)
            flows.append(prev)
        return [flows[3], flows[2], flows[1], flows[0]]            # model.py:140

    def prepare_weights(self):
        """Effective (BN-folded) weights of every conv for this step, then - on back-ends that keep packed copies -
        all of their tensor-core packings in one launch."""
        cache = getattr(self.backend, "pack_cache", None)
        if cache is None:
            return
        cache.invalidate()
        table = self._bn_job_table()
        if table is not None:
            L.check(L.load().of_bn_fold_fwd_batch(L.ptr(table), self._bn_njobs, self._bn_fwd_blocks, BN_EPS,
                                                  L.stream_ptr()), "of_bn_fold_fwd_batch")
        for layer in self.conv_layers:
            if layer.bn is not None and table is not None:
                layer.w_eff = self.buf(layer.name + ":wf", layer._p("/kernel").shape, torch.float32)
                layer.b_eff = self.buf(layer.name + ":bf", (layer.cout,), torch.float32)
            else:
                layer.w_eff, layer.b_eff = layer.effective_weights()
            layer.weights_ready = True
        cache.refresh()

    def _bn_job_table(self):
        """Device table of of_bn_job, one per conv + BN pair (all buffers persistent): both folds of the whole
        model are then one launch each.  Built once."""
        if self.bn_mode == "train":
            return None                               # nothing is folded: BN runs as a layer
        if getattr(self, "_bn_table", None) is not None:
            return self._bn_table
        layers = [l for l in self.conv_layers if l.bn is not None]
        if not layers:
            return None
        arr = (L.BnJob * len(layers))()
        fb = bb = 0
        st = self.store
        for k, l in enumerate(layers):
            w = l._p("/kernel")
            K = l.k * l.k * l.cin
            j = arr[k]
            j.w, j.b = w.data_ptr(), (l._p("/bias").data_ptr() if l.use_bias else None)
            j.gamma, j.beta = st.p[l.bn + "/gamma"].data_ptr(), st.p[l.bn + "/beta"].data_ptr()
            j.mean, j.var = st.stats[l.bn + "/moving_mean"].data_ptr(), st.stats[l.bn + "/moving_variance"].data_ptr()
            j.wf = self.buf(l.name + ":wf", w.shape, torch.float32).data_ptr()
            j.bf = self.buf(l.name + ":bf", (l.cout,), torch.float32).data_ptr()
            j.dwf = self.buf(l.name + ":dwf", w.shape, torch.float32).data_ptr()
            j.dbf = self.buf(l.name + ":dbf", (l.cout,), torch.float32).data_ptr()
            j.dw, j.db = l._g("/kernel").data_ptr(), (l._g("/bias").data_ptr() if l.use_bias else None)
            j.dgamma, j.dbeta = st.g[l.bn + "/gamma"].data_ptr(), st.g[l.bn + "/beta"].data_ptr()
            j.K, j.Cout, j.fwd_block_begin, j.bwd_block_begin = K, l.cout, fb, bb
            fb += (K * l.cout + L.OF_BN_FWD_BLOCK - 1) // L.OF_BN_FWD_BLOCK
            bb += l.cout
        self._bn_njobs, self._bn_fwd_blocks, self._bn_bwd_blocks = len(layers), fb, bb
        import ctypes
        raw = torch.frombuffer(bytearray(ctypes.string_at(ctypes.addressof(arr), ctypes.sizeof(arr))), dtype=torch.uint8)
        self._bn_table = raw.to(self.device)
        return self._bn_table

    def backend_split(self, batch_imgs):
        if hasattr(self.backend, "split"):
            return self.backend.split(batch_imgs)
        B, H, W, _ = batch_imgs.shape
        out = torch.empty((2 * B, H, W, 3), dtype=torch.float32, device=batch_imgs.device)
        L.check(L.load().of_split_image_pair(L.ptr(batch_imgs), L.ptr(out), B, H, W, L.OF_F32, L.stream_ptr()),
                "of_split_image_pair")
        return out

    def backward(self, d_flows, after_decoder=None):
        """d_flows: [d_flow3, d_flow2, d_flow1, d_flow0] (finest first, like the outputs).  ``after_decoder``: called
        once the four flow modules' gradients are final and before the encoder's backward starts (the data-parallel
        trainer launches the decoder bucket's all-reduce there, under the encoder backward)."""
        B = self.B
        # with a packed-weight back-end every BN fold is un-folded by ONE launch after the last wgrad
        table = self._bn_job_table() if getattr(self.backend, "pack_cache", None) is not None else None
        self._bn_bwd_deferred = table is not None
        d_by_level = [d_flows[3], d_flows[2], d_flows[1], d_flows[0]]
        d_enc = [None] * 4
        carry = None
        for lvl in range(3, -1, -1):
            d = d_by_level[lvl]
            if carry is not None:
                _accumulate(carry, d)      # carry is our own buffer: d(loss)/d(flow) + d(finer level)/d(flow)
                d = carry
            e = self.enc_outs[3 - lvl]
            de = torch.empty(e.shape, dtype=self.backend.act_dtype, device=e.device)
            d_f1, d_f2, carry = self.flow_modules[lvl].backward(d, de[:B], de[B:])
            if d_f1.data_ptr() != de.data_ptr():              # back-end did not write in place
                _copy_channels(d_f1, 0, de[:B], 0, d_f1.shape[-1])
            if d_f2.data_ptr() != de[B:].data_ptr():
                _copy_channels(d_f2, 0, de[B:], 0, d_f2.shape[-1])
            d_enc[3 - lvl] = de
        if after_decoder is not None:
            after_decoder()
        try:
            self.encoder.backward(d_enc)
        finally:
            self._bn_bwd_deferred = False
        if table is not None:
            L.check(L.load().of_bn_fold_bwd_batch(L.ptr(table), self._bn_njobs, self._bn_bwd_blocks, BN_EPS,
                                                  L.stream_ptr()), "of_bn_fold_bwd_batch")

    # -- I/O (.npz with the reference's layer names) -------------
    def save_weights(self, path):
        np.savez(path if path.endswith(".npz") else path + ".npz", **self.store.to_numpy())

    def load_weights(self, path, encoder_only=False):
        data = np.load(path if path.endswith(".npz") else path + ".npz")
        for name in self.store.specs:
            if encoder_only and name.startswith("flow"):
                continue
            assert name in data, "checkpoint lacks " + name      # assert_existing_objects_matched
            self.store.p[name].copy_(torch.from_numpy(data[name]))
        for name in self.store.stats:
            # BN runs in inference mode folded into the convs: a checkpoint without the moving statistics would load
            # and silently run with mean 0 / variance 1 (the reference's assert_existing_objects_matched would raise)
            assert name in data, "checkpoint lacks " + name
            self.store.stats[name].copy_(torch.from_numpy(data[name]))

    def summary(self):
        lines = ["Model: flow_net (%dx%d, conv_mode=%s)" % (self.height, self.width, self.conv_mode)]
        for name, shape in self.store.specs.items():
            lines.append("  %-36s %s" % (name, tuple(shape)))
        lines.append("Trainable params: %d" % self.store.num_trainable())
        txt = "\n".join(lines)
        print(txt)
        return txt


def build_flow_net(height, width, pretrained_weights_path, max_disp=3, **kwargs):
    """model.py:119-143."""
    return FlowNet(height, width, pretrained_weights_path, max_disp=max_disp, **kwargs)


def reset18_encoder(height, width, name, net=None):
    """model.py:10-26: returns the encoder (callable on (N,h,w,3), 4 outputs)."""
    net = net or FlowNet(height, width, None)
    return net.encoder


def flow_module(features1, features2, previous_flow, max_disp, net=None, level=None):
    """model.py:80-116 as a function: needs the net (weights) and which level's convs to use."""
    assert net is not None and level is not None, "pass net= and level= (the reference creates fresh Keras layers here)"
    assert max_disp == net.max_disp
    return net.flow_modules[level].forward(features1, features2, previous_flow)
