"""bf16 tensor-core convolution back-end (tcgen05 / TMEM implicit GEMM in conv_tc.cu and
conv_tc_wgrad.cu) for FlowNet, plus thin functional wrappers used by the parity tests.

Activations are bf16 NHWC; weights stay fp32 masters (BN folded per step) and are re-packed
to bf16 GEMM layouts every step; accumulation is fp32 in TMEM.  Layers that are not
tensor-core shaped (Cin = 3 stem, Cout = 2 flow heads' dgrad/wgrad) use the SIMT kernels.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib as L
from . import functional as F


def _ceil(a, b):
    return (a + b - 1) // b * b


class Source:
    """One K-source of a conv: bf16 NHWC tensor, `valid` real channels that map to original
    input channels [ci0, ci0 + valid)."""

    def __init__(self, t, valid=None, ci0=0):
        assert t.dtype == torch.bfloat16 and t.is_contiguous()
        self.t, self.valid, self.ci0 = t, (t.shape[-1] if valid is None else valid), ci0


class MultiSource:
    """Stands for tf.concat([...], axis=-1) (model.py:100-102) without materialising it."""

    def __init__(self, sources):
        self.sources = sources
        self.shape = tuple(sources[0].t.shape[:3]) + (sum(s.valid for s in sources),)
        self.device = sources[0].t.device


def _sources_of(x):
    return x.sources if isinstance(x, MultiSource) else [Source(x)]


class PackCache:
    """Weight packings of a whole model, refreshed by ONE launch per step (of_conv_tc_pack_weights_batch).

    The first request for a (weight tensor, layout) packs it on the spot and registers the job; after
    ``refresh()`` (called once the step's effective weights exist, before the first conv) every registered
    packing is current and ``get`` returns the persistent buffer without launching anything.  Only meant for
    weights that live at fixed addresses (the model's parameter / folded-weight buffers)."""
    BLOCK = 2048

    def __init__(self):
        self.jobs = {}            # key -> (packed tensor, PackJob fields)
        self.table = None         # device copy of the job table
        self.total_blocks = 0
        self.fresh = False
        self.zero_upsampled = {}  # (layer key, shape) -> persistent zero-inserted gradient buffer (stride-2 dgrads)

    def get(self, w_hwio, segs, rows, cols, transpose):
        key = (w_hwio.data_ptr(), tuple(w_hwio.shape), tuple(tuple(s) for s in segs), rows, cols, bool(transpose))
        hit = self.jobs.get(key)
        if hit is not None and self.fresh:
            return hit[0]
        if hit is None:
            out = torch.empty((w_hwio.shape[0] * w_hwio.shape[1], rows, cols), dtype=torch.bfloat16, device=w_hwio.device)
            self.jobs[key] = (out, w_hwio)            # keeps the source tensor (and so its address) alive
            self.table = None
        return pack_weights(w_hwio, segs, rows, cols, transpose, out=self.jobs[key][0])

    def refresh(self):
        if not self.jobs:
            return
        if self.table is None:
            arr = (L.PackJob * len(self.jobs))()
            blk = 0
            for k, (key, (out, w)) in enumerate(self.jobs.items()):
                _, shape, segs, rows, cols, transpose = key
                j = arr[k]
                j.w, j.out = w.data_ptr(), out.data_ptr()
                j.taps, j.Cin, j.Cout, j.rows, j.cols = shape[0] * shape[1], shape[2], shape[3], rows, cols
                j.transpose, j.nseg = int(transpose), len(segs)
                for i, sg in enumerate(segs):
                    j.seg_off[i], j.seg_cnt[i], j.seg_base[i] = sg
                j.block_begin = blk
                blk += (j.taps * rows * cols + self.BLOCK - 1) // self.BLOCK
            self.total_blocks = blk
            raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            dev = next(iter(self.jobs.values()))[0].device
            self.table = raw.to(dev)
        L.check(L.load().of_conv_tc_pack_weights_batch(L.ptr(self.table), len(self.jobs), self.total_blocks, L.stream_ptr()),
                "of_conv_tc_pack_weights_batch")
        self.fresh = True

    def invalidate(self):
        self.fresh = False


def pack_weights(w_hwio, segs, rows, cols, transpose, out=None, cache=None):
    """fp32 HWIO -> bf16 [taps][rows][cols]; segs = [(orig_ci_off, count, packed_base), ...]."""
    if cache is not None:
        return cache.get(w_hwio, segs, rows, cols, transpose)
    KH, KW, Cin, Cout = w_hwio.shape
    taps = KH * KW
    if out is None:
        out = torch.empty((taps, rows, cols), dtype=torch.bfloat16, device=w_hwio.device)
    n = len(segs)
    L.check(L.load().of_conv_tc_pack_weights(L.ptr(w_hwio), L.ptr(out), taps, Cin, Cout, rows, cols,
                                             1 if transpose else 0, n, L.int_array([s[0] for s in segs]),
                                             L.int_array([s[1] for s in segs]), L.int_array([s[2] for s in segs]),
                                             L.stream_ptr()), "of_conv_tc_pack_weights")
    return out


def tc_run(sources, wpacked, n_mma, n_valid, N, H, W, Ho, Wo, KH, KW, stride, pad_t, pad_l, sign, out,
           bias=None, act=L.OF_ACT_NONE, add=None, mask=None, mask_act=L.OF_ACT_NONE, sign_out=None, mask_bits=None, out2=None, out_split=0,
           add_mode=0):
    d = L.TcConv()
    d.add_mode = add_mode
    for k, s in enumerate(sources):
        d.src[k] = s.t.data_ptr()
        d.src_c[k] = s.t.shape[-1]
        d.src_stride[k] = s.t.shape[-1]
    d.num_src = len(sources)
    d.wpacked = wpacked.data_ptr()
    d.n_rows, d.k_pad = wpacked.shape[1], wpacked.shape[2]
    d.N, d.H, d.W, d.Ho, d.Wo = N, H, W, Ho, Wo
    d.KH, d.KW, d.stride, d.pad_t, d.pad_l, d.sign = KH, KW, stride, pad_t, pad_l, sign
    d.n_mma, d.n_valid = n_mma, n_valid
    d.out, d.out_dtype, d.out_stride = out.data_ptr(), L.dtype_code(out), out.shape[-1]
    d.bias = L.ptr(bias)
    d.act = act
    if add is not None:
        d.add, d.add_dtype, d.add_stride = add.data_ptr(), L.dtype_code(add), add.shape[-1]
    if mask is not None:
        d.mask, d.mask_dtype, d.mask_stride, d.mask_act = mask.data_ptr(), L.dtype_code(mask), mask.shape[-1], mask_act
    if mask_bits is not None:                    # 16 B / pixel sign bits of the producing layer's output
        assert mask is None and mask_bits.dtype == torch.uint8 and mask_bits.shape[-1] == 16
        d.mask, d.mask_dtype, d.mask_stride, d.mask_act = mask_bits.data_ptr(), L.OF_MASK_BITS, 0, mask_act
    if out2 is not None:
        d.out2, d.out2_stride, d.out_split = out2.data_ptr(), out2.shape[-1], out_split
    if sign_out is not None:
        assert sign_out.dtype == torch.uint8 and sign_out.shape[-1] == 16
        d.sign_out = sign_out.data_ptr()
    L.check(L.load().of_conv_tc_run(C.byref(d), L.stream_ptr()), "of_conv_tc_run")
    return out


# ------------------------------------------------------------------ functional --

def sign_bits_ok(ksize, stride, cout):
    """Layers whose fprop can emit (and whose consumers' dgrad can read) the 16 B / pixel sign-bit mask."""
    return ksize == 3 and stride == 1 and 16 <= cout <= 128 and cout % 16 == 0


def conv2d_fprop_tc(x, w_hwio, bias, stride=1, act=L.OF_ACT_NONE, out_dtype=torch.bfloat16, add=None, sign_out=None,
                    cache=None, add_mode=0):
    """Keras Conv2D(padding='same') forward on tensor cores.  x: bf16 tensor or MultiSource."""
    srcs = _sources_of(x)
    N, H, W = srcs[0].t.shape[:3]
    KH, KW, Cin, Cout = w_hwio.shape
    Ho, pt, _ = F.same_pad(H, KH, stride)
    Wo, pl, _ = F.same_pad(W, KW, stride)
    segs, kb = [], 0
    for s in srcs:
        segs.append((s.ci0, s.valid, kb))
        kb += _ceil(s.t.shape[-1], 64)
    n_mma = _ceil(Cout, 16)
    wp = pack_weights(w_hwio, segs, n_mma, kb, transpose=False, cache=cache)
    out = torch.empty((N, Ho, Wo, Cout), dtype=out_dtype, device=w_hwio.device)
    return tc_run(srcs, wp, n_mma, Cout, N, H, W, Ho, Wo, KH, KW, stride, pt, pl, +1, out, bias=bias, act=act, add=add,
                  sign_out=sign_out, add_mode=add_mode)


def zero_upsample2(dy, H, W, cache=None, key=None):
    """Zero-inserted dy for a stride-2 dgrad.  With a cache (the network's PackCache) the buffer is allocated and
    zero-filled once per layer (`key`) and a step only rewrites the non-zero quarter."""
    N, Ho, Wo, Cc = dy.shape
    if cache is not None and key is not None:
        bufs = cache.zero_upsampled
        k = (key, N, H, W, Cc, dy.dtype)
        out = bufs.get(k)
        if out is None:
            out = bufs[k] = torch.zeros((N, H, W, Cc), dtype=dy.dtype, device=dy.device)
        L.check(L.load().of_scatter_upsample2(L.ptr(dy), L.ptr(out), N, Ho, Wo, H, W, Cc, L.dtype_code(dy),
                                              L.stream_ptr()), "of_scatter_upsample2")
        return out
    out = torch.empty((N, H, W, Cc), dtype=dy.dtype, device=dy.device)
    L.check(L.load().of_zero_upsample2(L.ptr(dy), L.ptr(out), N, Ho, Wo, H, W, Cc, L.dtype_code(dy), L.stream_ptr()),
            "of_zero_upsample2")
    return out


def conv2d_dgrad_tc(dz, w_hwio, in_hw, stride=1, seg=None, out_dtype=torch.bfloat16, mask=None,
                    mask_act=L.OF_ACT_NONE, out_channels=None, add=None, mask_bits=None, cache=None):
    """dx for input channels [seg[0], seg[0]+seg[1]) (default: all).  dz: bf16 (N,Ho,Wo,Cout).
    stride 2 is handled by zero-insertion of dz followed by the stride-1 kernel."""
    KH, KW, Cin, Cout = w_hwio.shape
    H, W = in_hw
    N = dz.shape[0]
    _, pt, _ = F.same_pad(H, KH, stride)
    _, pl, _ = F.same_pad(W, KW, stride)
    if stride == 2:
        dz = zero_upsample2(dz, H, W, cache=cache, key=w_hwio.data_ptr())
    ci0, cnt = seg if seg is not None else (0, Cin)
    n_out = out_channels or cnt
    n_mma = _ceil(n_out, 16)
    kpad = _ceil(dz.shape[-1], 64)
    wp = pack_weights(w_hwio, [(ci0, cnt, 0)], n_mma, kpad, transpose=True, cache=cache)
    out = torch.empty((N, H, W, n_out), dtype=out_dtype, device=dz.device)
    assert add is None or tuple(add.shape) == tuple(out.shape)
    return tc_run([Source(dz)], wp, n_mma, n_out, N, H, W, H, W, KH, KW, 1, pt, pl, -1, out, mask=mask, mask_act=mask_act,
                  add=add, mask_bits=mask_bits)


def conv2d_wgrad_tc(x, dz, ksize, stride=1, dw=None, db=None, need_bias_grad=True):
    srcs = _sources_of(x)
    N, H, W = srcs[0].t.shape[:3]
    KH, KW = ksize
    Cout = dz.shape[-1]
    Cin = sum(s.valid for s in srcs)
    Ho, pt, _ = F.same_pad(H, KH, stride)
    Wo, pl, _ = F.same_pad(W, KW, stride)
    assert tuple(dz.shape[:3]) == (N, Ho, Wo) and dz.dtype == torch.bfloat16
    if dw is None:
        dw = torch.empty((KH, KW, Cin, Cout), dtype=torch.float32, device=dz.device)
    d = L.TcWgrad()
    for k, s in enumerate(srcs):
        d.x[k] = s.t.data_ptr()
        d.x_c[k] = s.t.shape[-1]
        d.x_stride[k] = s.t.shape[-1]
        d.x_valid[k] = s.valid
        d.x_ci0[k] = s.ci0
    d.num_src = len(srcs)
    d.dy, d.dy_c, d.dy_stride = dz.data_ptr(), Cout, Cout
    d.dw = dw.data_ptr()
    d.N, d.H, d.W, d.Ho, d.Wo = N, H, W, Ho, Wo
    d.KH, d.KW, d.stride, d.pad_t, d.pad_l, d.Cin, d.Cout = KH, KW, stride, pt, pl, Cin, Cout
    if need_bias_grad and db is None:
        db = torch.empty((Cout,), dtype=torch.float32, device=dz.device)
    # bias gradient fused into the wgrad launch (transposed ones-operand MMAs: 32 TMEM columns, so every Cout <= 256
    # qualifies; at Cout = 256 the kernel then runs one tap per pass instead of two, which costs less than the separate
    # column-sum pass did: step 8.98 -> 8.90 ms, round 2 call 58)
    fuse_db = need_bias_grad and (Cout + 15) // 16 * 16 <= 256
    if fuse_db:
        d.db = db.data_ptr()
    L.check(L.load().of_conv_tc_wgrad(C.byref(d), L.stream_ptr()), "of_conv_tc_wgrad")
    if need_bias_grad and not fuse_db:
        L.check(L.load().of_colsum(L.ptr(dz), L.ptr(db), N * Ho * Wo, Cout, Cout, L.dtype_code(dz), L.stream_ptr()),
                "of_colsum")
    return dw, db


# ------------------------------------------------------- stem (Cin = 3) on tensor cores --

def pack_image_pair8(batch_imgs):
    """(B,H,W,6) f32 -> bf16 (2B, (H+6)/2, W+6, 8): frame 1s then frame 2s (model.py:122-123,131-132), TF-SAME
    zero border of the 7x7/2 stem; a texel holds the pixels of two consecutive rows x 4 channels (3 used), so 8
    consecutive texels = one 64-wide K chunk covering two kh taps."""
    B, H, W, _ = batch_imgs.shape
    assert H % 2 == 0 and W % 2 == 0
    out = torch.empty((2 * B, (H + 6) // 2, W + 6, 8), dtype=torch.bfloat16, device=batch_imgs.device)
    L.check(L.load().of_pack_image_pair8(L.ptr(batch_imgs), L.ptr(out), B, H, W, L.stream_ptr()), "of_pack_image_pair8")
    return out


def _stem_view(d, xp, H, W):
    """Describe the packed image as the strided view the stem convolves: windows of 8 pixels every 2."""
    Hp, Wp = (H + 6) // 2, W + 6                      # row pairs, padded columns
    d.N, d.H, d.W, d.Ho, d.Wo = xp.shape[0], Hp, W // 2, H // 2, W // 2
    d.KH, d.KW, d.stride, d.pad_t, d.pad_l = 4, 1, 1, 0, 0
    d.stride_w, d.stride_h = 1, 1
    d.view_w_stride, d.view_h_stride, d.view_n_stride = 16, Wp * 8, Hp * Wp * 8


def conv1_fprop_tc(xp, H, W, w_hwio, bias, act=L.OF_ACT_RELU, out_dtype=torch.bfloat16):
    """model.py:12-15 stem: Conv2D(64, 7, strides=2, 'same') (+ folded BN + ReLU) as a 4-tap (row pair) conv over the view."""
    Cout = w_hwio.shape[3]
    assert tuple(w_hwio.shape[:3]) == (7, 7, 3) and Cout % 16 == 0
    wp = torch.empty((4, Cout, 64), dtype=torch.bfloat16, device=xp.device)
    L.check(L.load().of_conv1_pack_weights(L.ptr(w_hwio), L.ptr(wp), Cout, L.stream_ptr()), "of_conv1_pack_weights")
    out = torch.empty((xp.shape[0], H // 2, W // 2, Cout), dtype=out_dtype, device=xp.device)
    d = L.TcConv()
    d.src[0], d.src_c[0], d.src_stride[0], d.num_src = xp.data_ptr(), 64, 16, 1
    d.wpacked, d.n_rows, d.k_pad = wp.data_ptr(), Cout, 64
    _stem_view(d, xp, H, W)
    d.sign, d.n_mma, d.n_valid = 1, Cout, Cout
    d.out, d.out_dtype, d.out_stride = out.data_ptr(), L.dtype_code(out), Cout
    d.bias, d.act = L.ptr(bias), act
    L.check(L.load().of_conv_tc_run(C.byref(d), L.stream_ptr()), "of_conv_tc_run(stem)")
    return out


def conv1_wgrad_tc(xp, H, W, dz, dw, db):
    """dw (7,7,3,Cout) f32, db (Cout) f32 of the stem from the packed image and bf16 dz."""
    Cout = dz.shape[-1]
    dwv = torch.empty((4, 64, Cout), dtype=torch.float32, device=dz.device)
    d = L.TcWgrad()
    d.x[0], d.x_c[0], d.x_stride[0], d.x_valid[0], d.x_ci0[0], d.num_src = xp.data_ptr(), 64, 16, 64, 0, 1
    d.dy, d.dy_c, d.dy_stride = dz.data_ptr(), Cout, Cout
    d.dw = dwv.data_ptr()
    _stem_view(d, xp, H, W)
    d.Cin, d.Cout = 64, Cout
    if db is not None:
        d.db = db.data_ptr()                               # bias gradient fused into the same launch
    L.check(L.load().of_conv_tc_wgrad(C.byref(d), L.stream_ptr()), "of_conv_tc_wgrad(stem)")
    L.check(L.load().of_conv1_unpack_wgrad(L.ptr(dwv), L.ptr(dw), Cout, L.stream_ptr()), "of_conv1_unpack_wgrad")
    return dw, db


# --------------------------------------------------------------------- backend --

class Bf16Backend:
    """FlowNet convolution back-end: bf16 activations, tcgen05 kernels."""
    act_dtype = torch.bfloat16

    def __init__(self, net):
        self.net = net
        self.pack_cache = None if os.environ.get("OF_NO_PACK_CACHE") else PackCache()

    @staticmethod
    def _tc_ok(layer):
        return layer.cin >= 32 and layer.cout >= 16

    def split(self, batch_imgs):
        """model.py:122-123,131-132 + the stem's input layout: the packed bf16 pair (see pack_image_pair8)."""
        self.img_hw = (batch_imgs.shape[1], batch_imgs.shape[2])
        return pack_image_pair8(batch_imgs)

    def can_post_add(self, layer):
        """resnet/models.py:24-35 in one launch: relu(relu(bn2(conv2(.))) + shortcut) from the conv's epilogue (the inner
        activation survives as the sign-bit mask its backward reads)."""
        return (self.net.fuse_residual_epilogues and layer.act == L.OF_ACT_RELU
                and sign_bits_ok(layer.k, layer.stride, layer.cout) and layer.cin >= 32)

    def fprop(self, layer, x, w, b, post_add=None):
        if layer.cin < 32:          # conv1 (Cin = 3, 7x7/2): 4-tap (row pair) tensor-core conv over the packed image
            return conv1_fprop_tc(x, self.img_hw[0], self.img_hw[1], w, b, act=layer.act)
        out_dtype = torch.float32 if layer.cout < 16 else torch.bfloat16     # flow heads stay fp32
        layer.ybits = None
        if layer.act != L.OF_ACT_NONE and sign_bits_ok(layer.k, layer.stride, layer.cout) and layer.cin >= 32:
            t0 = x.sources[0].t if isinstance(x, MultiSource) else x
            layer.ybits = torch.empty((t0.shape[0], t0.shape[1], t0.shape[2], 16), dtype=torch.uint8, device=t0.device)
        if post_add is not None:
            assert layer.ybits is not None and self.can_post_add(layer)
            return conv2d_fprop_tc(x, w, b, stride=layer.stride, act=layer.act, out_dtype=out_dtype, sign_out=layer.ybits,
                                   cache=self.pack_cache, add=post_add, add_mode=1)
        return conv2d_fprop_tc(x, w, b, stride=layer.stride, act=layer.act, out_dtype=out_dtype, sign_out=layer.ybits,
                               cache=self.pack_cache)

    def act_bwd(self, layer, y, dy):
        if layer.act == L.OF_ACT_NONE:
            return dy
        bits = getattr(layer, "ybits", None)
        if bits is not None and dy.shape[-1] % 8 == 0 and tuple(bits.shape[:3]) == tuple(dy.shape[:3]):
            return F.act_bwd_bits(bits, dy, layer.act)  # 1/16 of the mask bytes (and all there is after a fused post-add)
        assert y is not None
        return F.act_bwd(y, dy, layer.act)

    def wgrad(self, layer, x, dz, dw, db):
        if layer.cin < 32:
            conv1_wgrad_tc(x, self.img_hw[0], self.img_hw[1], dz, dw, db)
        elif self._tc_ok(layer):
            conv2d_wgrad_tc(x, dz, (layer.k, layer.k), stride=layer.stride, dw=dw, db=db, need_bias_grad=db is not None)
        else:
            assert not isinstance(x, MultiSource)
            F.conv2d_wgrad(x, dz, (layer.k, layer.k), stride=layer.stride, dw=dw, db=db, need_bias_grad=db is not None)

    def head_bwd(self, layer, dy):
        """model.py:114 flow head backward in one launch (csrc/head.cu): fills the layer's dw / db and
        returns the previous conv's dz (its activation derivative applied from x = its output)."""
        x = layer.x
        N, H, W, Cin = x.shape
        chain = layer.prev_chain
        act = chain.act if chain is not None else L.OF_ACT_NONE
        dz = torch.empty_like(x)
        L.check(L.load().of_flow_head_bwd(L.ptr(x), L.ptr(dy), L.ptr(layer.w_eff), L.ptr(layer._g("/kernel")),
                                          L.ptr(layer._g("/bias")), L.ptr(dz), N, H, W, Cin, act, L.dtype_code(x),
                                          L.stream_ptr()), "of_flow_head_bwd")
        if chain is not None and act != L.OF_ACT_NONE:
            chain.incoming_is_dz = True
        return dz

    dgrad_can_add = True

    def dgrad(self, layer, dz, w, in_hw, chain=None, add=None):
        """Returns dx (+ ``add``, summed in the epilogue); with `chain` (the ConvLayer producing this layer's input, sole consumer)
        the epilogue multiplies by act'(chain.y) and the result is already chain's dz."""
        x = layer.x
        if not self._tc_ok(layer):                       # flow heads: dz is the fp32 d_flow
            return F.conv2d_dgrad(dz, w, in_hw, stride=layer.stride, out_dtype=torch.bfloat16)
        mask = mask_act = mask_bits = None
        if chain is not None and chain.act != L.OF_ACT_NONE:
            mask, mask_act = chain.y, chain.act
            chain.incoming_is_dz = True
            # the producer's fprop left a 16 B / pixel sign-bit mask: read that instead of the bf16 activation
            if getattr(chain, "ybits", None) is not None and layer.k == 3 and layer.stride == 1 and not isinstance(x, MultiSource):
                mask, mask_bits = None, chain.ybits
        if (isinstance(x, MultiSource) and len(x.sources) == 2 and layer.k == 3 and layer.stride == 1
                and all(s.t.shape[-1] % 16 == 0 for s in x.sources) and sum(s.t.shape[-1] for s in x.sources) <= 128
                and not os.environ.get("OF_DGRAD_NO_SPLIT")):
            # both concat segments from ONE launch: dz is read once, each segment's gradient goes to its own tensor
            s0, s1 = x.sources
            n0, n1 = s0.t.shape[-1], s1.t.shape[-1]
            N = dz.shape[0]
            H, W = in_hw
            _, pt, _ = F.same_pad(H, 3, 1)
            _, pl, _ = F.same_pad(W, 3, 1)
            wp = pack_weights(w, [(s0.ci0, s0.valid, 0), (s1.ci0, s1.valid, n0)], n0 + n1, _ceil(dz.shape[-1], 64), transpose=True,
                              cache=self.pack_cache)
            o0 = torch.empty((N, H, W, n0), dtype=torch.bfloat16, device=dz.device)
            o1 = torch.empty((N, H, W, n1), dtype=torch.bfloat16, device=dz.device)
            tc_run([Source(dz)], wp, n0 + n1, n0 + n1, N, H, W, H, W, 3, 3, 1, pt, pl, -1, o0, out2=o1, out_split=n0)
            return (o0, o1)
        if isinstance(x, MultiSource):
            outs = []
            for s in x.sources:
                n_out = s.t.shape[-1]                    # padded channels of the cv/flow buffer get zeros
                dt = torch.bfloat16                       # features segment too: summed in fp32 inside concat_bwd
                outs.append(conv2d_dgrad_tc(dz, w, in_hw, 1, seg=(s.ci0, s.valid), out_dtype=dt, out_channels=n_out,
                                            cache=self.pack_cache))
            return tuple(outs)
        assert add is None or self._tc_ok(layer)
        return conv2d_dgrad_tc(dz, w, in_hw, layer.stride, out_dtype=torch.bfloat16, mask=mask,
                               mask_act=mask_act if (mask is not None or mask_bits is not None) else L.OF_ACT_NONE,
                               add=add, mask_bits=mask_bits, cache=self.pack_cache)

    # concat [features1, cost_volume, flow_up] (model.py:97-102) as two K-sources
    def concat_fwd(self, fm, f1, f2w, flow_up):
        net = self.net
        B, h, w, Cc = f1.shape
        cvflow = net.zero_buf("cvflow%d" % fm.level, (B, h, w, 64), torch.bfloat16)
        L.check(L.load().of_cost_volume_fwd(L.ptr(f1), L.ptr(f2w), L.ptr(cvflow), B, h, w, Cc, net.max_disp, 64,
                                            L.OF_BF16, L.stream_ptr()), "of_cost_volume_fwd")
        valid = fm.ncv
        if flow_up is not None:
            L.check(L.load().of_copy_channels(L.ptr(flow_up), 2, cvflow.data_ptr() + fm.ncv * 2, 64, 2, B * h * w,
                                              L.OF_F32, L.OF_BF16, 0, L.stream_ptr()), "of_copy_channels")
            valid += 2
        fm.cvflow = cvflow
        return MultiSource([Source(f1, Cc, 0), Source(cvflow, valid, Cc)])

    def concat_bwd(self, fm, d, out_f1=None, out_f2=None):
        """Gradient split of the concat (model.py:100-102) + cost-volume backward, bf16 end to end:
        d_f1 = conv segment + dcv/df1 written straight into ``out_f1`` (the encoder's incoming
        gradient); d_f2w goes to ``out_f2`` when there is no warp in between (coarsest level)."""
        d_f1_conv, d_cvflow = d                          # bf16 (B,h,w,C), bf16 (B,h,w,64)
        B, h, w, Cc = d_f1_conv.shape
        net = self.net
        d_f1 = out_f1 if out_f1 is not None else torch.empty_like(d_f1_conv)
        d_f2w = out_f2 if (out_f2 is not None and not fm.has_prev) else torch.empty_like(d_f1_conv)
        assert d_f1.is_contiguous() and d_f2w.is_contiguous() and d_f1.dtype == torch.bfloat16
        L.check(L.load().of_cost_volume_bwd_bf16(L.ptr(fm.f1), L.ptr(fm.f2w), L.ptr(d_cvflow), L.ptr(d_f1), L.ptr(d_f2w),
                                                 L.ptr(d_f1_conv), B, h, w, Cc, net.max_disp, 64, L.stream_ptr()),
                "of_cost_volume_bwd_bf16")
        d_flow_up = None
        if fm.has_prev:
            d_flow_up = torch.empty((B, h, w, 2), dtype=torch.float32, device=d_f1.device)
            L.check(L.load().of_copy_channels(d_cvflow.data_ptr() + fm.ncv * 2, 64, L.ptr(d_flow_up), 2, 2, B * h * w,
                                              L.OF_BF16, L.OF_F32, 0, L.stream_ptr()), "of_copy_channels")
        return d_f1, d_f2w, d_flow_up


# ------------------------------------------------------------------ tensor-core parity mode --

def split_bf16(t, pad_to=8, as_f32=False):
    """(hi, lo) with t = hi + lo + O(2^-16 |t|); the last dimension is padded with zeros to a multiple of ``pad_to``
    (TMA needs 16-byte pixel strides).  bf16 tensors (operands), or fp32 tensors holding bf16 values (weights)."""
    assert t.dtype == torch.float32 and t.is_contiguous()
    C = t.shape[-1]
    Cp = _ceil(C, pad_to)
    dt = torch.float32 if as_f32 else torch.bfloat16
    hi = torch.empty(tuple(t.shape[:-1]) + (Cp,), dtype=dt, device=t.device)
    lo = torch.empty_like(hi)
    L.check(L.load().of_split_bf16(L.ptr(t), L.ptr(hi), L.ptr(lo), t.numel() // C, C, Cp, L.dtype_code(hi), L.stream_ptr()),
            "of_split_bf16")
    return hi, lo


class Bf16x3Backend:
    """conv_mode="bf16x3": the fp32 convolutions of the reference (layers.Conv2D, model.py:12-15,104-114;
    resnet/models.py:17-34) on the tensor cores at fp32-class accuracy.  Storage stays fp32 (every non-conv kernel is
    the fp32 mode's); a convolution splits both operands into bf16 hi + lo parts and runs THREE passes of the bf16
    tcgen05 kernels with fp32 accumulation chained through the epilogue's add input:
        x * w  ~  x_hi * w_hi + x_hi * w_lo + x_lo * w_hi        (the dropped lo * lo term is 2^-16 relative)
    - closer to fp32 than a tf32 mode would be (tf32 keeps 10 mantissa bits, hi + lo keeps 16), with the kernels the
    bf16 mode has already validated.  The stem (Cin = 3) and the flow heads (Cout = 2) use the SIMT fp32 kernels."""
    act_dtype = torch.float32

    def __init__(self, net):
        self.net = net

    @staticmethod
    def _tc_ok(layer):
        return layer.cin >= 32 and layer.cout >= 16 and layer.cout % 8 == 0

    @staticmethod
    def _src(t, valid):
        return MultiSource([Source(t, valid, 0)])

    # Splits are shared inside a step: fprop's split of its input and of the weights serves the layer's wgrad / dgrad (same
    # tensors, refreshed by every fprop), wgrad's split of dz serves the dgrad that follows it.
    @staticmethod
    def _cached(layer, slot, t, make):
        c = getattr(layer, slot, None)
        if c is not None and c[0] == t.data_ptr() and c[1] == tuple(t.shape):
            return c[2]
        v = make()
        setattr(layer, slot, (t.data_ptr(), tuple(t.shape), v))
        return v

    def fprop(self, layer, x, w, b):
        if not self._tc_ok(layer):
            return F.conv2d_fprop(x, w, b, stride=layer.stride, act=layer.act, out_dtype=torch.float32)
        layer._x3 = layer._w3 = layer._dz3 = None
        xh, xl = self._cached(layer, "_x3", x, lambda: split_bf16(x))
        wh, wl = self._cached(layer, "_w3", w, lambda: split_bf16(w, pad_to=1, as_f32=True))
        cin = layer.cin
        y = conv2d_fprop_tc(self._src(xh, cin), wh, b, stride=layer.stride, out_dtype=torch.float32)
        y = conv2d_fprop_tc(self._src(xh, cin), wl, None, stride=layer.stride, out_dtype=torch.float32, add=y)
        return conv2d_fprop_tc(self._src(xl, cin), wh, None, stride=layer.stride, act=layer.act, out_dtype=torch.float32, add=y)

    def act_bwd(self, layer, y, dy):
        if layer.act == L.OF_ACT_NONE:
            return dy
        return F.act_bwd(y, dy, layer.act)

    def wgrad(self, layer, x, dz, dw, db):
        if not self._tc_ok(layer):
            F.conv2d_wgrad(x, dz, (layer.k, layer.k), stride=layer.stride, dw=dw, db=db, need_bias_grad=db is not None)
            return
        from .model import _accumulate
        xh, xl = self._cached(layer, "_x3", x, lambda: split_bf16(x))
        zh, zl = self._cached(layer, "_dz3", dz, lambda: split_bf16(dz))
        layer._x3 = None                                   # last use of the input's split
        ks, cin = (layer.k, layer.k), layer.cin
        conv2d_wgrad_tc(self._src(xh, cin), zh, ks, stride=layer.stride, dw=dw, need_bias_grad=False)
        for a, g in ((xh, zl), (xl, zh)):
            part, _ = conv2d_wgrad_tc(self._src(a, cin), g, ks, stride=layer.stride, need_bias_grad=False)
            _accumulate(dw, part)
        if db is not None:
            N, Ho, Wo, Cout = dz.shape
            L.check(L.load().of_colsum(L.ptr(dz), L.ptr(db), N * Ho * Wo, Cout, Cout, L.dtype_code(dz), L.stream_ptr()),
                    "of_colsum")

    def dgrad(self, layer, dz, w, in_hw, chain=None):
        if not self._tc_ok(layer):
            return F.conv2d_dgrad(dz, w, in_hw, stride=layer.stride, out_dtype=torch.float32)
        zh, zl = self._cached(layer, "_dz3", dz, lambda: split_bf16(dz))
        wh, wl = self._cached(layer, "_w3", w, lambda: split_bf16(w, pad_to=1, as_f32=True))
        layer._dz3 = layer._w3 = None                      # last uses
        cin = layer.cin
        from .model import _copy_channels
        out = None
        for c0 in range(0, cin, 256):                    # the kernels write at most 256 output channels per launch
            cnt = min(256, cin - c0)
            cp = _ceil(cnt, 8)
            kw = dict(stride=layer.stride, seg=(c0, cnt), out_dtype=torch.float32, out_channels=cp)
            dx = conv2d_dgrad_tc(zh, wh, in_hw, **kw)
            dx = conv2d_dgrad_tc(zh, wl, in_hw, add=dx, **kw)
            dx = conv2d_dgrad_tc(zl, wh, in_hw, add=dx, **kw)
            if c0 == 0 and cnt == cin and cp == cin:
                return dx
            if out is None:
                out = torch.empty(tuple(dx.shape[:3]) + (cin,), dtype=torch.float32, device=dx.device)
            _copy_channels(dx, 0, out, c0, cnt)
        return out
