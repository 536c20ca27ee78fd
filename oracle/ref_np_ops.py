"""NumPy restatement, part 2: resize, flow upscale, cost volume, the two
pyramid losses (forward + hand-derived backward), Adam, TF SAME padding.

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py (pinned to tests/golden/ops.npz).
"""
from __future__ import annotations

import numpy as np

from . import ref_np as R

F32 = np.float32


# ---------------------------------------------------------------------------
# tf.image.resize (bilinear, half_pixel_centers=True, antialias=False)
# used by model.py:77, loss.py:18, old/loss.py:15-22
# ---------------------------------------------------------------------------

def resize_coeffs(out_size, in_size):
    """TF compute_interpolation_weights with half-pixel centres:
    in = (o + 0.5) * scale - 0.5; lower = max(floor(in), 0);
    upper = min(ceil(in), in_size - 1); lerp = in - floor(in)."""
    scale = F32(in_size) / F32(out_size)
    o = np.arange(out_size, dtype=F32)
    src = ((o + F32(0.5)) * scale - F32(0.5)).astype(F32)
    fl = np.floor(src)
    lower = np.maximum(fl, 0).astype(np.int32)
    upper = np.minimum(np.ceil(src), in_size - 1).astype(np.int32)
    lerp = (src - fl).astype(F32)
    return lower, upper, lerp


def resize_bilinear(x, out_h, out_w):
    x = np.ascontiguousarray(x, dtype=F32)
    B, H, W, C = x.shape
    if out_h == H and out_w == W:
        return x.copy()
    ylo, yhi, yl = resize_coeffs(out_h, H)
    xlo, xhi, xl = resize_coeffs(out_w, W)
    top = x[:, ylo]            # (B, oh, W, C)
    bot = x[:, yhi]
    xl_ = xl.reshape(1, 1, -1, 1)
    yl_ = yl.reshape(1, -1, 1, 1)
    tl, tr = top[:, :, xlo], top[:, :, xhi]
    bl, br = bot[:, :, xlo], bot[:, :, xhi]
    t = tl + (tr - tl) * xl_
    b = bl + (br - bl) * xl_
    return (t + (b - t) * yl_).astype(F32)


def resize_bilinear_bwd(d_out, in_h, in_w):
    """Adjoint of resize_bilinear w.r.t. its input (float64 accumulate)."""
    g = np.asarray(d_out, dtype=np.float64)
    B, oh, ow, C = g.shape
    ylo, yhi, yl = resize_coeffs(oh, in_h)
    xlo, xhi, xl = resize_coeffs(ow, in_w)
    yl = yl.astype(np.float64).reshape(1, -1, 1, 1)
    xl = xl.astype(np.float64).reshape(1, 1, -1, 1)
    d_in = np.zeros((B, in_h, in_w, C), dtype=np.float64)
    for yy, wy in ((ylo, 1 - yl), (yhi, yl)):
        for xx, wx in ((xlo, 1 - xl), (xhi, xl)):
            contrib = g * wy * wx
            np.add.at(d_in, (slice(None), yy[:, None], xx[None, :]), contrib)
    return d_in.astype(F32)


def upscale_flow(flow):
    """model.py:76-77."""
    flow = np.ascontiguousarray(flow, dtype=F32)
    return (resize_bilinear(flow, flow.shape[1] * 2, flow.shape[2] * 2) * F32(2.0)).astype(F32)


def upscale_flow_bwd(d_up):
    d_up = np.asarray(d_up, dtype=F32)
    return resize_bilinear_bwd(d_up * F32(2.0), d_up.shape[1] // 2, d_up.shape[2] // 2)


# ---------------------------------------------------------------------------
# model.py:29-42 create_cost_volume
# ---------------------------------------------------------------------------

def create_cost_volume(features1, features2, max_disp):
    f1 = np.ascontiguousarray(features1, dtype=F32)
    f2 = np.ascontiguousarray(features2, dtype=F32)
    assert f1.shape == f2.shape
    _, height, width, _ = f1.shape
    f2p = np.pad(f2, [[0, 0], [max_disp, max_disp], [max_disp, max_disp], [0, 0]])  # :34
    corr = []
    for i in range(2 * max_disp + 1):           # :37
        for j in range(2 * max_disp + 1):       # :38
            corr.append(np.sum(f1 * f2p[:, i:i + height, j:j + width, :], axis=-1, dtype=F32))
    return np.stack(corr, axis=-1).astype(F32)  # :40


def create_cost_volume_bwd(features1, features2, max_disp, d_cv):
    f1 = np.asarray(features1, dtype=np.float64)
    f2 = np.asarray(features2, dtype=np.float64)
    g = np.asarray(d_cv, dtype=np.float64)
    _, height, width, _ = f1.shape
    m = max_disp
    f2p = np.pad(f2, [[0, 0], [m, m], [m, m], [0, 0]])
    d_f1 = np.zeros_like(f1)
    d_f2p = np.zeros_like(f2p)
    k = 0
    for i in range(2 * m + 1):
        for j in range(2 * m + 1):
            gk = g[..., k:k + 1]
            d_f1 += gk * f2p[:, i:i + height, j:j + width, :]
            d_f2p[:, i:i + height, j:j + width, :] += gk * f1
            k += 1
    d_f2 = d_f2p[:, m:m + height, m:m + width, :]
    return d_f1.astype(F32), d_f2.astype(F32)


# ---------------------------------------------------------------------------
# loss.py:5-32 LossLayer
# ---------------------------------------------------------------------------

def loss_layer(batch_imgs, flows, return_grads=False, return_diffs=False):
    """loss.py:5-32.  Returns loss (float32 scalar) and optionally d_flows (and, with return_diffs, the per-scale
    photometric differences im1 - warped: |diff| ~ 1 ulp marks the pixels whose sign(diff) - hence whose whole
    d_flow - is an fp32 tie-break, which large-shape tests exclude from max-abs bounds).

    Forward in fp32 like the reference; the mean/sum and the backward are
    accumulated in float64 (the "exact" value the CUDA reduction is held to)."""
    imgs = np.ascontiguousarray(batch_imgs, dtype=F32)
    num_scales = len(flows)
    loss = 0.0
    d_flows = []
    diffs = []
    for s in range(num_scales):
        sh = int(imgs.shape[1] / (2.0 ** (s + 1)))   # :11
        sw = int(imgs.shape[2] / (2.0 ** (s + 1)))   # :12
        flow = np.ascontiguousarray(flows[s], dtype=F32)
        assert flow.shape[1] == sh and flow.shape[2] == sw  # :14-15
        resized = resize_bilinear(imgs, sh, sw)      # :18 (always from full res)
        im1 = resized[..., :3]
        im2 = np.ascontiguousarray(resized[..., 3:])
        warped = R.warp_features(flow, im2)          # :26
        diff = (im1 - warped).astype(F32)
        n = diff.size
        diffs.append(diff)
        loss += float(np.sum(np.abs(diff), dtype=np.float64)) / n   # :28 reduce_mean
        if return_grads:
            # d|d|/dwarped = -sign(diff) (tf.abs grad = sign, sign(0)=0); /n mean; /num_scales (:30)
            d_warped = (-np.sign(diff).astype(np.float64) / n / num_scales)
            _, d_flow = R.warp_features_bwd(flow, im2, d_warped)
            d_flows.append(d_flow)
    loss = F32(loss / float(num_scales))             # :30
    if return_grads and return_diffs:
        return loss, d_flows, diffs
    if return_grads:
        return loss, d_flows
    return loss


# ---------------------------------------------------------------------------
# old/loss.py:7-105 MyLoss (photometric Charbonnier + smoothness [+ motion])
# ---------------------------------------------------------------------------

def my_loss(y_true, flows_coarse_to_fine, motions=None, return_grads=False):
    """old/loss.py:7-105.  flows_coarse_to_fine = [flow6 (H/16), flow7 (H/8),
    flow8 (H/4), flow9 (H/2), flow10 (H)] exactly like y_pred[0] (:24-25).
    Warps image 1 by the flow and compares with image 2 (:40-68)."""
    y = np.ascontiguousarray(y_true, dtype=F32)
    im1 = np.ascontiguousarray(y[..., :3])
    im2 = np.ascontiguousarray(y[..., 3:])
    B, H, W, _ = im1.shape
    eps = 1e-4
    photo = 0.0
    smooth = 0.0
    d_flows = []
    for k, flow in enumerate(flows_coarse_to_fine):
        div = 16 >> k                                 # 16, 8, 4, 2, 1
        h, w = H // div, W // div                     # :15-22
        flow = np.ascontiguousarray(flow, dtype=F32)
        assert flow.shape == (B, h, w, 2)             # :28-32
        a = resize_bilinear(im1, h, w) if div > 1 else im1
        b = resize_bilinear(im2, h, w) if div > 1 else im2
        warped = R.image_warp(a, flow)                # :40 etc.
        diff = (warped - b).astype(F32)               # :41
        sq = np.sum(np.square(diff), axis=-1, dtype=F32)   # :42
        err = np.sqrt(sq + F32(eps)).astype(F32)      # :43
        photo += float(np.sum(err, dtype=np.float64))  # :44
        di = (flow[:, :-1] - flow[:, 1:]).astype(F32)      # :71
        dj = (flow[:, :, :-1] - flow[:, :, 1:]).astype(F32)  # :72
        si = np.sqrt(np.square(di) + F32(eps)).astype(F32)
        sj = np.sqrt(np.square(dj) + F32(eps)).astype(F32)
        smooth += float(np.sum(si, dtype=np.float64)) + float(np.sum(sj, dtype=np.float64))  # :73-74
        if return_grads:
            d_warped = diff.astype(np.float64) / err.astype(np.float64)[..., None] / B
            _, d_flow = R.image_warp_bwd(a, flow, d_warped)
            d_flow = d_flow.astype(np.float64)
            gi = 0.2 * di.astype(np.float64) / si.astype(np.float64) / B
            gj = 0.2 * dj.astype(np.float64) / sj.astype(np.float64) / B
            d_flow[:, :-1] += gi
            d_flow[:, 1:] -= gi
            d_flow[:, :, :-1] += gj
            d_flow[:, :, 1:] -= gj
            d_flows.append(d_flow.astype(F32))
    motion = 0.0
    if motions is not None:
        for mo in motions:                            # :97-101 (sqrt without eps)
            mo = np.asarray(mo, dtype=F32)
            motion += float(np.sum(np.sqrt(np.sum(np.square(mo), axis=-1, dtype=F32)), dtype=np.float64))
    loss = F32((photo + 0.2 * smooth + motion) / B)   # :103-105
    if return_grads:
        return loss, d_flows
    return loss


# ---------------------------------------------------------------------------
# train.py:34,56  tf.keras.optimizers.Adam(1e-4): beta 0.9/0.999, eps 1e-7
# ---------------------------------------------------------------------------

def adam_step(param, grad, m, v, step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-7):
    """Keras/TF2 Adam (non-amsgrad): lr_t = lr*sqrt(1-b2^t)/(1-b1^t);
    m += (g-m)(1-b1); v += (g^2-v)(1-b2); p -= lr_t*m/(sqrt(v)+eps).
    ``step`` is the 1-based iteration count t.  float32 state like TF."""
    p = np.asarray(param, dtype=F32).copy()
    g = np.asarray(grad, dtype=F32)
    m = np.asarray(m, dtype=F32).copy()
    v = np.asarray(v, dtype=F32).copy()
    lr_t = F32(lr * np.sqrt(1.0 - beta2 ** step) / (1.0 - beta1 ** step))
    # TF's ApplyAdam kernel forms (1 - beta) in the variable dtype: T(1) - T(beta) in fp32
    m += (g - m) * (F32(1.0) - F32(beta1))
    v += (np.square(g) - v) * (F32(1.0) - F32(beta2))
    p -= lr_t * m / (np.sqrt(v) + F32(eps))
    return p, m, v


# ---------------------------------------------------------------------------
# Keras Conv2D / MaxPool padding='same' arithmetic
# ---------------------------------------------------------------------------

def tf_same_pad(in_size, k, s):
    """Returns (out_size, pad_lo, pad_hi) as TF computes 'SAME'."""
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    lo = total // 2
    return out, lo, total - lo
