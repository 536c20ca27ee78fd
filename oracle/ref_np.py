"""Literal NumPy restatement of the reference hot path (fp32 unless noted).

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py (pinned to tests/golden/ops.npz: the
reference's own Python run over a TF stand-in); every function cites the reference lines it follows.
All citations are relative to the reference tree (xianlopez/optical_flow).

Layout: NHWC float32, exactly like the reference.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def _f32(a):
    return np.ascontiguousarray(a, dtype=F32)


# ---------------------------------------------------------------------------
# transformations.py:70-129  bilinear_interpolation / evaluate_tensor_on_xy_grid
# ---------------------------------------------------------------------------

def sample_indices(sampling_points, height, width):
    """transformations.py:93-107.  Returns clipped int32 (x0, x1, y0, y1).

    x = pts[...,0] is clipped to [0, width-1], y = pts[...,1] to [0, height-1]
    where (height, width) are the INPUT tensor's sizes.
    """
    pts = _f32(sampling_points)
    x = pts[..., 0]
    y = pts[..., 1]
    x0 = np.floor(x).astype(np.int32)          # :98
    x1 = x0 + 1                                # :99
    y0 = np.floor(y).astype(np.int32)          # :100
    y1 = y0 + 1                                # :101
    x0 = np.clip(x0, 0, width - 1)             # :104
    x1 = np.clip(x1, 0, width - 1)             # :105
    y0 = np.clip(y0, 0, height - 1)            # :106
    y1 = np.clip(y1, 0, height - 1)            # :107
    return x0, x1, y0, y1


def _gather_bxy(inp, x, y):
    """transformations.py:70-81 evaluate_tensor_on_xy_grid (gather_nd on (b,y,x))."""
    b = np.arange(inp.shape[0], dtype=np.int64).reshape(-1, 1, 1)
    return inp[b, y, x]


def bilinear_interpolation(input_tensor, sampling_points, return_indices=False):
    """transformations.py:85-129.

    Weights are computed from the CLIPPED x1 / y1 (:116-125); the four products
    are summed in the order x0y0, x0y1, x1y0, x1y1 (:128-129).
    """
    inp = _f32(input_tensor)
    pts = _f32(sampling_points)
    _, height, width, _ = inp.shape
    assert pts.shape[:3] == inp.shape[:3] and pts.shape[3] == 2
    x = pts[..., 0]
    y = pts[..., 1]
    x0, x1, y0, y1 = sample_indices(pts, height, width)
    v00 = _gather_bxy(inp, x0, y0)             # :110
    v01 = _gather_bxy(inp, x0, y1)             # :111
    v10 = _gather_bxy(inp, x1, y0)             # :112
    v11 = _gather_bxy(inp, x1, y1)             # :113
    x1f = x1.astype(F32)                       # :116
    y1f = y1.astype(F32)                       # :117
    a = (x1f - x).astype(F32)                  # :120  x1minusx
    b = (y1f - y).astype(F32)                  # :121  y1minusy
    one = F32(1.0)
    w00 = (a * b)[..., None]                   # :122
    w01 = (a * (one - b))[..., None]           # :123
    w10 = ((one - a) * b)[..., None]           # :124
    w11 = ((one - a) * (one - b))[..., None]   # :125
    out = w00 * v00
    out = out + w01 * v01
    out = out + w10 * v10
    out = out + w11 * v11                      # :128-129 accumulate_n
    out = out.astype(F32)
    if return_indices:
        return out, (x0, x1, y0, y1)
    return out


def bilinear_interpolation_bwd(input_tensor, sampling_points, d_out, return_abs=False):
    """Reverse mode of transformations.py:85-129 as TF autodiff derives it.

    return_abs=True additionally returns sum |w_k| * |d_out| scattered the same way: the
    magnitude an fp32 scatter-add (TF's scatter_nd or CUDA atomics, any order) rounds
    against.  Clipped taps have weights far outside [0,1] that cancel, so
    this - not max|d_input| - is the scale of the legitimate fp32 reordering error.

    * d_input: scatter-add (gather_nd gradient) of w_k * d_out at the 4 taps.
    * d_points: only through a = x1c - x and b = y1c - y (floor/cast/clip carry
      no gradient): d/dx = -d/da, d/dy = -d/db.
    Accumulated in float64 then rounded, so it is the "exact" reference for
    the atomics-based CUDA backward.
    """
    inp = _f32(input_tensor)
    pts = _f32(sampling_points)
    g = np.asarray(d_out, dtype=np.float64)
    B, H, W, C = inp.shape
    x = pts[..., 0].astype(np.float64)
    y = pts[..., 1].astype(np.float64)
    x0, x1, y0, y1 = sample_indices(pts, H, W)
    # a, b are fp32 quantities in the forward pass
    a = (x1.astype(F32) - pts[..., 0]).astype(np.float64)
    b = (y1.astype(F32) - pts[..., 1]).astype(np.float64)
    inp64 = inp.astype(np.float64)
    v00 = _gather_bxy(inp64, x0, y0)
    v01 = _gather_bxy(inp64, x0, y1)
    v10 = _gather_bxy(inp64, x1, y0)
    v11 = _gather_bxy(inp64, x1, y1)
    d_a = np.sum(g * (b[..., None] * (v00 - v10) + (1 - b)[..., None] * (v01 - v11)), axis=-1)
    d_b = np.sum(g * (a[..., None] * (v00 - v01) + (1 - a)[..., None] * (v10 - v11)), axis=-1)
    d_pts = np.stack([-d_a, -d_b], axis=-1)
    d_inp = np.zeros((B * H * W, C), dtype=np.float64)
    bidx = np.arange(B, dtype=np.int64).reshape(-1, 1, 1) * (H * W)
    for xx, yy, w in ((x0, y0, a * b), (x0, y1, a * (1 - b)),
                      (x1, y0, (1 - a) * b), (x1, y1, (1 - a) * (1 - b))):
        flat = (bidx + yy.astype(np.int64) * W + xx).reshape(-1)
        np.add.at(d_inp, flat, (w[..., None] * g).reshape(-1, C))
    if return_abs:
        d_abs = np.zeros((B * H * W, C), dtype=np.float64)
        for xx, yy, w in ((x0, y0, a * b), (x0, y1, a * (1 - b)),
                          (x1, y0, (1 - a) * b), (x1, y1, (1 - a) * (1 - b))):
            flat = (bidx + yy.astype(np.int64) * W + xx).reshape(-1)
            np.add.at(d_abs, flat, (np.abs(w)[..., None] * np.abs(g)).reshape(-1, C))
        return d_inp.reshape(B, H, W, C).astype(F32), d_pts.astype(F32), d_abs.reshape(B, H, W, C)
    return d_inp.reshape(B, H, W, C).astype(F32), d_pts.astype(F32)


# ---------------------------------------------------------------------------
# model.py:55-73  warp_features (row/col swap quirk)
# ---------------------------------------------------------------------------

def warp_grid(height, width):
    """model.py:65-69: meshgrid(range(h), range(w), 'ij') stacked as [row, col]."""
    ii, jj = np.meshgrid(np.arange(height), np.arange(width), indexing="ij")
    return np.stack([ii, jj], axis=-1).astype(F32)[None]


def warp_features(flow_1_to_2, features2, return_indices=False):
    """model.py:55-73.  sampling point = [row, col] + flow, handed to
    bilinear_interpolation which reads channel 0 as x and channel 1 as y."""
    flow = _f32(flow_1_to_2)
    feats = _f32(features2)
    _, height, width, _ = feats.shape
    assert flow.shape[1] == height and flow.shape[2] == width and flow.shape[3] == 2
    pts = (warp_grid(height, width) + flow).astype(F32)   # :71 fp32 add
    return bilinear_interpolation(feats, pts, return_indices=return_indices)


def warp_features_bwd(flow_1_to_2, features2, d_out, return_abs=False):
    flow = _f32(flow_1_to_2)
    feats = _f32(features2)
    _, height, width, _ = feats.shape
    pts = (warp_grid(height, width) + flow).astype(F32)
    return bilinear_interpolation_bwd(feats, pts, d_out, return_abs=return_abs)  # d_flow == d_pts


def warp_xy(flow_xy, features2, return_indices=False):
    """The corrected 'XY' mode: sampling point = [col, row] + flow (x,y order),
    same sampler.  Not in the reference's current model; offered as a mode."""
    flow = _f32(flow_xy)
    feats = _f32(features2)
    _, height, width, _ = feats.shape
    g = warp_grid(height, width)[..., ::-1]
    pts = (np.ascontiguousarray(g) + flow).astype(F32)
    return bilinear_interpolation(feats, pts, return_indices=return_indices)


def warp_xy_bwd(flow_xy, features2, d_out, return_abs=False):
    flow = _f32(flow_xy)
    feats = _f32(features2)
    _, height, width, _ = feats.shape
    g = np.ascontiguousarray(warp_grid(height, width)[..., ::-1])
    pts = (g + flow).astype(F32)
    return bilinear_interpolation_bwd(feats, pts, d_out, return_abs=return_abs)


# ---------------------------------------------------------------------------
# old/image_warp.py:7-78  UnFlow backward warp
# ---------------------------------------------------------------------------

def image_warp_indices(flow, height, width):
    """old/image_warp.py:28-56: int(floor(flow)) + pos, each tap clamped."""
    flow = _f32(flow)
    ff = np.floor(flow)
    fl = ff.astype(np.int32)                                   # :28
    wts = (flow - ff).astype(F32)                              # :29
    pos_x = np.arange(width, dtype=np.int32).reshape(1, 1, -1)  # :32
    pos_y = np.arange(height, dtype=np.int32).reshape(1, -1, 1)  # :33-34
    x0 = pos_x + fl[..., 0]                                    # :48
    x1 = x0 + 1
    y0 = pos_y + fl[..., 1]                                    # :50
    y1 = y0 + 1
    x0 = np.clip(x0, 0, width - 1)                             # :53-56
    x1 = np.clip(x1, 0, width - 1)
    y0 = np.clip(y0, 0, height - 1)
    y1 = np.clip(y1, 0, height - 1)
    return (x0, x1, y0, y1), wts


def image_warp(im, flow, return_indices=False):
    """old/image_warp.py:7-78 (== image_warp_tf :81-153)."""
    im = _f32(im)
    B, H, W, C = im.shape
    (x0, x1, y0, y1), wts = image_warp_indices(flow, H, W)
    xw = wts[..., 0]
    yw = wts[..., 1]
    one = F32(1.0)
    wa = ((one - xw) * (one - yw))[..., None]   # :43 top-left
    wb = ((one - xw) * yw)[..., None]           # :44 bottom-left
    wc = (xw * (one - yw))[..., None]           # :45 top-right
    wd = (xw * yw)[..., None]                   # :46 bottom-right
    Ia = _gather_bxy(im, x0, y0)                # idx_a = base_y0 + x0
    Ib = _gather_bxy(im, x0, y1)                # idx_b = base_y1 + x0
    Ic = _gather_bxy(im, x1, y0)
    Id = _gather_bxy(im, x1, y1)
    out = wa * Ia
    out = out + wb * Ib
    out = out + wc * Ic
    out = out + wd * Id                         # :75 add_n
    out = out.astype(F32)
    if return_indices:
        return out, (x0, x1, y0, y1)
    return out


def image_warp_bwd(im, flow, d_out):
    """Reverse mode of old/image_warp.py: d_flow only through the fractional
    weights (d(flow - floor(flow))/dflow = 1), d_im by 4-tap scatter-add."""
    im = _f32(im)
    B, H, W, C = im.shape
    g = np.asarray(d_out, dtype=np.float64)
    (x0, x1, y0, y1), wts = image_warp_indices(flow, H, W)
    xw = wts[..., 0].astype(np.float64)
    yw = wts[..., 1].astype(np.float64)
    im64 = im.astype(np.float64)
    Ia = _gather_bxy(im64, x0, y0)
    Ib = _gather_bxy(im64, x0, y1)
    Ic = _gather_bxy(im64, x1, y0)
    Id = _gather_bxy(im64, x1, y1)
    d_xw = np.sum(g * ((1 - yw)[..., None] * (Ic - Ia) + yw[..., None] * (Id - Ib)), axis=-1)
    d_yw = np.sum(g * ((1 - xw)[..., None] * (Ib - Ia) + xw[..., None] * (Id - Ic)), axis=-1)
    d_flow = np.stack([d_xw, d_yw], axis=-1)
    d_im = np.zeros((B * H * W, C), dtype=np.float64)
    bidx = np.arange(B, dtype=np.int64).reshape(-1, 1, 1) * (H * W)
    for xx, yy, w in ((x0, y0, (1 - xw) * (1 - yw)), (x0, y1, (1 - xw) * yw),
                      (x1, y0, xw * (1 - yw)), (x1, y1, xw * yw)):
        flat = (bidx + yy.astype(np.int64) * W + xx).reshape(-1)
        np.add.at(d_im, flat, (w[..., None] * g).reshape(-1, C))
    return d_im.reshape(B, H, W, C).astype(F32), d_flow.astype(F32)
