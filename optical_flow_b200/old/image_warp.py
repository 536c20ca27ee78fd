"""Mirror of old/image_warp.py (UnFlow backward warp)."""
from __future__ import annotations

from .. import functional as F


def image_warp(im, flow, return_indices=False):
    """old/image_warp.py:7-78.  im (B,H,W,C) f32, flow (B,H,W,2) f32 (x, y)."""
    return F.image_warp_fwd(im, flow, return_indices=return_indices)


def image_warp_tf(im, flow, return_indices=False):
    """old/image_warp.py:81-153: identical math with static shapes."""
    return F.image_warp_fwd(im, flow, return_indices=return_indices)


def image_warp_backward(im, flow, d_out, need_im_grad=True, need_flow_grad=True):
    return F.image_warp_bwd(im, flow, d_out, need_im_grad, need_flow_grad)
