"""Mirror of old/loss.py MyLoss (photometric Charbonnier + flow smoothness)."""
from __future__ import annotations

import torch

from .. import functional as F


def MyLoss(y_true, y_pred, return_grads=False):
    """old/loss.py:7-105.  y_true (B,H,W,6); y_pred = [flows, motions, depths, ego]
    (:24) or just the list of 5 flows [flow6..flow10] coarse -> fine.

    The photometric and smoothness terms run as one fused CUDA launch.  The
    motion term (:96-101) belongs to the superseded depth/ego-motion network
    (old/model.py, out of scope); if motions are passed it is added with a
    plain reduction so the returned value matches the reference's formula.
    Returns the scalar loss (0-d tensor); with return_grads=True also the list
    of d loss / d flow.
    """
    if len(y_pred) == 4 and isinstance(y_pred[0], (list, tuple)):
        flows, motions = y_pred[0], y_pred[1]
    else:
        flows, motions = y_pred, None
    B, H, W, _ = y_true.shape
    assert len(flows) == 5
    for k, f in enumerate(flows):   # old/loss.py:28-32
        div = 16 >> k
        assert tuple(f.shape) == (B, H // div, W // div, 2)
    loss, d_flows = F.loss_charbonnier_smooth(y_true, flows, need_grads=return_grads)
    loss = loss.reshape(())
    if motions is not None:
        extra = sum(torch.sqrt((m.float() ** 2).sum(dim=-1)).sum() for m in motions) / B
        loss = loss + extra
    return (loss, d_flows) if return_grads else loss
