"""Mirror of the reference's loss.py."""
from __future__ import annotations

from . import functional as F


class LossLayer:
    """loss.py:5-32.  ``LossLayer()(batch_imgs, flows)`` -> scalar loss.

    One fused CUDA launch resizes, warps (model.py:55-73 semantics), reduces
    and - unlike the reference, where tf.GradientTape does it - also produces
    d loss / d flow for every scale; the gradients of the last call are kept in
    ``self.d_flows`` for the explicit backward pass.
    """

    def __init__(self, warp_mode="ref_ij"):
        self.warp_mode = warp_mode
        self.d_flows = None

    def __call__(self, batch_imgs, flows, need_grads=True):
        num_scales = len(flows)
        for scale_idx in range(num_scales):
            scaled_height = int(batch_imgs.shape[1] / (2.0 ** (scale_idx + 1)))   # loss.py:11
            scaled_width = int(batch_imgs.shape[2] / (2.0 ** (scale_idx + 1)))    # loss.py:12
            flow = flows[scale_idx]
            assert flow.shape[1] == scaled_height   # loss.py:14
            assert flow.shape[2] == scaled_width    # loss.py:15
        loss, self.d_flows = F.loss_l1_pyramid(batch_imgs, flows, need_grads=need_grads, mode=self.warp_mode)
        return loss.reshape(())
