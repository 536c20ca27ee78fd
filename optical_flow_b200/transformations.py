"""Mirror of the reference's transformations.py hot functions.

Only ``bilinear_interpolation`` (transformations.py:85-129, with the
``evaluate_tensor_on_xy_grid`` gather of :70-81 folded into the kernel) is on
the hot path; the depth/pose helpers of the same file are out of scope
(SURVEY 2.1).
"""
from __future__ import annotations

from . import functional as F


def bilinear_interpolation(input_tensor, sampling_points, return_indices=False):
    """input_tensor (B,H,W,C) float32/bfloat16 CUDA tensor, sampling_points
    (B,H,W,2) float32 absolute 'xy' coordinates.  Returns (B,H,W,C)."""
    return F.bilinear_sample_fwd(input_tensor, sampling_points, return_indices=return_indices)


def bilinear_interpolation_backward(input_tensor, sampling_points, d_out, need_input_grad=True,
                                    need_points_grad=True):
    """Explicit reverse mode (replaces tf.GradientTape over :85-129)."""
    return F.bilinear_sample_bwd(input_tensor, sampling_points, d_out, need_input_grad, need_points_grad)
