"""Mirror of the reference's transformations.py: same names, argument order and shapes.

``bilinear_interpolation`` (transformations.py:85-129, with the ``evaluate_tensor_on_xy_grid`` gather of
:70-81 folded into the kernel) is on the training hot path.  The depth / pose helpers of the same file
(``BackprojectLayer`` :12-30, ``ProjectLayer`` :33-44, ``WarpLayer`` :47-58, ``transform3d`` :61-67,
``rotation_from_axisangle`` :133-173, ``make_transformation_matrix`` :176-191) are SURVEY 8f rank 3: they
reuse the same sampler kernel behind a fused "transform + project" coordinate kernel (csrc/geometry.cu).
TensorFlow's tape is replaced by explicit ``backward`` methods that call the C ABI in reverse order.
"""
from __future__ import annotations

import numpy as np
import torch

from . import functional as F


def concat_images(imgs1, imgs2):
    """transformations.py:5-9 (a buffer concat, no arithmetic)."""
    return torch.cat([imgs1, imgs2], dim=-1)


def bilinear_interpolation(input_tensor, sampling_points, return_indices=False):
    """input_tensor (B,H,W,C) float32/bfloat16 CUDA tensor, sampling_points
    (B,H,W,2) float32 absolute 'xy' coordinates.  Returns (B,H,W,C)."""
    return F.bilinear_sample_fwd(input_tensor, sampling_points, return_indices=return_indices)


def bilinear_interpolation_backward(input_tensor, sampling_points, d_out, need_input_grad=True,
                                    need_points_grad=True):
    """Explicit reverse mode (replaces tf.GradientTape over :85-129)."""
    return F.bilinear_sample_bwd(input_tensor, sampling_points, d_out, need_input_grad, need_points_grad)


def _intrinsics(K):
    if isinstance(K, torch.Tensor):
        K = K.detach().cpu().numpy()
    K = np.asarray(K)
    assert K.shape == (3, 3), "K: intrinsics matrix (3, 3)"
    return K


class BackprojectLayer:
    """transformations.py:12-30.  The (batch_size, h, w, 3) ray tensor of :21 is never materialised: the
    kernel rebuilds Kinv @ [x, y, 1] per pixel."""

    def __init__(self, K, height, width, batch_size):
        self.Kinv = np.linalg.inv(_intrinsics(K)).astype(np.float32)   # :15
        self.shape = (batch_size, height, width, 1)

    def __call__(self, depth):
        assert tuple(depth.shape) == self.shape                          # :26
        return F.backproject_fwd(depth, self.Kinv)

    def backward(self, d_points3d_hom):
        return F.backproject_bwd(d_points3d_hom, self.Kinv)


class ProjectLayer:
    """transformations.py:33-44 (the tiled (batch_size, h, w, 3, 4) K_ext of :38 is a kernel constant)."""

    def __init__(self, K, height, width, batch_size):
        self.K = _intrinsics(K).astype(np.float32)
        self.shape = (batch_size, height, width, 4)

    def __call__(self, points3d_hom):
        assert tuple(points3d_hom.shape) == self.shape
        return F.project_fwd(points3d_hom, self.K)

    def backward(self, points3d_hom, d_pixel_coords):
        return F.project_bwd(points3d_hom, self.K, d_pixel_coords)


class WarpLayer:
    """transformations.py:47-58: image_target = bilinear_interpolation(image_source,
    project(transform3d(T_source_target, points3d_target))).  Two launches: coordinates, sampler."""

    def __init__(self, K, height, width, batch_size):
        self.project = ProjectLayer(K, height, width, batch_size)
        self._saved = None

    def __call__(self, image_source, points3d_target, T_source_target):
        pix = F.view_coords_fwd(self.project.K, T_source_target, points3d_target)
        self._saved = (image_source, points3d_target, T_source_target, pix)
        return F.bilinear_sample_fwd(image_source, pix)

    def backward(self, d_image_target, need_image_grad=True):
        """-> (d_image_source, d_points3d_target, d_T_source_target)."""
        image_source, points3d_target, T, pix = self._saved
        d_image, d_pix = F.bilinear_sample_bwd(image_source, pix, d_image_target, need_image_grad, True)
        d_points, d_T = F.view_coords_bwd(self.project.K, T, points3d_target, d_pix)
        return d_image, d_points, d_T


def transform3d(T21, points3d_hom_1):
    """transformations.py:61-67."""
    return F.transform3d_fwd(T21, points3d_hom_1)


def transform3d_backward(T21, points3d_hom_1, d_points3d_hom_2):
    return F.transform3d_bwd(T21, points3d_hom_1, d_points3d_hom_2)


def make_transformation_matrix(raw_transformation, invert):
    """transformations.py:176-191: (batch_size, 6) axis-angle + translation -> (batch_size, 4, 4)."""
    return F.make_transformation_matrix_fwd(raw_transformation, invert)


def make_transformation_matrix_backward(raw_transformation, invert, d_matrix):
    return F.make_transformation_matrix_bwd(raw_transformation, invert, d_matrix)


def rotation_from_axisangle(axisangle):
    """transformations.py:133-173: the rotation block of make_transformation_matrix with zero translation."""
    raw = torch.cat([axisangle, torch.zeros_like(axisangle)], dim=1).contiguous()
    return F.make_transformation_matrix_fwd(raw, False)[:, :3, :3].contiguous()
