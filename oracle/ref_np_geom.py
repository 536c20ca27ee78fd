"""NumPy restatement, part 3: the depth / pose re-projection helpers of transformations.py
(BackprojectLayer :12-30, ProjectLayer :33-44, WarpLayer :47-58, transform3d :61-67,
rotation_from_axisangle :133-173, make_transformation_matrix :176-191), forward plus hand-derived
reverse mode.

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py.  Pinned to tests/golden/geom.npz (the reference's own
transformations.py executed over the torch-backed TF stand-in, tests/golden/make_golden_geom.py).
"""
from __future__ import annotations

import numpy as np

from . import ref_np as R

F32 = np.float32


def _rays(kinv, h, w):
    # :16-20  meshgrid(x, y) -> [X, Y, 1]; rays = matvec(Kinv, .)
    X, Y = np.meshgrid(np.arange(w, dtype=F32), np.arange(h, dtype=F32))
    hom = np.stack([X, Y, np.ones_like(X)], axis=-1)
    return np.einsum("rc,hwc->hwr", np.asarray(kinv, F32), hom).astype(F32)


def backproject(depth, K):
    """:24-30  depth (B,h,w,1) -> (B,h,w,4)."""
    B, h, w, _ = depth.shape
    rays = _rays(np.linalg.inv(np.asarray(K)).astype(F32), h, w)
    pts = depth.astype(F32) * rays[None]
    return np.concatenate([pts, np.ones((B, h, w, 1), F32)], axis=-1)


def backproject_bwd(d_points, K):
    B, h, w, _ = d_points.shape
    rays = _rays(np.linalg.inv(np.asarray(K)).astype(F32), h, w)
    return (d_points[..., :3] * rays[None]).sum(-1, keepdims=True).astype(F32)


def project(points, K):
    """:39-44."""
    K_ext = np.concatenate([np.asarray(K, F32), np.zeros((3, 1), F32)], axis=1)
    hom = np.einsum("rc,bhwc->bhwr", K_ext, points.astype(F32)).astype(F32)
    return (hom[..., :2] / (hom[..., 2:3] + F32(1e-7))).astype(F32)


def project_bwd(points, K, d_pix):
    K_ext = np.concatenate([np.asarray(K, F32), np.zeros((3, 1), F32)], axis=1).astype(np.float64)
    hom = np.einsum("rc,bhwc->bhwr", K_ext, points.astype(np.float64))
    den = hom[..., 2:3] + 1e-7
    g = d_pix.astype(np.float64)
    d_hom = np.concatenate([g / den, -(g * hom[..., :2]).sum(-1, keepdims=True) / den ** 2], axis=-1)
    return np.einsum("rc,bhwr->bhwc", K_ext, d_hom).astype(F32)


def transform3d(T, points):
    """:61-67."""
    return np.einsum("brc,bhwc->bhwr", T.astype(F32), points.astype(F32)).astype(F32)


def transform3d_bwd(T, points, d_out):
    g = d_out.astype(np.float64)
    d_points = np.einsum("brc,bhwr->bhwc", T.astype(np.float64), g).astype(F32)
    d_T = np.einsum("bhwr,bhwc->brc", g, points.astype(np.float64)).astype(F32)
    return d_points, d_T


def rotation_from_axisangle(axisangle):
    """:133-173 (literal)."""
    a = axisangle.astype(F32)
    theta = np.sqrt((a * a).sum(1, keepdims=True)).astype(F32)
    axis = a / (theta + F32(1e-7))
    theta = theta[:, 0]
    c, s = np.cos(theta), np.sin(theta)
    C = F32(1.0) - c
    x, y, z = axis[:, 0], axis[:, 1], axis[:, 2]
    xs, ys, zs = x * s, y * s, z * s
    xC, yC, zC = x * C, y * C, z * C
    xyC, yzC, zxC = x * yC, y * zC, z * xC
    col0 = np.stack([x * xC + c, xyC + zs, zxC - ys], axis=1)
    col1 = np.stack([xyC - zs, y * yC + c, yzC + xs], axis=1)
    col2 = np.stack([zxC + ys, yzC - xs, z * zC + c], axis=1)
    return np.stack([col0, col1, col2], axis=2).astype(F32)


def make_transformation_matrix(raw, invert):
    """:176-191."""
    raw = raw.astype(F32)
    rot = rotation_from_axisangle(raw[:, :3])
    t = raw[:, 3:]
    B = raw.shape[0]
    if invert:
        rot = rot.transpose(0, 2, 1)
        t = -np.einsum("brc,bc->br", rot, t)
    aux1 = np.concatenate([rot, np.zeros((B, 1, 3), F32)], axis=1)
    aux2 = np.concatenate([t, np.ones((B, 1), F32)], axis=1)
    return np.concatenate([aux1, aux2[:, :, None]], axis=2).astype(F32)


def make_transformation_matrix_bwd(raw, invert, d_out, eps=1e-6):
    """Reverse mode by fp64 central differences of the literal forward (6 inputs per sample: exact enough
    to check a hand-derived kernel at 1e-5, and independent of it)."""
    raw64 = raw.astype(np.float64)

    def fwd(r):
        a, t = r[:, :3], r[:, 3:]
        th = np.sqrt((a * a).sum(1, keepdims=True))
        ax = a / (th + 1e-7)
        th = th[:, 0]
        c, s = np.cos(th), np.sin(th)
        C = 1.0 - c
        x, y, z = ax[:, 0], ax[:, 1], ax[:, 2]
        rot = np.stack([np.stack([x * x * C + c, x * y * C - z * s, z * x * C + y * s], 1),
                        np.stack([x * y * C + z * s, y * y * C + c, y * z * C - x * s], 1),
                        np.stack([z * x * C - y * s, y * z * C + x * s, z * z * C + c], 1)], 1)
        if invert:
            rot = rot.transpose(0, 2, 1)
            t = -np.einsum("brc,bc->br", rot, t)
        m = np.zeros((r.shape[0], 4, 4))
        m[:, :3, :3] = rot
        m[:, :3, 3] = t
        m[:, 3, 3] = 1.0
        return m

    g = d_out.astype(np.float64)
    d_raw = np.zeros_like(raw64)
    for k in range(6):
        e = np.zeros_like(raw64)
        e[:, k] = eps
        d_raw[:, k] = ((fwd(raw64 + e) - fwd(raw64 - e)) * g).sum((1, 2)) / (2 * eps)
    return d_raw.astype(F32)


def warp_layer(image_source, points3d_target, T, K):
    """:52-58 -> (image_target, pix_coords_source)."""
    pix = project(transform3d(T, points3d_target), K)
    return R.bilinear_interpolation(image_source, pix), pix
