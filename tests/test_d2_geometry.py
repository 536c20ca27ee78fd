"""Depth / pose re-projection helpers (transformations.py:12-67,133-191): the oracle against
vectors produced by the reference's own transformations.py (CPU), the CUDA kernels against both (GPU).
Tolerance: 1e-5 relative to the tensor's max (fp32; the reference's matvec summation order is BLAS-defined)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_np as R, ref_np_geom as Gm  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "geom.npz"))
TOL = 1e-5


def close(a, b, tol=TOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    err = np.abs(a - b).max()
    assert err <= tol * max(np.abs(b).max(), 1e-30), (err, np.abs(b).max())


# ------------------------------------------------------------------ oracle vs the reference's own run --
@pytest.mark.parametrize("invert", [False, True])
def test_oracle_pose_matrix(invert):
    tag = "inv" if invert else "fwd"
    close(Gm.make_transformation_matrix(G["raw"], invert), G["M_" + tag])
    close(Gm.make_transformation_matrix_bwd(G["raw"], invert, G["gM_" + tag]), G["d_raw_" + tag], 1e-4)
    close(Gm.rotation_from_axisangle(G["raw"][:, :3]), G["R"])


def test_oracle_pose_properties():
    raw = np.random.default_rng(1).standard_normal((5, 6)).astype(np.float32)
    M, Mi = Gm.make_transformation_matrix(raw, False), Gm.make_transformation_matrix(raw, True)
    eye = np.broadcast_to(np.eye(4, dtype=np.float32), M.shape)
    assert np.abs(M @ Mi - eye).max() < 1e-5                       # invert=True is the inverse rigid transform
    Rm = M[:, :3, :3]
    assert np.abs(Rm @ Rm.transpose(0, 2, 1) - eye[:, :3, :3]).max() < 1e-5
    zero = Gm.make_transformation_matrix(np.zeros((1, 6), np.float32), False)
    assert np.array_equal(zero[0], np.eye(4, dtype=np.float32))    # theta = 0: axis = 0 / 1e-7 = 0, R = I


def test_oracle_backproject_project_transform():
    K = G["K"]
    close(Gm.backproject(G["depth"], K), G["points"])
    close(Gm.backproject_bwd(G["g_points"], K), G["d_depth"])
    close(Gm.transform3d(G["T"], G["points"]), G["tpoints"])
    d_p, d_T = Gm.transform3d_bwd(G["T"], G["points"], G["g_tpoints"])
    close(d_p, G["d_points_t3d"])
    close(d_T, G["d_T_t3d"])
    close(Gm.project(G["tpoints"], K), G["pix"])
    close(Gm.project_bwd(G["tpoints"], K, G["g_pix"]), G["d_points_proj"])


def test_oracle_roundtrip_identity_pose():
    """project(backproject(depth)) with T = I lands on the pixel grid, so WarpLayer is the identity."""
    K = G["K"]
    pts = Gm.backproject(G["depth"], K)
    pix = Gm.project(pts, K)
    B, h, w, _ = pix.shape
    X, Y = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    assert np.abs(pix[..., 0] - X).max() < 1e-4 and np.abs(pix[..., 1] - Y).max() < 1e-4


def test_oracle_warp_layer():
    K = G["K"]
    warped, pix = Gm.warp_layer(G["image"], G["points"], G["T"], K)
    close(warped, G["warped"], 1e-4)      # 4-tap blend of values whose coordinates carry 1e-6 noise
    d_img, d_pix = R.bilinear_interpolation_bwd(G["image"], pix, G["g_warped"])
    close(d_img, G["d_image"], 1e-4)
    d_q = Gm.project_bwd(Gm.transform3d(G["T"], G["points"]), K, d_pix)
    d_p, d_T = Gm.transform3d_bwd(G["T"], G["points"], d_q)
    close(d_p, G["d_points_warp"], 1e-4)
    close(d_T, G["d_T_warp"], 1e-4)


# ------------------------------------------------------------------------------ CUDA vs golden / oracle --
def dev(a, cuda):
    return torch.tensor(np.ascontiguousarray(a), device=cuda)


@pytest.mark.gpu
@pytest.mark.parametrize("invert", [False, True])
def test_cuda_pose_matrix(cuda, invert):
    from optical_flow_b200 import transformations as Tr
    tag = "inv" if invert else "fwd"
    raw = dev(G["raw"], cuda)
    close(Tr.make_transformation_matrix(raw, invert).cpu().numpy(), G["M_" + tag])
    close(Tr.make_transformation_matrix_backward(raw, invert, dev(G["gM_" + tag], cuda)).cpu().numpy(),
          G["d_raw_" + tag])
    close(Tr.rotation_from_axisangle(dev(G["raw"][:, :3], cuda)).cpu().numpy(), G["R"])
    z = Tr.make_transformation_matrix(torch.zeros((3, 6), device=cuda), invert).cpu().numpy()
    assert np.array_equal(z, np.broadcast_to(np.eye(4, dtype=np.float32), z.shape))


@pytest.mark.gpu
def test_cuda_layers_match_reference_run(cuda):
    from optical_flow_b200 import transformations as Tr
    K = G["K"]
    B, h, w, _ = G["depth"].shape
    bp = Tr.BackprojectLayer(K, h, w, B)
    close(bp(dev(G["depth"], cuda)).cpu().numpy(), G["points"])
    close(bp.backward(dev(G["g_points"], cuda)).cpu().numpy(), G["d_depth"])
    T, pts = dev(G["T"], cuda), dev(G["points"], cuda)
    close(Tr.transform3d(T, pts).cpu().numpy(), G["tpoints"])
    d_p, d_T = Tr.transform3d_backward(T, pts, dev(G["g_tpoints"], cuda))
    close(d_p.cpu().numpy(), G["d_points_t3d"])
    close(d_T.cpu().numpy(), G["d_T_t3d"])
    pr = Tr.ProjectLayer(K, h, w, B)
    tp = dev(G["tpoints"], cuda)
    close(pr(tp).cpu().numpy(), G["pix"])
    close(pr.backward(tp, dev(G["g_pix"], cuda)).cpu().numpy(), G["d_points_proj"])
    wl = Tr.WarpLayer(K, h, w, B)
    close(wl(dev(G["image"], cuda), pts, T).cpu().numpy(), G["warped"], 1e-4)
    d_img, d_p, d_T = wl.backward(dev(G["g_warped"], cuda))
    close(d_img.cpu().numpy(), G["d_image"], 1e-4)
    close(d_p.cpu().numpy(), G["d_points_warp"], 1e-4)
    close(d_T.cpu().numpy(), G["d_T_warp"], 1e-4)
    with pytest.raises(AssertionError):
        bp(dev(G["depth"][:, :-1], cuda))                         # transformations.py:26 shape assert


@pytest.mark.gpu
@pytest.mark.parametrize("B,h,w", [(1, 7, 5), (3, 96, 320), (8, 384, 1280)])
def test_cuda_view_chain_vs_oracle(cuda, B, h, w):
    """KITTI-shaped sizes (and a ragged one) against the NumPy oracle; at full size also the identity-pose
    round trip (size-independent property)."""
    from optical_flow_b200 import transformations as Tr
    r = np.random.default_rng(B * 1000 + h)
    K = np.array([[0.58 * w, 0, 0.5 * w], [0, 1.92 * h, 0.5 * h], [0, 0, 1]], np.float32)
    depth = (1.0 + r.random((B, h, w, 1)) * 20).astype(np.float32)
    raw = np.concatenate([r.standard_normal((B, 3)) * 0.02, r.standard_normal((B, 3)) * 0.1], 1).astype(np.float32)
    T = Tr.make_transformation_matrix(dev(raw, cuda), False)
    close(T.cpu().numpy(), Gm.make_transformation_matrix(raw, False))
    pts = Tr.BackprojectLayer(K, h, w, B)(dev(depth, cuda))
    pts_np = Gm.backproject(depth, K)
    close(pts.cpu().numpy(), pts_np)
    from optical_flow_b200 import functional as F
    pix = F.view_coords_fwd(K, T, pts)
    T_np = T.cpu().numpy()
    close(pix.cpu().numpy(), Gm.project(Gm.transform3d(T_np, pts_np), K), 2e-5)
    g = r.standard_normal((B, h, w, 2)).astype(np.float32)
    d_p, d_T = F.view_coords_bwd(K, T, pts, dev(g, cuda))
    e_p, e_T = Gm.transform3d_bwd(T_np, pts_np, Gm.project_bwd(Gm.transform3d(T_np, pts_np), K, g))
    close(d_p.cpu().numpy(), e_p, 2e-5)
    # d_T sums h*w products of mixed sign: bound by the sum of magnitudes (fp32 accumulation in another order)
    mag = np.einsum("bhwr,bhwc->brc", np.abs(Gm.project_bwd(Gm.transform3d(T_np, pts_np), K, g)), np.abs(pts_np))
    assert (np.abs(d_T.cpu().numpy() - e_T) <= 1e-5 * mag + 1e-6).all()
    eye = torch.eye(4, device=cuda).repeat(B, 1, 1).contiguous()
    pix0 = F.view_coords_fwd(K, eye, pts).cpu().numpy()
    X, Y = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    assert np.abs(pix0[..., 0] - X).max() < 2e-3 and np.abs(pix0[..., 1] - Y).max() < 2e-3
