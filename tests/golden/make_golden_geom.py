#!/usr/bin/env python
"""Golden vectors for the depth / pose re-projection helpers: EXECUTES THE REFERENCE'S OWN
transformations.py (/root/reference, imported unmodified) over the torch-backed TensorFlow stand-in in
tf_shim/ and writes tests/golden/geom.npz (inputs, outputs and tape gradients at a small shape).

Run in the build container only (/root/reference does not exist on the GPU box):

    python tests/golden/make_golden_geom.py
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(HERE, "tf_shim"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import tensorflow as tf  # noqa: E402  (the stand-in)

import transformations as ref_tr  # noqa: E402


def leaf(a):
    return tf.constant(np.ascontiguousarray(a, dtype=np.float32)).requires_grad_(True)


def npy(t):
    return t.detach().numpy().copy()


def main():
    r = np.random.default_rng(77)
    B, h, w = 2, 12, 20
    K = np.array([[18.0, 0.0, 9.5], [0.0, 17.0, 5.5], [0.0, 0.0, 1.0]], np.float32)
    out = {"K": K}

    # make_transformation_matrix (+ rotation_from_axisangle), both directions
    raw = np.concatenate([r.standard_normal((B, 3)) * 0.2, r.standard_normal((B, 3)) * 0.3], 1).astype(np.float32)
    out["raw"] = raw
    for invert in (False, True):
        x = leaf(raw)
        M = ref_tr.make_transformation_matrix(x, invert)
        g = r.standard_normal((B, 4, 4)).astype(np.float32)
        (d_raw,) = torch.autograd.grad((M * torch.tensor(g)).sum(), [x])
        tag = "inv" if invert else "fwd"
        out["M_" + tag], out["gM_" + tag], out["d_raw_" + tag] = npy(M), g, npy(d_raw)
    out["R"] = npy(ref_tr.rotation_from_axisangle(tf.constant(raw[:, :3])))

    # BackprojectLayer
    depth = (2.0 + r.random((B, h, w, 1)) * 3.0).astype(np.float32)
    d = leaf(depth)
    pts = ref_tr.BackprojectLayer(K, h, w, B)(d)
    g = r.standard_normal((B, h, w, 4)).astype(np.float32)
    (d_depth,) = torch.autograd.grad((pts * torch.tensor(g)).sum(), [d])
    out["depth"], out["points"], out["g_points"], out["d_depth"] = depth, npy(pts), g, npy(d_depth)

    # transform3d
    T = out["M_fwd"]
    p, Tt = leaf(out["points"]), leaf(T)
    q = ref_tr.transform3d(Tt, p)
    g = r.standard_normal((B, h, w, 4)).astype(np.float32)
    d_p, d_T = torch.autograd.grad((q * torch.tensor(g)).sum(), [p, Tt])
    out["T"], out["tpoints"], out["g_tpoints"] = T, npy(q), g
    out["d_points_t3d"], out["d_T_t3d"] = npy(d_p), npy(d_T)

    # ProjectLayer
    p = leaf(out["tpoints"])
    pix = ref_tr.ProjectLayer(tf.constant(K), h, w, B)(p)
    g = r.standard_normal((B, h, w, 2)).astype(np.float32)
    (d_p,) = torch.autograd.grad((pix * torch.tensor(g)).sum(), [p])
    out["pix"], out["g_pix"], out["d_points_proj"] = npy(pix), g, npy(d_p)

    # WarpLayer: image, points, T -> warped image; gradients to all three
    img = r.standard_normal((B, h, w, 3)).astype(np.float32)
    im, p, Tt = leaf(img), leaf(out["points"]), leaf(T)
    warped = ref_tr.WarpLayer(tf.constant(K), h, w, B)(im, p, Tt)
    g = r.standard_normal((B, h, w, 3)).astype(np.float32)
    d_im, d_p, d_T = torch.autograd.grad((warped * torch.tensor(g)).sum(), [im, p, Tt])
    out["image"], out["warped"], out["g_warped"] = img, npy(warped), g
    out["d_image"], out["d_points_warp"], out["d_T_warp"] = npy(d_im), npy(d_p), npy(d_T)

    np.savez_compressed(os.path.join(HERE, "geom.npz"), **out)
    print("wrote geom.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
