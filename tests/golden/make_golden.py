#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by EXECUTING THE REFERENCE'S OWN PYTHON
(/root/reference: transformations.py, model.py, loss.py, resnet/models.py, old/image_warp.py,
old/loss.py - imported unmodified) over the torch-backed TensorFlow stand-in in tf_shim/.

Run in the build container only (/root/reference does not exist on the GPU box):

    python tests/golden/make_golden.py

Outputs (committed): ops.npz (op-level inputs/outputs/gradients at small shapes) and
net_c1.npz (whole train-step graph at BASELINE configs[0]: batch 1, 128x416 - loss, the four
flows, and a digest of every weight gradient).  tests/test_a2_golden.py checks the oracle
(oracle/ref_np*.py, oracle/ref_torch.py) against them on CPU and the CUDA path on GPU.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(REF, "old"))      # old/loss.py does `from image_warp import image_warp`
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(HERE, "tf_shim"))

import importlib.util  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402
import tensorflow as tf  # noqa: E402  (the stand-in)
from tensorflow.keras import layers  # noqa: E402

import loss as ref_loss  # noqa: E402
import model as ref_model  # noqa: E402
import transformations as ref_tr  # noqa: E402


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_image_warp = _load("image_warp", os.path.join(REF, "old", "image_warp.py"))
sys.modules["image_warp"] = ref_image_warp
ref_old_loss = _load("old_loss", os.path.join(REF, "old", "loss.py"))

from oracle import ref_torch as T  # noqa: E402  (only for seeded inputs / parameter init)


def leaf(a):
    return tf.constant(np.ascontiguousarray(a, dtype=np.float32)).requires_grad_(True)


def npy(t):
    return t.detach().numpy().copy()


def adversarial_flow(r, shape, sigma):
    fl = (r.standard_normal(shape) * sigma).astype(np.float32)
    flat = fl.reshape(-1)
    k = max(flat.size // 16, 1)
    pos = r.choice(flat.size, size=4 * k, replace=False)
    ints = r.integers(-6, 7, size=k).astype(np.float32)
    flat[pos[:k]] = ints
    flat[pos[k:2 * k]] = np.nextafter(ints, np.float32(np.inf))
    flat[pos[2 * k:3 * k]] = np.nextafter(ints, np.float32(-np.inf))
    flat[pos[3 * k:]] = r.choice(np.array([-1000.5, 1000.5, -37.25, 41.75], np.float32), size=k)
    return fl


# ------------------------------------------------------------------------- op level --
def gen_ops():
    r = np.random.default_rng(20260101)
    out = {}

    # transformations.bilinear_interpolation + model.warp_features (fwd, grads)
    for tag, shape in (("a", (2, 12, 20, 8)), ("b", (1, 9, 7, 3))):          # W > H and H > W
        B, H, W, C = shape
        feats = r.standard_normal(shape).astype(np.float32)
        flow = adversarial_flow(r, (B, H, W, 2), 3.0)
        g = r.standard_normal(shape).astype(np.float32)
        ft, fl = leaf(feats), leaf(flow)
        y = ref_model.warp_features(fl, ft)
        d_feats, d_flow = torch.autograd.grad(y, [ft, fl], torch.as_tensor(g))
        out.update({"warp_%s_feats" % tag: feats, "warp_%s_flow" % tag: flow, "warp_%s_g" % tag: g,
                    "warp_%s_out" % tag: npy(y), "warp_%s_dfeats" % tag: npy(d_feats),
                    "warp_%s_dflow" % tag: npy(d_flow)})
        base = np.stack(np.meshgrid(np.arange(W), np.arange(H), indexing="xy"), -1).astype(np.float32)[None]
        pts = (base + flow).astype(np.float32)
        out["bil_%s_pts" % tag] = pts
        out["bil_%s_out" % tag] = npy(ref_tr.bilinear_interpolation(tf.constant(feats), tf.constant(pts)))

    # model.create_cost_volume, model.upscale_flow
    f1 = r.standard_normal((2, 7, 10, 16)).astype(np.float32)
    f2 = r.standard_normal((2, 7, 10, 16)).astype(np.float32)
    gcv = r.standard_normal((2, 7, 10, 49)).astype(np.float32)
    t1, t2 = leaf(f1), leaf(f2)
    cv = ref_model.create_cost_volume(t1, t2, 3)
    d1, d2 = torch.autograd.grad(cv, [t1, t2], torch.as_tensor(gcv))
    out.update(cv_f1=f1, cv_f2=f2, cv_g=gcv, cv_out=npy(cv), cv_df1=npy(d1), cv_df2=npy(d2))
    fl = r.standard_normal((2, 5, 6, 2)).astype(np.float32)
    gup = r.standard_normal((2, 10, 12, 2)).astype(np.float32)
    tfl = leaf(fl)
    up = ref_model.upscale_flow(tfl)
    out.update(up_flow=fl, up_g=gup, up_out=npy(up), up_dflow=npy(torch.autograd.grad(up, tfl, torch.as_tensor(gup))[0]))

    # loss.LossLayer (4 scales) with d loss / d flow_s
    B, H, W = 2, 32, 48
    imgs = T.synthetic_pairs(B, H, W, seed=5)
    flows = [(r.standard_normal((B, H >> (s + 1), W >> (s + 1), 2)) * (3.0 / 2 ** s)).astype(np.float32) for s in range(4)]
    tfl = [leaf(f) for f in flows]
    lv = ref_loss.LossLayer()(tf.constant(imgs), tfl)
    gs = torch.autograd.grad(lv, tfl)
    out.update(loss_imgs=imgs, loss_value=np.float32(lv.item()))
    for s in range(4):
        out["loss_flow%d" % s] = flows[s]
        out["loss_dflow%d" % s] = npy(gs[s])

    # old/image_warp.image_warp and image_warp_tf (fwd, grads)
    im = r.standard_normal((2, 10, 14, 3)).astype(np.float32)
    flow = adversarial_flow(r, (2, 10, 14, 2), 2.5)
    g = r.standard_normal(im.shape).astype(np.float32)
    ti, tfw = leaf(im), leaf(flow)
    y = ref_image_warp.image_warp(ti, tfw)
    d_im, d_fl = torch.autograd.grad(y, [ti, tfw], torch.as_tensor(g))
    y2 = ref_image_warp.image_warp_tf(tf.constant(im), tf.constant(flow))
    assert np.array_equal(npy(y), npy(y2)), "image_warp and image_warp_tf disagree"
    out.update(iw_im=im, iw_flow=flow, iw_g=g, iw_out=npy(y), iw_dim=npy(d_im), iw_dflow=npy(d_fl))

    # old/loss.MyLoss (5 scales coarse -> fine, photometric + smoothness; zero motions)
    B, H, W = 2, 32, 48
    imgs = T.synthetic_pairs(B, H, W, seed=6)
    flows = [(r.standard_normal((B, H >> k, W >> k, 2)) * (2.0 / 2 ** k)).astype(np.float32) for k in (4, 3, 2, 1, 0)]
    tfl = [leaf(f) for f in flows]
    eps_m = [tf.constant(np.zeros((B, H >> k, W >> k, 3), np.float32)) for k in (4, 3, 2, 1, 0)]
    lv = ref_old_loss.MyLoss(tf.constant(imgs), [tfl, eps_m, None, None])
    lv_nomotion = lv          # motions are zero -> motion_loss == 0 exactly (old/loss.py:96-101)
    gs = torch.autograd.grad(lv, tfl)
    out.update(myloss_imgs=imgs, myloss_value=np.float32(lv_nomotion.item()))
    for k in range(5):
        out["myloss_flow%d" % k] = flows[k]
        out["myloss_dflow%d" % k] = npy(gs[k])
    return out


# ------------------------------------------------------------------ whole network --
def grad_digest(g):
    """Compact fingerprint of one gradient tensor: sum, L2 norm, and up to 256 strided samples."""
    flat = g.reshape(-1).astype(np.float64)
    step = max(flat.size // 256, 1)
    return np.float64(flat.sum()), np.float64(np.sqrt((flat * flat).sum())), flat[::step][:256].astype(np.float32)


def gen_net(B=1, H=128, W=416, seed=42, data_seed=11):
    """train.py:50-55 on the reference's own model.py / loss.py graph: flows = flow_net(batch);
    loss = LossLayer()(batch, flows); grads = tape.gradient(loss, trainable_weights)."""
    layers.reset(batch=B)
    net = ref_model.build_flow_net(H, W, None)
    params, stats = T.init_params(seed=seed, randomize_bn=True)
    # decoder convs are unnamed in model.py:104-114 -> Keras auto-names conv2d, conv2d_1, ... in
    # creation order: flow0's six convs first (model.py:134-137)
    rename = {}
    dec = [n for n in layers.REGISTRY if n.startswith("conv2d")]
    assert len(dec) == 24
    for k, n in enumerate(dec):
        rename[n] = "flow%d_conv%d" % (k // 6, k % 6 + 1)
    weights = {}
    for lname, layer in layers.REGISTRY.items():
        ours = rename.get(lname, lname)
        for wname, w, trainable in layer.weights():
            key = ours + "/" + wname
            src = params[key] if trainable else stats[key]
            assert tuple(w.shape) == tuple(src.shape), (key, w.shape, src.shape)
            with torch.no_grad():
                w.copy_(torch.as_tensor(src))
            if trainable:
                weights[key] = w
    assert set(weights) == set(params), set(params) ^ set(weights)
    imgs = T.synthetic_pairs(B, H, W, seed=data_seed)
    x = tf.constant(imgs)
    with tf.GradientTape() as tape:
        flows = net(x)
        loss_value = ref_loss.LossLayer()(x, flows)
    names = list(weights)
    grads = tape.gradient(loss_value, [weights[n] for n in names])
    out = {"meta": np.array([B, H, W, seed, data_seed], np.int64), "loss": np.float32(loss_value.item())}
    for s, f in enumerate(flows):
        out["flow%d" % s] = npy(f)                   # finest first (model.py:140)
    for n, g in zip(names, grads):
        gs, gn, smp = grad_digest(npy(g) if g is not None else np.zeros(tuple(weights[n].shape), np.float32))
        out["gsum:" + n], out["gnorm:" + n], out["gsmp:" + n] = gs, gn, smp
    return out


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    ops = gen_ops()
    np.savez_compressed(os.path.join(HERE, "ops.npz"), **ops)
    print("ops.npz:", len(ops), "arrays")
    netv = gen_net()
    np.savez_compressed(os.path.join(HERE, "net_c1.npz"), **netv)
    print("net_c1.npz: loss", float(netv["loss"]), "arrays", len(netv))
