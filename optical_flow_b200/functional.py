"""Thin Python wrappers over the C ABI: one function per exported kernel entry.

Tensors are NHWC, contiguous, on a CUDA device; they are only buffers
(``data_ptr()``), all arithmetic happens in liboflow_b200.so on the current
CUDA stream.  There is no autograd: every op has an explicit ``*_bwd``.
"""
from __future__ import annotations

import torch

from . import _lib as L


def _empty_like_shape(t, shape, dtype=None):
    return torch.empty(shape, dtype=dtype or t.dtype, device=t.device)


# ------------------------------------------------------------------ sampler --

def bilinear_sample_fwd(inp, pts, return_indices=False):
    """transformations.py:85-129."""
    B, H, W, Cc = inp.shape
    assert pts.dtype == torch.float32, "sampling points must be float32 (transformations.py:95)"
    assert tuple(pts.shape) == (B, H, W, 2), "sampling_points must be (B,H,W,2) of the input's size"
    out = torch.empty_like(inp)
    idx = torch.empty((B, H, W, 4), dtype=torch.int32, device=inp.device) if return_indices else None
    L.check(L.load().of_bilinear_sample_fwd(L.ptr(inp), L.ptr(pts), L.ptr(out), L.ptr(idx), B, H, W, Cc,
                                            L.dtype_code(inp), L.stream_ptr()), "of_bilinear_sample_fwd")
    return (out, idx) if return_indices else out


def bilinear_sample_bwd(inp, pts, d_out, need_input_grad=True, need_points_grad=True):
    B, H, W, Cc = inp.shape
    assert d_out.dtype == torch.float32 and tuple(d_out.shape) == tuple(inp.shape)
    d_in = torch.empty((B, H, W, Cc), dtype=torch.float32, device=inp.device) if need_input_grad else None
    d_pts = torch.empty((B, H, W, 2), dtype=torch.float32, device=inp.device) if need_points_grad else None
    L.check(L.load().of_bilinear_sample_bwd(L.ptr(inp), L.ptr(pts), L.ptr(d_out), L.ptr(d_in), L.ptr(d_pts),
                                            B, H, W, Cc, L.dtype_code(inp), L.stream_ptr()), "of_bilinear_sample_bwd")
    return d_in, d_pts


def _mode_code(mode):
    if mode in ("ref_ij", L.OF_WARP_REF_IJ):
        return L.OF_WARP_REF_IJ
    if mode in ("xy", L.OF_WARP_XY):
        return L.OF_WARP_XY
    raise AssertionError("warp mode must be 'ref_ij' or 'xy'")


def warp_features_fwd(flow, feats, mode="ref_ij", return_indices=False, out=None):
    """model.py:55-73."""
    B, H, W, Cc = feats.shape
    assert flow.dtype == torch.float32
    assert flow.shape[1] == H and flow.shape[2] == W and flow.shape[3] == 2  # model.py:62-64
    assert flow.shape[0] == B
    if out is None:
        out = torch.empty_like(feats)
    idx = torch.empty((B, H, W, 4), dtype=torch.int32, device=feats.device) if return_indices else None
    L.check(L.load().of_warp_features_fwd(L.ptr(feats), L.ptr(flow), L.ptr(out), L.ptr(idx), B, H, W, Cc,
                                          _mode_code(mode), L.dtype_code(feats), L.stream_ptr()),
            "of_warp_features_fwd")
    return (out, idx) if return_indices else out


def warp_features_bwd(flow, feats, d_out, mode="ref_ij", need_feats_grad=True, need_flow_grad=True):
    B, H, W, Cc = feats.shape
    assert tuple(d_out.shape) == tuple(feats.shape)
    assert d_out.dtype == torch.float32 or (d_out.dtype == torch.bfloat16 and feats.dtype == torch.bfloat16)
    d_feats = torch.empty((B, H, W, Cc), dtype=torch.float32, device=feats.device) if need_feats_grad else None
    d_flow = torch.empty((B, H, W, 2), dtype=torch.float32, device=feats.device) if need_flow_grad else None
    L.check(L.load().of_warp_features_bwd_g(L.ptr(feats), L.ptr(flow), L.ptr(d_out), L.dtype_code(d_out), L.ptr(d_feats),
                                            L.ptr(d_flow), B, H, W, Cc, _mode_code(mode), L.dtype_code(feats),
                                            L.stream_ptr()), "of_warp_features_bwd_g")
    return d_feats, d_flow


def image_warp_fwd(im, flow, return_indices=False):
    """old/image_warp.py:7-78."""
    B, H, W, Cc = im.shape
    assert im.dtype == torch.float32 and flow.dtype == torch.float32
    assert tuple(flow.shape) == (B, H, W, 2)
    out = torch.empty_like(im)
    idx = torch.empty((B, H, W, 4), dtype=torch.int32, device=im.device) if return_indices else None
    L.check(L.load().of_image_warp_fwd(L.ptr(im), L.ptr(flow), L.ptr(out), L.ptr(idx), B, H, W, Cc,
                                       L.stream_ptr()), "of_image_warp_fwd")
    return (out, idx) if return_indices else out


def image_warp_bwd(im, flow, d_out, need_im_grad=True, need_flow_grad=True):
    B, H, W, Cc = im.shape
    d_im = torch.empty_like(im) if need_im_grad else None
    d_flow = torch.empty_like(flow) if need_flow_grad else None
    L.check(L.load().of_image_warp_bwd(L.ptr(im), L.ptr(flow), L.ptr(d_out), L.ptr(d_im), L.ptr(d_flow),
                                       B, H, W, Cc, L.stream_ptr()), "of_image_warp_bwd")
    return d_im, d_flow


# ------------------------------------------------------------------- resize --

def resize_bilinear(x, out_h, out_w):
    B, H, W, Cc = x.shape
    assert x.dtype == torch.float32
    out = torch.empty((B, out_h, out_w, Cc), dtype=torch.float32, device=x.device)
    L.check(L.load().of_resize_bilinear(L.ptr(x), L.ptr(out), B, H, W, Cc, out_h, out_w, L.stream_ptr()),
            "of_resize_bilinear")
    return out


def upscale_flow_fwd(flow):
    """model.py:76-77."""
    B, h, w, c = flow.shape
    assert c == 2 and flow.dtype == torch.float32
    out = torch.empty((B, 2 * h, 2 * w, 2), dtype=torch.float32, device=flow.device)
    L.check(L.load().of_upscale_flow_fwd(L.ptr(flow), L.ptr(out), B, h, w, L.stream_ptr()), "of_upscale_flow_fwd")
    return out


def upscale_flow_bwd(d_up):
    B, H2, W2, c = d_up.shape
    assert c == 2 and H2 % 2 == 0 and W2 % 2 == 0
    d_flow = torch.empty((B, H2 // 2, W2 // 2, 2), dtype=torch.float32, device=d_up.device)
    L.check(L.load().of_upscale_flow_bwd(L.ptr(d_up), L.ptr(d_flow), B, H2 // 2, W2 // 2, L.stream_ptr()),
            "of_upscale_flow_bwd")
    return d_flow


# -------------------------------------------------------------- cost volume --

def cost_volume_fwd(f1, f2, max_disp, out=None, cv_stride=None):
    """model.py:29-42."""
    assert f1.shape[1] == f2.shape[1]   # model.py:31-33
    assert f1.shape[2] == f2.shape[2]
    assert f1.shape[3] == f2.shape[3]
    assert f1.dtype == f2.dtype and f1.shape[0] == f2.shape[0]
    B, H, W, Cc = f1.shape
    D2 = (2 * max_disp + 1) ** 2
    if out is None:
        cv_stride = D2
        out = torch.empty((B, H, W, D2), dtype=f1.dtype, device=f1.device)
    else:
        cv_stride = cv_stride or out.shape[3]
    L.check(L.load().of_cost_volume_fwd(L.ptr(f1), L.ptr(f2), L.ptr(out), B, H, W, Cc, max_disp, cv_stride,
                                        L.dtype_code(f1), L.stream_ptr()), "of_cost_volume_fwd")
    return out


def cost_volume_bwd(f1, f2, d_cv, max_disp):
    B, H, W, Cc = f1.shape
    assert d_cv.dtype == f1.dtype
    d_f1 = torch.empty((B, H, W, Cc), dtype=torch.float32, device=f1.device)
    d_f2 = torch.empty((B, H, W, Cc), dtype=torch.float32, device=f1.device)
    L.check(L.load().of_cost_volume_bwd(L.ptr(f1), L.ptr(f2), L.ptr(d_cv), L.ptr(d_f1), L.ptr(d_f2), B, H, W, Cc,
                                        max_disp, d_cv.shape[3], L.dtype_code(f1), L.stream_ptr()),
            "of_cost_volume_bwd")
    return d_f1, d_f2


# ------------------------------------------------------------- fused losses --

_loss_ws = {}


def _workspace(device, B, H, W, num_scales):
    """Zero-initialised once, grown when a larger problem arrives (the kernels re-arm it)."""
    key = (device.type, device.index)
    need = L.load().of_loss_workspace_bytes(B, H, W, num_scales)
    ws = _loss_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(need, dtype=torch.uint8, device=device)
        _loss_ws[key] = ws
    return ws


def loss_l1_pyramid(batch_imgs, flows, need_grads=True, mode="ref_ij", loss_out=None, d_flows=None):
    """loss.py:5-32 in one launch.  Returns (loss[1] device tensor, d_flows or None)."""
    B, H, W, c6 = batch_imgs.shape
    assert c6 == 6 and batch_imgs.dtype == torch.float32
    for f in flows:
        assert f.dtype == torch.float32 and f.shape[0] == B and f.shape[3] == 2
    if loss_out is None:
        loss_out = torch.empty(1, dtype=torch.float32, device=batch_imgs.device)
    if need_grads and d_flows is None:
        d_flows = [torch.empty_like(f) for f in flows]
    ws = _workspace(batch_imgs.device, B, H, W, len(flows))
    L.check(L.load().of_loss_l1_pyramid_fwdbwd(
        L.ptr(batch_imgs), L.ptr_array(flows), L.ptr_array(d_flows) if need_grads else None,
        L.int_array([f.shape[1] for f in flows]), L.int_array([f.shape[2] for f in flows]), len(flows),
        L.ptr(loss_out), B, H, W, _mode_code(mode), L.ptr(ws), L.stream_ptr()), "of_loss_l1_pyramid_fwdbwd")
    return loss_out, (d_flows if need_grads else None)


def loss_charbonnier_smooth(y_true, flows_coarse_to_fine, need_grads=True, loss_out=None, d_flows=None):
    """old/loss.py:7-105 (photometric + smoothness) in one launch."""
    B, H, W, c6 = y_true.shape
    assert c6 == 6 and y_true.dtype == torch.float32
    flows = list(flows_coarse_to_fine)
    if loss_out is None:
        loss_out = torch.empty(1, dtype=torch.float32, device=y_true.device)
    if need_grads and d_flows is None:
        d_flows = [torch.empty_like(f) for f in flows]
    ws = _workspace(y_true.device, B, H, W, len(flows))
    L.check(L.load().of_loss_charbonnier_smooth_fwdbwd(
        L.ptr(y_true), L.ptr_array(flows), L.ptr_array(d_flows) if need_grads else None,
        L.int_array([f.shape[1] for f in flows]), L.int_array([f.shape[2] for f in flows]), len(flows),
        L.ptr(loss_out), B, H, W, L.ptr(ws), L.stream_ptr()), "of_loss_charbonnier_smooth_fwdbwd")
    return loss_out, (d_flows if need_grads else None)


# ---------------------------------------------------- pooling / elementwise --

def maxpool2_fwd(x):
    B, H, W, Cc = x.shape
    out = torch.empty((B, H // 2, W // 2, Cc), dtype=x.dtype, device=x.device)
    L.check(L.load().of_maxpool2_fwd(L.ptr(x), L.ptr(out), B, H, W, Cc, L.dtype_code(x), L.stream_ptr()),
            "of_maxpool2_fwd")
    return out


def maxpool2_bwd(x, d_out):
    B, H, W, Cc = x.shape
    assert d_out.dtype == x.dtype
    d_in = torch.empty_like(x)
    L.check(L.load().of_maxpool2_bwd(L.ptr(x), L.ptr(d_out), L.ptr(d_in), B, H, W, Cc, L.dtype_code(x),
                                     L.stream_ptr()), "of_maxpool2_bwd")
    return d_in


def maxpool2_bwd_add_act(x, d_out, add, act):
    """dz = (maxpool2_bwd(x, d_out) + add) * act'(x) in one pass (x = the pool input, post-activation)."""
    B, H, W, Cc = x.shape
    assert d_out.dtype == x.dtype and add.dtype == x.dtype and tuple(add.shape) == tuple(x.shape)
    dz = torch.empty_like(x)
    L.check(L.load().of_maxpool2_bwd_add_act(L.ptr(x), L.ptr(d_out), L.ptr(add), L.ptr(dz), B, H, W, Cc, act,
                                             L.dtype_code(x), L.stream_ptr()), "of_maxpool2_bwd_add_act")
    return dz


def add_relu_fwd(a, b):
    out = torch.empty_like(a)
    L.check(L.load().of_add_relu_fwd(L.ptr(a), L.ptr(b), L.ptr(out), a.numel(), L.dtype_code(a), L.stream_ptr()),
            "of_add_relu_fwd")
    return out


def act_bwd(y, dy, act, out=None):
    assert y.dtype == dy.dtype
    if out is None:
        out = torch.empty_like(dy)
    L.check(L.load().of_act_bwd(L.ptr(y), L.ptr(dy), L.ptr(out), y.numel(), act, L.dtype_code(y), L.stream_ptr()),
            "of_act_bwd")
    return out


def act_bwd_bits(bits, dy, act, out=None):
    """act_bwd from the 16 B / pixel sign-bit mask a conv fprop left (bits: uint8 (..., 16)); dy (..., C)."""
    C = dy.shape[-1]
    assert bits.dtype == torch.uint8 and bits.shape[-1] == 16 and bits.numel() // 16 == dy.numel() // C
    if out is None:
        out = torch.empty_like(dy)
    L.check(L.load().of_act_bwd_bits(L.ptr(bits), L.ptr(dy), L.ptr(out), dy.numel() // C, C, act, L.dtype_code(dy),
                                     L.stream_ptr()), "of_act_bwd_bits")
    return out


# -------------------------------------------------------------- convolution --

def same_pad(in_size, k, s):
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    return out, total // 2, total - total // 2


def conv2d_fprop(x, w, bias, stride=1, act=L.OF_ACT_NONE, out_dtype=None):
    """Keras Conv2D(padding='same'), NHWC x HWIO (fp32 SIMT path)."""
    N, H, W, Cin = x.shape
    KH, KW, Ci2, Cout = w.shape
    assert Ci2 == Cin and w.dtype == torch.float32
    Ho, Wo = same_pad(H, KH, stride)[0], same_pad(W, KW, stride)[0]
    out_dtype = out_dtype or x.dtype
    y = torch.empty((N, Ho, Wo, Cout), dtype=out_dtype, device=x.device)
    L.check(L.load().of_conv2d_fprop(L.ptr(x), L.ptr(w), L.ptr(bias), L.ptr(y), N, H, W, Cin, KH, KW, Cout, stride,
                                     act, L.dtype_code(x), L.dtype_code(y), L.stream_ptr()), "of_conv2d_fprop")
    return y


def conv2d_dgrad(dy, w, in_hw, stride=1, out_dtype=None):
    KH, KW, Cin, Cout = w.shape
    N = dy.shape[0]
    H, W = in_hw
    dx = torch.empty((N, H, W, Cin), dtype=out_dtype or dy.dtype, device=dy.device)
    L.check(L.load().of_conv2d_dgrad(L.ptr(dy), L.ptr(w), L.ptr(dx), N, H, W, Cin, KH, KW, Cout, stride,
                                     L.dtype_code(dy), L.dtype_code(dx), L.stream_ptr()), "of_conv2d_dgrad")
    return dx


def conv2d_wgrad(x, dy, ksize, stride=1, dw=None, db=None, need_bias_grad=True):
    N, H, W, Cin = x.shape
    Cout = dy.shape[3]
    KH, KW = ksize
    if dw is None:
        dw = torch.empty((KH, KW, Cin, Cout), dtype=torch.float32, device=x.device)
    if db is None and need_bias_grad:
        db = torch.empty((Cout,), dtype=torch.float32, device=x.device)
    L.check(L.load().of_conv2d_wgrad(L.ptr(x), L.ptr(dy), L.ptr(dw), L.ptr(db) if need_bias_grad else None,
                                     N, H, W, Cin, KH, KW, Cout, stride, L.dtype_code(x), L.dtype_code(dy),
                                     L.stream_ptr()), "of_conv2d_wgrad")
    return dw, db


# ---------------------------------------------------------------- optimiser --

def adam_step(param, grad, m, v, step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-7, grad_scale=1.0):
    """train.py:34,56 (in place on param, m, v)."""
    assert param.dtype == grad.dtype == m.dtype == v.dtype == torch.float32
    assert param.numel() == grad.numel() == m.numel() == v.numel()
    L.check(L.load().of_adam_step(L.ptr(param), L.ptr(grad), L.ptr(m), L.ptr(v), param.numel(), int(step),
                                  float(lr), float(beta1), float(beta2), float(eps), float(grad_scale),
                                  L.stream_ptr()), "of_adam_step")


def adam_step_dev(param, grad, m, v, state, beta1=0.9, beta2=0.999, eps=1e-7, grad_scale=1.0):
    """The same update with (step, lr) read from the 8-byte device ``state`` (of_adam_state) and the step count
    advanced on the stream: graph-capturable."""
    assert param.dtype == grad.dtype == m.dtype == v.dtype == torch.float32
    assert param.numel() == grad.numel() == m.numel() == v.numel() and state.numel() * state.element_size() == 8
    L.check(L.load().of_adam_step_dev(L.ptr(param), L.ptr(grad), L.ptr(m), L.ptr(v), param.numel(), L.ptr(state),
                                      float(beta1), float(beta2), float(eps), float(grad_scale), L.stream_ptr()),
            "of_adam_step_dev")


def l2_regularizer(param, grad, l2, loss=None):
    """Keras ``l2(l)`` on one tensor (opt-in): grad += 2 l w; loss (0-d / 1-element f32 tensor) += l sum(w^2)."""
    assert param.dtype == grad.dtype == torch.float32 and param.numel() == grad.numel()
    assert param.is_contiguous() and grad.is_contiguous()
    L.check(L.load().of_l2_regularizer(L.ptr(param), L.ptr(grad), param.numel(), float(l2), L.ptr(loss), L.stream_ptr()),
            "of_l2_regularizer")


# ------------------------------------------- depth / pose re-projection (8f rank 3) --

def _f32(shape, like):
    return torch.empty(shape, dtype=torch.float32, device=like.device)


def backproject_fwd(depth, kinv):
    """transformations.py:24-30; depth (B,h,w,1) -> points3d_hom (B,h,w,4)."""
    B, H, W, one = depth.shape
    assert one == 1 and depth.dtype == torch.float32
    out = _f32((B, H, W, 4), depth)
    L.check(L.load().of_backproject_fwd(L.ptr(depth), L.mat3_host(kinv), L.ptr(out), B, H, W, L.stream_ptr()),
            "of_backproject_fwd")
    return out


def backproject_bwd(d_points, kinv):
    B, H, W, four = d_points.shape
    assert four == 4 and d_points.dtype == torch.float32
    d_depth = _f32((B, H, W, 1), d_points)
    L.check(L.load().of_backproject_bwd(L.ptr(d_points), L.mat3_host(kinv), L.ptr(d_depth), B, H, W, L.stream_ptr()),
            "of_backproject_bwd")
    return d_depth


def project_fwd(points, k):
    """transformations.py:39-44; points3d_hom (B,h,w,4) -> pixel_coords (B,h,w,2)."""
    B, H, W, four = points.shape
    assert four == 4 and points.dtype == torch.float32
    pix = _f32((B, H, W, 2), points)
    L.check(L.load().of_project_fwd(L.mat3_host(k), L.ptr(points), L.ptr(pix), B, H, W, L.stream_ptr()),
            "of_project_fwd")
    return pix


def project_bwd(points, k, d_pix):
    B, H, W, _ = points.shape
    assert tuple(d_pix.shape) == (B, H, W, 2) and d_pix.dtype == torch.float32
    d_points = _f32((B, H, W, 4), points)
    L.check(L.load().of_project_bwd(L.mat3_host(k), L.ptr(points), L.ptr(d_pix), L.ptr(d_points), B, H, W,
                                    L.stream_ptr()), "of_project_bwd")
    return d_points


def _check_T(T, B):
    assert T.dtype == torch.float32 and tuple(T.shape) == (B, 4, 4), "T must be (batch_size, 4, 4) float32"


def transform3d_fwd(T, points):
    """transformations.py:61-67."""
    B, H, W, four = points.shape
    assert four == 4 and points.dtype == torch.float32
    _check_T(T, B)
    out = torch.empty_like(points)
    L.check(L.load().of_transform3d_fwd(L.ptr(T), L.ptr(points), L.ptr(out), B, H, W, L.stream_ptr()),
            "of_transform3d_fwd")
    return out


def transform3d_bwd(T, points, d_out, need_points_grad=True, need_T_grad=True):
    B, H, W, _ = points.shape
    _check_T(T, B)
    assert tuple(d_out.shape) == tuple(points.shape) and d_out.dtype == torch.float32
    d_points = torch.empty_like(points) if need_points_grad else None
    d_T = _f32((B, 4, 4), points) if need_T_grad else None
    L.check(L.load().of_transform3d_bwd(L.ptr(T), L.ptr(points), L.ptr(d_out), L.ptr(d_points), L.ptr(d_T), B, H, W,
                                        L.stream_ptr()), "of_transform3d_bwd")
    return d_points, d_T


def view_coords_fwd(k, T, points):
    """transformations.py:56-57 (transform3d + ProjectLayer) in one pass."""
    B, H, W, four = points.shape
    assert four == 4 and points.dtype == torch.float32
    _check_T(T, B)
    pix = _f32((B, H, W, 2), points)
    L.check(L.load().of_view_coords_fwd(L.mat3_host(k), L.ptr(T), L.ptr(points), L.ptr(pix), B, H, W,
                                        L.stream_ptr()), "of_view_coords_fwd")
    return pix


def view_coords_bwd(k, T, points, d_pix, need_points_grad=True, need_T_grad=True):
    B, H, W, _ = points.shape
    _check_T(T, B)
    assert tuple(d_pix.shape) == (B, H, W, 2) and d_pix.dtype == torch.float32
    d_points = torch.empty_like(points) if need_points_grad else None
    d_T = _f32((B, 4, 4), points) if need_T_grad else None
    L.check(L.load().of_view_coords_bwd(L.mat3_host(k), L.ptr(T), L.ptr(points), L.ptr(d_pix), L.ptr(d_points),
                                        L.ptr(d_T), B, H, W, L.stream_ptr()), "of_view_coords_bwd")
    return d_points, d_T


def make_transformation_matrix_fwd(raw, invert):
    """transformations.py:176-191."""
    B, six = raw.shape
    assert six == 6 and raw.dtype == torch.float32
    out = _f32((B, 4, 4), raw)
    L.check(L.load().of_make_transformation_matrix_fwd(L.ptr(raw), int(bool(invert)), L.ptr(out), B, L.stream_ptr()),
            "of_make_transformation_matrix_fwd")
    return out


def make_transformation_matrix_bwd(raw, invert, d_out):
    B, _ = raw.shape
    assert tuple(d_out.shape) == (B, 4, 4) and d_out.dtype == torch.float32
    d_raw = torch.empty_like(raw)
    L.check(L.load().of_make_transformation_matrix_bwd(L.ptr(raw), int(bool(invert)), L.ptr(d_out), L.ptr(d_raw), B,
                                                       L.stream_ptr()), "of_make_transformation_matrix_bwd")
    return d_raw
