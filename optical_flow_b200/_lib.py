"""ctypes binding of liboflow_b200.so (the C ABI declared in include/oflow_b200.h).

The product path has NO CPU fallback: if the shared library is missing or a
call fails, this module raises.  torch tensors are used only as device buffers
(``data_ptr()``) and for the current CUDA stream.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# OF_PROFILE_LIB=1 selects the -DOF_PROFILE_HOOKS build (scripts/ only; tests, smoke and bench use the shipped library)
# (OF_PROFILE_LIB=<tag> selects an A/B variant liboflow_b200_<tag>.so built by `build.py --tag=<tag> -D...`)
_variant = os.environ.get("OF_PROFILE_LIB", "")
LIB_PATH = os.path.join(_HERE, "liboflow_b200.so" if not _variant else
                        ("liboflow_b200_prof.so" if _variant == "1" else "liboflow_b200_%s.so" % _variant))

OF_F32, OF_BF16 = 0, 1
OF_MASK_BITS = 2      # of_tc_conv.mask_dtype: sign-bit mask written by a conv_tc3 fprop (sign_out)
OF_WARP_REF_IJ, OF_WARP_XY = 0, 1
OF_ACT_NONE, OF_ACT_RELU, OF_ACT_LEAKY = 0, 1, 2

_vp, _i, _sz, _f = C.c_void_p, C.c_int, C.c_size_t, C.c_float
_pp = C.POINTER(C.c_void_p)
_ip = C.POINTER(C.c_int)
_fp9 = C.POINTER(C.c_float)   # 9 host floats (row-major 3x3 intrinsics)

class TcConv(C.Structure):
    """of_tc_conv (include/oflow_b200.h)."""
    _fields_ = [("src", C.c_void_p * 2), ("src_c", C.c_int * 2), ("src_stride", C.c_int * 2), ("num_src", C.c_int),
                ("wpacked", C.c_void_p), ("n_rows", C.c_int), ("k_pad", C.c_int),
                ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Ho", C.c_int), ("Wo", C.c_int),
                ("KH", C.c_int), ("KW", C.c_int), ("stride", C.c_int), ("pad_t", C.c_int), ("pad_l", C.c_int),
                ("sign", C.c_int),
                ("view_w_stride", C.c_longlong), ("view_h_stride", C.c_longlong), ("view_n_stride", C.c_longlong),
                ("stride_w", C.c_int), ("stride_h", C.c_int),
                ("n_mma", C.c_int), ("n_valid", C.c_int),
                ("out", C.c_void_p), ("out_dtype", C.c_int), ("out_stride", C.c_int),
                ("bias", C.c_void_p), ("act", C.c_int),
                ("add", C.c_void_p), ("add_dtype", C.c_int), ("add_stride", C.c_int),
                ("mask", C.c_void_p), ("mask_dtype", C.c_int), ("mask_stride", C.c_int), ("mask_act", C.c_int),
                ("sign_out", C.c_void_p),
                ("out2", C.c_void_p), ("out2_stride", C.c_int), ("out_split", C.c_int), ("add_mode", C.c_int)]


class TcWgrad(C.Structure):
    """of_tc_wgrad (include/oflow_b200.h)."""
    _fields_ = [("x", C.c_void_p * 2), ("x_c", C.c_int * 2), ("x_stride", C.c_int * 2), ("x_valid", C.c_int * 2),
                ("x_ci0", C.c_int * 2), ("num_src", C.c_int),
                ("dy", C.c_void_p), ("dy_c", C.c_int), ("dy_stride", C.c_int),
                ("dw", C.c_void_p),
                ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Ho", C.c_int), ("Wo", C.c_int),
                ("KH", C.c_int), ("KW", C.c_int), ("stride", C.c_int), ("pad_t", C.c_int), ("pad_l", C.c_int),
                ("Cin", C.c_int), ("Cout", C.c_int),
                ("view_w_stride", C.c_longlong), ("view_h_stride", C.c_longlong), ("view_n_stride", C.c_longlong),
                ("stride_w", C.c_int), ("stride_h", C.c_int), ("db", C.c_void_p)]


class PackJob(C.Structure):
    """of_pack_job (include/oflow_b200.h)."""
    _fields_ = [("w", C.c_void_p), ("out", C.c_void_p),
                ("taps", C.c_int), ("Cin", C.c_int), ("Cout", C.c_int), ("rows", C.c_int), ("cols", C.c_int),
                ("transpose", C.c_int), ("nseg", C.c_int),
                ("seg_off", C.c_int * 3), ("seg_cnt", C.c_int * 3), ("seg_base", C.c_int * 3),
                ("block_begin", C.c_int), ("reserved", C.c_int)]


class BnJob(C.Structure):
    """of_bn_job (include/oflow_b200.h)."""
    _fields_ = [(n, C.c_void_p) for n in ("w", "b", "gamma", "beta", "mean", "var", "wf", "bf", "dwf", "dbf", "dw", "db",
                                          "dgamma", "dbeta")] + \
               [("K", C.c_int), ("Cout", C.c_int), ("fwd_block_begin", C.c_int), ("bwd_block_begin", C.c_int)]


class ProfRec(C.Structure):
    """of_prof_rec (include/oflow_b200.h)."""
    _fields_ = [("name", C.c_char * 40), ("work", C.c_double), ("ms", C.c_float)]


OF_BN_FWD_BLOCK = 2048

# name -> (restype, argtypes); must list every symbol include/oflow_b200.h declares
PROTOTYPES = {
    "of_conv_tc_run": (C.c_int, [C.POINTER(TcConv), C.c_void_p]),
    "of_conv_tc_pack_weights_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "of_conv_tc_pack_weights": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                          C.c_void_p]),
    "of_conv_tc_wgrad": (C.c_int, [C.POINTER(TcWgrad), C.c_void_p]),
    "of_pack_image_pair8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "of_conv1_pack_weights": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "of_conv1_unpack_wgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "of_zero_upsample2": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_void_p]),
    "of_scatter_upsample2": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_void_p]),
    "of_colsum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "of_version": (_i, []),
    "of_last_error": (C.c_char_p, []),
    "of_launch_count": (C.c_ulonglong, []),
    "of_set_option": (C.c_int, [C.c_char_p, C.c_char_p]),
    "of_prof_enable": (C.c_int, [C.c_int, C.c_void_p]),
    "of_prof_collect": (C.c_int, [C.POINTER(ProfRec), C.c_int]),
    "of_prof_marker": (C.c_int, [C.c_char_p]),
    "of_bilinear_sample_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "of_bilinear_sample_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "of_warp_features_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_warp_features_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_warp_features_bwd_g": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_image_warp_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "of_image_warp_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "of_resize_bilinear": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_upscale_flow_fwd": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "of_upscale_flow_bwd": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "of_cost_volume_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "of_cost_volume_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "of_preprocess_pair": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _i, _vp]),
    "of_preprocess_batch": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "of_flow_head_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_cost_volume_bwd_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_loss_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "of_loss_l1_pyramid_fwdbwd": (_i, [_vp, _pp, _pp, _ip, _ip, _i, _vp, _i, _i, _i, _i, _vp, _vp]),
    "of_loss_charbonnier_smooth_fwdbwd": (_i, [_vp, _pp, _pp, _ip, _ip, _i, _vp, _i, _i, _i, _vp, _vp]),
    "of_maxpool2_fwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "of_maxpool2_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "of_maxpool2_bwd_add_act": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "of_add_relu_fwd": (_i, [_vp, _vp, _vp, _sz, _i, _vp]),
    "of_relu_bwd": (_i, [_vp, _vp, _vp, _sz, _i, _vp]),
    "of_act_bwd": (_i, [_vp, _vp, _vp, _sz, _i, _i, _vp]),
    "of_act_bwd_bits": (_i, [_vp, _vp, _vp, _sz, _i, _i, _i, _vp]),
    "of_split_bf16": (_i, [_vp, _vp, _vp, _sz, _i, _i, _i, _vp]),
    "of_conv2d_fprop": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "of_conv2d_dgrad": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "of_conv2d_wgrad": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "of_split_image_pair": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "of_copy_channels": (_i, [_vp, _i, _vp, _i, _i, _sz, _i, _i, _i, _vp]),
    "of_bn_fold_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _i, _i, _vp]),
    "of_bn_fold_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "of_backproject_fwd": (_i, [_vp, _fp9, _vp, _i, _i, _i, _vp]),
    "of_backproject_bwd": (_i, [_vp, _fp9, _vp, _i, _i, _i, _vp]),
    "of_project_fwd": (_i, [_fp9, _vp, _vp, _i, _i, _i, _vp]),
    "of_project_bwd": (_i, [_fp9, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "of_transform3d_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "of_transform3d_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "of_view_coords_fwd": (_i, [_fp9, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "of_view_coords_bwd": (_i, [_fp9, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "of_make_transformation_matrix_fwd": (_i, [_vp, _i, _vp, _i, _vp]),
    "of_make_transformation_matrix_bwd": (_i, [_vp, _i, _vp, _vp, _i, _vp]),
    "of_bn_fold_fwd_batch": (_i, [_vp, _i, _i, _f, _vp]),
    "of_bn_fold_bwd_batch": (_i, [_vp, _i, _i, _f, _vp]),
    "of_adam_step": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _f, _f, _f, _f, _f, _vp]),
    "of_adam_step_dev": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _f, _f, _f, _f, _vp]),
    "of_bn_train_workspace_bytes": (_sz, [_i]),
    "of_bn_train_stats": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _i, _vp, _vp]),
    "of_bn_train_stats_static": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _i, _vp]),
    "of_bn_train_apply": (_i, [_vp, _vp, _vp, _vp, _f, _vp, _vp, _f, _i, _vp, _vp, _vp, _f, _sz, _i, _i, _vp]),
    "of_bn_train_bwd_reduce": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _i, _vp, _vp, _sz, _i, _i, _vp, _vp]),
    "of_bn_train_bwd_apply": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _f, _i, _vp, _vp, _vp, _sz, _i, _i, _vp]),
    "of_l2_regularizer": (_i, [_vp, _vp, _sz, _f, _vp, _vp]),
    "of_comm_unique_id": (_i, [_vp]),
    "of_comm_init": (_i, [_vp, _i, _i, _pp]),
    "of_comm_allreduce_sum_f32": (_i, [_vp, _vp, _sz, _vp]),
    "of_comm_destroy": (_i, [_vp]),
}

_lib = None


def load():
    """Load the shared library (once) and set every prototype.  Raises if the
    library is absent: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "liboflow_b200.so is missing (%s). Build it with `python -m optical_flow_b200.build`; "
            "there is no CPU / PyTorch fallback for this path." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)   # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class option:
    """``with option("OF_CV_NO_TC", "1"): ...`` - select a kernel variant for the duration of the block
    (of_set_option; the library reads no environment variables)."""

    def __init__(self, name, value="1"):
        self.name, self.value = name.encode(), None if value is None else str(value).encode()

    def __enter__(self):
        check(load().of_set_option(self.name, self.value), "of_set_option")
        return self

    def __exit__(self, *exc):
        load().of_set_option(self.name, None)
        return False


def timeline(fn, max_recs=8192):
    """Run ``fn()`` with the per-launch timeline on (current stream); returns [(kernel name, ms, work), ...]."""
    lib = load()
    check(lib.of_prof_enable(1, stream_ptr()), "of_prof_enable")
    try:
        fn()
    finally:
        recs = (ProfRec * max_recs)()
        n = lib.of_prof_collect(recs, max_recs)
        lib.of_prof_enable(0, None)
    return [(recs[i].name.decode(), float(recs[i].ms), float(recs[i].work)) for i in range(n)]


def check(rc: int, what: str = ""):
    """Error convention of the C ABI -> Python exceptions:
    < 0 argument/shape error == the reference's trace-time asserts -> AssertionError,
    > 0 cudaError_t -> RuntimeError."""
    if rc == 0:
        return
    msg = load().of_last_error().decode("utf-8", "replace")
    if rc < 0:
        raise AssertionError("%s: %s (code %d)" % (what, msg, rc))
    raise RuntimeError("%s: CUDA error %d: %s" % (what, rc, msg))


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return OF_F32
    if t.dtype == torch.bfloat16:
        return OF_BF16
    raise AssertionError("unsupported dtype %s (float32 / bfloat16 only)" % t.dtype)


def ptr(t):
    if t is None:
        return None
    assert t.is_cuda, "tensor must live on a CUDA device (no CPU path)"
    assert t.is_contiguous(), "tensor must be contiguous NHWC"
    return t.data_ptr()


def ptr_array(tensors):
    arr = (C.c_void_p * len(tensors))()
    for k, t in enumerate(tensors):
        arr[k] = ptr(t)
    return arr


def mat3_host(m):
    """3x3 intrinsics (numpy / nested lists / CPU tensor) -> 9 host floats, row-major."""
    import numpy as np
    a = np.ascontiguousarray(np.asarray(m, dtype=np.float32)).reshape(-1)
    assert a.size == 9, "intrinsics matrix must be (3, 3)"
    return (C.c_float * 9)(*[float(v) for v in a])


def int_array(vals):
    arr = (C.c_int * len(vals))()
    for k, v in enumerate(vals):
        arr[k] = int(v)
    return arr
