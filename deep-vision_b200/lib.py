"""ctypes binding of libdv_b200.so — the C ABI declared in include/dv.h.

The structures below mirror include/dv.h field for field.  There is no CPU fallback: if the shared library is not
built (``python __graft_entry__.py build`` or ``make -C deep-vision_b200/csrc``) importing this module raises, and every
compute entry point fails on a machine without an sm_100 GPU.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DV_B200_LIB") or os.path.join(_HERE, "libdv_b200.so")   # override: A/B builds of the same ABI

DV_F32, DV_BF16 = 0, 1
DV_ALGO_AUTO, DV_ALGO_SIMT, DV_ALGO_TC = 0, 1, 2
DV_CONV_FPROP, DV_CONV_DGRAD, DV_CONV_WGRAD = 0, 1, 2
DV_OK, DV_ERR_ARG, DV_ERR_UNSUPPORTED, DV_ERR_CUDA, DV_ERR_ARCH, DV_ERR_WORKSPACE = 0, 1, 2, 3, 4, 5


class ConvParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p), ("w_t", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("stat_sums", C.c_void_p),
        ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32), ("K", C.c_int32),
        ("R", C.c_int32), ("S", C.c_int32), ("stride", C.c_int32), ("pad", C.c_int32),
        ("dtype", C.c_int32), ("algo", C.c_int32), ("y_f32", C.c_int32), ("accumulate", C.c_int32),
        ("ep_scale", C.c_void_p), ("ep_shift", C.c_void_p), ("ep_residual", C.c_void_p), ("ep_relu", C.c_int32),
        ("bnb_x", C.c_void_p), ("bnb_gamma", C.c_void_p), ("bnb_beta", C.c_void_p), ("bnb_mean", C.c_void_p),
        ("bnb_invstd", C.c_void_p), ("bnb_dgamma", C.c_void_p), ("bnb_dbeta", C.c_void_p), ("ep_residual_mask", C.c_void_p),
    ]


class StemParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p), ("dw", C.c_void_p), ("stat_sums", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32), ("K", C.c_int32),
        ("R", C.c_int32), ("S", C.c_int32), ("stride", C.c_int32), ("pad", C.c_int32), ("accumulate", C.c_int32),
    ]


class DwConvParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32), ("stride", C.c_int32),
        ("dtype", C.c_int32), ("ep_shift", C.c_void_p), ("ep_relu", C.c_int32), ("stat_sums", C.c_void_p),
        ("bnb_x", C.c_void_p), ("bnb_gamma", C.c_void_p), ("bnb_beta", C.c_void_p), ("bnb_mean", C.c_void_p),
        ("bnb_invstd", C.c_void_p), ("bnb_dgamma", C.c_void_p), ("bnb_dbeta", C.c_void_p),
    ]


class AugmentParams(C.Structure):
    _fields_ = [
        ("src", C.c_void_p), ("dst", C.c_void_p), ("geom", C.c_void_p), ("factors", C.c_void_p), ("widths", C.c_void_p),
        ("gray_sums", C.c_void_p),
        ("B", C.c_int32), ("Hs", C.c_int32), ("Ws", C.c_int32), ("crop", C.c_int32), ("dtype", C.c_int32), ("nchw", C.c_int32),
        ("mean", C.c_float * 3), ("std", C.c_float * 3),
    ]


class RescaleParams(C.Structure):
    _fields_ = [
        ("src", C.c_void_p), ("src_off", C.c_void_p), ("src_hw", C.c_void_p), ("dst", C.c_void_p), ("dst_hw", C.c_void_p),
        ("B", C.c_int32), ("Hs", C.c_int32), ("Ws", C.c_int32),
    ]


class BnParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("y", C.c_void_p), ("residual", C.c_void_p),
        ("gamma", C.c_void_p), ("beta", C.c_void_p),
        ("running_mean", C.c_void_p), ("running_var", C.c_void_p),
        ("save_mean", C.c_void_p), ("save_invstd", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("M", C.c_int64), ("C", C.c_int32), ("dtype", C.c_int32), ("relu", C.c_int32), ("eval_mode", C.c_int32),
        ("eps", C.c_float), ("momentum", C.c_float), ("relu_mask", C.c_void_p), ("reverse", C.c_int32),
    ]


class BnPoolParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("y", C.c_void_p), ("idx", C.c_void_p), ("dy", C.c_void_p), ("dx", C.c_void_p),
        ("gamma", C.c_void_p), ("beta", C.c_void_p), ("save_mean", C.c_void_p), ("save_invstd", C.c_void_p),
        ("dgamma", C.c_void_p), ("dbeta", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32), ("dtype", C.c_int32),
    ]


class BnBwdParams(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("y", C.c_void_p), ("dy", C.c_void_p), ("dx", C.c_void_p), ("dres", C.c_void_p),
        ("gamma", C.c_void_p), ("beta", C.c_void_p), ("save_mean", C.c_void_p), ("save_invstd", C.c_void_p),
        ("dgamma", C.c_void_p), ("dbeta", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("M", C.c_int64), ("C", C.c_int32), ("dtype", C.c_int32), ("relu", C.c_int32), ("atomic_sums", C.c_int32),
        ("relu_mask", C.c_void_p), ("reverse", C.c_int32),
    ]


# every symbol include/dv.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_SIG = {
    "dv_conv_fprop": (C.c_int, [C.POINTER(ConvParams), _P]),
    "dv_conv_dgrad": (C.c_int, [C.POINTER(ConvParams), _P]),
    "dv_conv_wgrad": (C.c_int, [C.POINTER(ConvParams), _P]),
    "dv_conv_workspace_bytes": (C.c_size_t, [C.POINTER(ConvParams), C.c_int]),
    "dv_conv_algo": (C.c_int, [C.POINTER(ConvParams), C.c_int]),
    "dv_filter_prep": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_im2col": (C.c_int, [_P, _P] + [C.c_int] * 10 + [_P]),
    "dv_pad_rows": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_unpad_rows": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, _P]),
    "dv_stem_supported": (C.c_int, [C.POINTER(StemParams), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "dv_stem_pad_input": (C.c_int, [C.POINTER(StemParams), _P, _P]),
    "dv_stem_pad_input_nchw": (C.c_int, [C.POINTER(StemParams), _P, _P]),
    "dv_stem_pack_filters": (C.c_int, [C.POINTER(StemParams), _P, _P]),
    "dv_stem_fprop": (C.c_int, [C.POINTER(StemParams), _P]),
    "dv_stem_wgrad": (C.c_int, [C.POINTER(StemParams), _P]),
    "dv_stem_workspace_bytes": (C.c_size_t, [C.POINTER(StemParams)]),
    "dv_dwconv_fuses_stats": (C.c_int, [C.POINTER(DwConvParams)]),
    "dv_dwconv_fuses_bn_bwd": (C.c_int, [C.POINTER(DwConvParams)]),
    "dv_dwconv_fprop": (C.c_int, [C.POINTER(DwConvParams), _P]),
    "dv_dwconv_dgrad": (C.c_int, [C.POINTER(DwConvParams), _P]),
    "dv_dwconv_wgrad": (C.c_int, [C.POINTER(DwConvParams), _P]),
    "dv_dwconv_workspace_bytes": (C.c_size_t, [C.POINTER(DwConvParams)]),
    "dv_augment": (C.c_int, [C.POINTER(AugmentParams), _P]),
    "dv_rescale": (C.c_int, [C.POINTER(RescaleParams), _P]),
    "dv_bn_fold": (C.c_int, [_P, _P, _P, _P, C.c_float, _P, _P, C.c_int, _P, _P, _P]),
    "dv_bn_stats": (C.c_int, [C.POINTER(BnParams), _P]),
    "dv_bn_apply": (C.c_int, [C.POINTER(BnParams), _P]),
    "dv_bn_stats_from_sums": (C.c_int, [C.POINTER(BnParams), _P, _P]),
    "dv_bn_stats_static": (C.c_int, [C.POINTER(BnParams), _P]),
    "dv_bn_bwd_reduce": (C.c_int, [C.POINTER(BnBwdParams), _P]),
    "dv_bn_bwd_apply": (C.c_int, [C.POINTER(BnBwdParams), _P]),
    "dv_bn_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int]),
    "dv_bn_relu_maxpool_fwd": (C.c_int, [C.POINTER(BnPoolParams), _P]),
    "dv_bn_relu_maxpool_bwd": (C.c_int, [C.POINTER(BnPoolParams), _P]),
    "dv_bn_pool_workspace_bytes": (C.c_size_t, [C.POINTER(BnPoolParams)]),
    "dv_relu_fwd": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P]),
    "dv_relu_bwd": (C.c_int, [_P, _P, _P, C.c_int64, C.c_int, _P]),
    "dv_maxpool_fwd": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_maxpool_bwd": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_avgpool_fwd": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_avgpool_bwd": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_nchw_to_nhwc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dv_cast": (C.c_int, [_P, C.c_int, _P, C.c_int, C.c_int64, _P]),
    "dv_bias_add": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P]),
    "dv_bias_grad": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P]),
    "dv_cross_entropy": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, _P]),
    "dv_scale_by": (C.c_int, [_P, _P, C.c_int64, _P]),
    "dv_sgd_step": (C.c_int, [_P, _P, _P, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, _P]),
    "dv_rmsprop_step": (C.c_int, [_P, _P, _P, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _P]),
    "dv_sgd_step_bf16": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, _P]),
    "dv_rmsprop_step_bf16": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _P]),
    "dv_version": (C.c_int, []),
    "dv_last_error": (C.c_char_p, []),
    "dv_device_check": (C.c_int, []),
    "dv_launch_count": (C.c_uint64, []),
}

EXPORTED_SYMBOLS = tuple(_SIG)

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is not built; run `python __graft_entry__.py build` (or `make -C deep-vision_b200/csrc`). "
        "deep_vision_b200 has no CPU or PyTorch fallback.")

lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in _SIG.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


class DvError(RuntimeError):
    pass


def check(rc, what=""):
    """Raise the library's message as a RuntimeError, the way torch ops surface their errors."""
    if rc != 0:
        msg = lib.dv_last_error()
        raise DvError(f"{what}: dv error {rc}: {msg.decode() if msg else ''}")
