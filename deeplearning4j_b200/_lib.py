"""ctypes binding of libdl4jb200.so -- the same C ABI (include/dl4jb200.h) the Java helpers bind through JavaCPP.

There is no fallback: if the shared library is missing or a call fails, this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libdl4jb200.so")

OK, NOT_SUPPORTED = 0, 1
ERR_INVALID_ARG, ERR_CUDA, ERR_INVALID_INPUT, ERR_INVALID_CONFIG = -1, -2, -3, -4
MODE_STRICT, MODE_TRUNCATE, MODE_SAME = 0, 1, 2
NCHW, NHWC = 0, 1
BUF_HOST_NEWER, BUF_SYNCED, BUF_DEVICE_NEWER = 0, 1, 2
BUF_READ, BUF_WRITE = 1, 2
POOL_MAX, POOL_AVG, POOL_SUM, POOL_PNORM = 0, 1, 2, 3
ACT_IDENTITY, ACT_RELU, ACT_SIGMOID, ACT_TANH, ACT_LEAKYRELU = 0, 1, 2, 3, 4
ACT_FROM_OUTPUT = 0x200
ALGO_TC_TF32, ALGO_FP32 = 0, 1
OPT_TF32_ACTIVATIONS = 1
OPT_PAIR_TILES = 2
UPD_SGD, UPD_NESTEROVS, UPD_RMSPROP, UPD_ADADELTA, UPD_NOOP = 0, 1, 2, 3, 4


class ConvDesc(C.Structure):
    """dl4jb200_conv_desc"""
    _fields_ = [(n, C.c_int32) for n in (
        "batch", "c_in", "in_h", "in_w", "c_out", "k_h", "k_w", "stride_h", "stride_w", "pad_top", "pad_left",
        "dil_h", "dil_w", "out_h", "out_w")]


class DL4JB200Error(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"dl4jb200 status {status}: {msg}")
        self.status = status


class DL4JInvalidInputException(ValueError):
    """org.deeplearning4j.exception.DL4JInvalidInputException"""


class DL4JInvalidConfigException(ValueError):
    """org.deeplearning4j.exception.DL4JInvalidConfigException"""


class HelperNotSupported(Exception):
    """status 1: the Java helper returns null and the layer runs its built-in path."""


_vp, _i, _sz, _d = C.c_void_p, C.c_int, C.c_size_t, C.c_double
_desc_p = C.POINTER(ConvDesc)

# name -> argtypes, mirrors include/dl4jb200.h declaration by declaration
PROTOTYPES = {
    "dl4jb200_ctx_create": [_i, _vp, C.POINTER(_vp)],
    "dl4jb200_ctx_destroy": [_vp],
    "dl4jb200_ctx_set_stream": [_vp, _vp],
    "dl4jb200_ctx_set_option": [_vp, _i, _i],
    "dl4jb200_ctx_sync": [_vp],
    "dl4jb200_ctx_workspace_freeze": [_vp, _i],
    "dl4jb200_malloc": [_vp, _sz, C.POINTER(_vp)],
    "dl4jb200_free": [_vp, _vp],
    "dl4jb200_memset": [_vp, _vp, _i, _sz],
    "dl4jb200_memcpy_h2d": [_vp, _vp, _vp, _sz],
    "dl4jb200_memcpy_d2h": [_vp, _vp, _vp, _sz],
    "dl4jb200_memcpy_d2d": [_vp, _vp, _vp, _sz],
    "dl4jb200_buf_register": [_vp, _vp, _sz, C.POINTER(_vp)],
    "dl4jb200_buf_acquire": [_vp, _vp, _sz, _i, C.POINTER(_vp)],
    "dl4jb200_buf_mark_host_dirty": [_vp, _vp],
    "dl4jb200_buf_sync_to_host": [_vp, _vp],
    "dl4jb200_buf_state": [_vp, _vp, C.POINTER(_i)],
    "dl4jb200_ctx_transfer_bytes": [_vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)],
    "dl4jb200_buf_lookup": [_vp, _vp, C.POINTER(_vp), C.POINTER(_sz)],
    "dl4jb200_buf_upload": [_vp, _vp],
    "dl4jb200_buf_download": [_vp, _vp],
    "dl4jb200_buf_release": [_vp, _vp],
    "dl4jb200_out_size": [_i, _i, _i, _i, _i, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "dl4jb200_nchw_to_nhwc": [_vp, _vp, _vp, _i, _i, _i, _i, _i],
    "dl4jb200_nhwc_to_nchw": [_vp, _vp, _vp, _i, _i, _i, _i, _i],
    "dl4jb200_conv_fwd": [_vp, _desc_p, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i],
    "dl4jb200_conv_bwd": [_vp, _desc_p, _vp, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i],
    "dl4jb200_bias_grad": [_vp, _vp, _i, _i, _i, _vp],
    "dl4jb200_nchw_to_rowpack4": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i],
    "dl4jb200_conv_rowpack_fwd": [_vp, _desc_p, _vp, _i, _vp, _vp, _vp, _i, _i],
    "dl4jb200_conv_rowpack_fwd_bnstats": [_vp, _desc_p, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, C.POINTER(C.c_int)],
    "dl4jb200_conv_rowpack_wgrad": [_vp, _desc_p, _vp, _i, _vp, _i, _vp, _vp],
    "dl4jb200_act_fwd": [_vp, _vp, _vp, _sz, _i],
    "dl4jb200_act_bwd": [_vp, _vp, _vp, _vp, _sz, _i],
    "dl4jb200_pool_fwd": [_vp, _desc_p, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "dl4jb200_pool_bwd": [_vp, _desc_p, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _d],
    "dl4jb200_bn_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _d, _d, _i, _i, _i, _vp],
    "dl4jb200_bn_fwd_noalloc": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _d, _d, _i, _i, _i],
    "dl4jb200_bn_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i],
    "dl4jb200_lrn_fwd": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _d, _d, _d, _d],
    "dl4jb200_lrn_bwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _d, _d, _d, _d],
    "dl4jb200_add_act_fwd": [_vp, _vp, _vp, _vp, _sz, _i],
    "dl4jb200_conv_fwd_bnstats_rows": [_vp, _desc_p, _i, _i, _i, C.POINTER(C.c_int)],
    "dl4jb200_conv_fwd_bnstats": [_vp, _desc_p, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _vp],
    "dl4jb200_bn_fwd_prestats": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _d, _d, _i, _i, _vp, _vp, _vp, _i],
    "dl4jb200_bn_fwd_prestats_maxpool": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _d, _d, _i, _i, _vp, _vp, _i,
                                         _desc_p],
    "dl4jb200_filter_register": [_vp, _vp, _vp, _i, _i, _i, _i],
    "dl4jb200_filters_repack": [_vp],
    "dl4jb200_filters_invalidate": [_vp],
    "dl4jb200_dropout_fwd": [_vp, _vp, _vp, _vp, _sz, _d],
    "dl4jb200_eps_sum_act_bwd": [_vp, _vp, _vp, _vp, _vp, _sz, _i],
    "dl4jb200_conv_bwd_data_fused": [_vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _vp, _i, _i],
    "dl4jb200_accumulate": [_vp, _vp, _vp, _sz],
    "dl4jb200_softmax_nll": [_vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i],
    "dl4jb200_l1l2_score": [_vp, _vp, _sz, _d, _d, _vp],
    "dl4jb200_score_combine": [_vp, _vp, _vp, _vp, _i],
    "dl4jb200_pool_bwd_bnstats": [_vp, _desc_p, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, C.POINTER(C.c_int)],
    "dl4jb200_copy2d": [_vp, _vp, _i, _vp, _i, _i, _i],
    "dl4jb200_grad_normalize": [_vp, _vp, _sz, _i, _d],
    "dl4jb200_updater_apply_segments": [_vp, _vp, _vp, _vp, _sz, _vp, _vp, _i, _vp, _i, _d, _vp],
    "dl4jb200_conv_bwd_data_bnstats_rows": [_vp, _desc_p, _i, _i, _i, C.POINTER(C.c_int)],
    "dl4jb200_conv_bwd_data_bnstats": [_vp, _desc_p, _vp, _vp, _i, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "dl4jb200_bn_bwd_prestats": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i],
    "dl4jb200_deconv2d_fwd": [_vp, _desc_p, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i],
    "dl4jb200_deconv2d_bwd": [_vp, _desc_p, _vp, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i],
    "dl4jb200_bias_act": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i],
    "dl4jb200_depthwise_fwd": [_vp, _desc_p, _i, _vp, _i, _vp, _vp, _i],
    "dl4jb200_depthwise_bwd": [_vp, _desc_p, _i, _vp, _i, _vp, _vp, _i, _vp, _vp, _i],
    "dl4jb200_comm_unique_id": [_vp, _sz],
    "dl4jb200_comm_init": [_vp, _i, _i, _vp, C.POINTER(_vp)],
    "dl4jb200_comm_destroy": [_vp],
    "dl4jb200_comm_allreduce_sum_f32": [_vp, _vp, _sz, _vp],
    "dl4jb200_comm_allreduce_avg_f32": [_vp, _vp, _sz, _vp],
    "dl4jb200_comm_allreduce_avg_spans_f32": [_vp, _vp, _vp, _vp, _i, _vp],
    "dl4jb200_comm_broadcast_f32": [_vp, _vp, _sz, _i, _vp],
    "dl4jb200_spans_scale_copy": [_vp, _vp, _vp, _vp, _vp, _i, _d],
    "dl4jb200_updater_apply_segments_range": [_vp, _vp, _vp, _vp, _sz, _vp, _vp, _i, _vp, _i, _d, _vp, _sz, _sz, _i],
    "dl4jb200_updater_apply": [_vp, _vp, _vp, _vp, _sz, _i, _d, _d, _d, _d, _d, _d],
}
# entry points whose return value is not a status
_SPECIAL = {
    "dl4jb200_last_error": (C.c_char_p, []),
    "dl4jb200_version": (_i, []),
    "dl4jb200_launch_count": (C.c_ulonglong, []),
    "dl4jb200_ctx_stream": (_vp, [_vp]),
    "dl4jb200_ctx_workspace_bytes": (_sz, [_vp]),
    "dl4jb200_ctx_sm_count": (_i, [_vp]),
    "dl4jb200_ctx_workspace_generation": (C.c_ulonglong, [_vp]),
    "dl4jb200_comm_nranks": (_i, [_vp]),
    "dl4jb200_comm_nccl_version": (_i, []),
}

_cdll = None


def cdll():
    """The loaded shared library (raises if it has not been built: there is no other path)."""
    global _cdll
    if _cdll is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m deeplearning4j_b200.build` "
                "(or __graft_entry__.build()). deeplearning4j_b200 has no CPU or eager fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, argt in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.argtypes = argt
            fn.restype = _i
        for name, (rest, argt) in _SPECIAL.items():
            fn = getattr(lib, name)
            fn.argtypes = argt
            fn.restype = rest
        _cdll = lib
    return _cdll


def exported_names():
    return list(PROTOTYPES) + list(_SPECIAL)


def last_error():
    return cdll().dl4jb200_last_error().decode()


def check(status):
    """Maps a C status to the exception the Java helper would raise (or HelperNotSupported for `return null`)."""
    if status == OK:
        return
    msg = last_error()
    if status == NOT_SUPPORTED:
        raise HelperNotSupported(msg)
    if status == ERR_INVALID_INPUT:
        raise DL4JInvalidInputException(msg)
    if status == ERR_INVALID_CONFIG:
        raise DL4JInvalidConfigException(msg)
    raise DL4JB200Error(status, msg)


def call(name, *args):
    check(getattr(cdll(), name)(*args))


def out_size(inp, k, s, p, d=1, mode=MODE_TRUNCATE):
    o, lo, hi = _i(), _i(), _i()
    call("dl4jb200_out_size", inp, k, s, p, d, mode, C.byref(o), C.byref(lo), C.byref(hi))
    return o.value, lo.value, hi.value
