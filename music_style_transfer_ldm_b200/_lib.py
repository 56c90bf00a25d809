"""ctypes binding of libmstk.so (the C ABI declared in include/mstk.h).

The library is built in-tree by ``__graft_entry__.build()`` /
``python -m music_style_transfer_ldm_b200.build``.  There is NO fallback: if the
shared object is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmstk.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_TANH = 0, 1, 2
BACKEND_AUTO, BACKEND_SIMT, BACKEND_TCGEN05 = 0, 1, 2

c_void_p, c_int, c_float, c_size_t, c_int64 = C.c_void_p, C.c_int32, C.c_float, C.c_size_t, C.c_int64


class MstkError(RuntimeError):
    pass


class DdimArgs(C.Structure):
    _fields_ = [("x_cur", c_void_p), ("x_next", c_void_p), ("x_next_lp", c_void_p), ("x0_pred", c_void_p),
                ("sab_t", c_float), ("s1m_t", c_float), ("sab_n", c_float), ("s1m_n", c_float), ("eta", c_float)]


class ConvArgs(C.Structure):
    _fields_ = [("x", c_void_p), ("wpacked", c_void_p), ("bias", c_void_p), ("y", c_void_p), ("add", c_void_p),
                ("add_nc", c_void_p), ("stats", c_void_p), ("ddim", C.POINTER(DdimArgs)),
                ("workspace", c_void_p), ("workspace_bytes", c_size_t),
                ("n", c_int), ("h", c_int), ("w", c_int), ("cin", c_int), ("cout", c_int),
                ("ksize", c_int), ("stride", c_int), ("transposed", c_int), ("act", c_int),
                ("out_scale", c_float), ("out_shift", c_float),
                ("dtype", c_int), ("out_dtype", c_int), ("backend", c_int)]


class BnArgs(C.Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p), ("stats", c_void_p), ("gamma", c_void_p), ("beta", c_void_p),
                ("running_mean", c_void_p), ("running_var", c_void_p), ("save_mean", c_void_p),
                ("save_invstd", c_void_p), ("count", c_int64), ("rows", c_int64), ("c", c_int), ("eps", c_float),
                ("momentum", c_float), ("act", c_int), ("dtype", c_int), ("out_dtype", c_int)]


class AttnArgs(C.Structure):
    _fields_ = [("q", c_void_p), ("k", c_void_p), ("v", c_void_p), ("out", c_void_p),
                ("b", c_int), ("l", c_int), ("s", c_int), ("e", c_int), ("heads", c_int),
                ("q_stride", c_int), ("k_stride", c_int), ("v_stride", c_int), ("out_stride", c_int),
                ("dtype", c_int)]


class XattnArgs(C.Structure):
    _fields_ = [("x", c_void_p), ("kvp", c_void_p), ("bias", c_void_p), ("out", c_void_p),
                ("b", c_int), ("l", c_int), ("s", c_int), ("e", c_int), ("heads", c_int), ("ctot", c_int),
                ("kv_static", c_int), ("dtype", c_int)]


class MelArgs(C.Structure):
    _fields_ = [("wave", c_void_p), ("out", c_void_p), ("db_out", c_void_p), ("mel_fb", c_void_p),
                ("work", c_void_p), ("work_bytes", c_size_t),
                ("b", c_int), ("len", c_int), ("n_mels", c_int), ("crop_frames", c_int), ("band_ranges", c_void_p)]


class MelToStftArgs(C.Structure):
    _fields_ = [("image", c_void_p), ("pinv", c_void_p), ("mel_fb", c_void_p), ("band_ranges", c_void_p),
                ("mag", c_void_p), ("pg_norm", c_void_p),
                ("b", c_int), ("n_mels", c_int), ("frames", c_int), ("max_db", c_float), ("pg_iters", c_int), ("pg_step", c_float)]


class GriffinLimArgs(C.Structure):
    _fields_ = [("mag", c_void_p), ("phase_u", c_void_p), ("audio", c_void_p), ("work", c_void_p), ("work_bytes", c_size_t),
                ("b", c_int), ("frames", c_int), ("n_iter", c_int), ("momentum", c_float)]


class PackEntry(C.Structure):
    _fields_ = [("src", c_void_p), ("dst", c_void_p), ("cout", c_int), ("cin", c_int), ("kk", c_int), ("mode", c_int),
                ("dtype", c_int), ("first_block", c_int)]


class PeerCtx(C.Structure):
    _fields_ = [("bufs", c_void_p * 8), ("rank", c_int), ("world", c_int)]


class WgradArgs(C.Structure):
    _fields_ = [("x", c_void_p), ("dy", c_void_p), ("dw", c_void_p),
                ("n", c_int), ("h", c_int), ("w", c_int), ("cin", c_int), ("cout", c_int), ("ksize", c_int),
                ("stride", c_int), ("transposed", c_int), ("x_dtype", c_int), ("dy_dtype", c_int)]


class ActBwdArgs(C.Structure):
    _fields_ = [("dy", c_void_p), ("dy2", c_void_p), ("y", c_void_p), ("g", c_void_p), ("dbias", c_void_p),
                ("rows", c_int64), ("c", c_int), ("act", c_int), ("y_scale", c_float), ("y_shift", c_float),
                ("dy_dtype", c_int), ("y_dtype", c_int), ("g_dtype", c_int)]


class BnBwdArgs(C.Structure):
    _fields_ = [("dy", c_void_p), ("v", c_void_p), ("dv", c_void_p), ("mean", c_void_p), ("invstd", c_void_p),
                ("gamma", c_void_p), ("beta", c_void_p), ("sums", c_void_p), ("dgamma", c_void_p), ("dbeta", c_void_p),
                ("rows", c_int64), ("count", c_int64), ("c", c_int), ("act", c_int), ("dy_dtype", c_int),
                ("v_dtype", c_int), ("dv_dtype", c_int)]


class AttnBwdArgs(C.Structure):
    _fields_ = [("q", c_void_p), ("k", c_void_p), ("v", c_void_p), ("d_out", c_void_p), ("dq", c_void_p), ("dk", c_void_p),
                ("dv", c_void_p), ("b", c_int), ("l", c_int), ("s", c_int), ("e", c_int), ("heads", c_int),
                ("q_stride", c_int), ("k_stride", c_int), ("v_stride", c_int), ("do_stride", c_int), ("dq_stride", c_int),
                ("dk_stride", c_int), ("dv_stride", c_int), ("dtype", c_int)]


# name -> (restype, argtypes); every symbol include/mstk.h declares
SIGNATURES = {
    "mstk_version": (c_int, []),
    "mstk_last_error": (C.c_char_p, []),
    "mstk_tcgen05_available": (c_int, []),
    "mstk_conv2d_fwd": (c_int, [C.POINTER(ConvArgs), c_void_p]),
    "mstk_conv2d_workspace_bytes": (c_size_t, [C.POINTER(ConvArgs)]),
    "mstk_pack_conv_weight": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_pack_convt_weight": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_nchw_to_nhwc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_nhwc_to_nchw": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_u8_gather_crop": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_cast": (c_int, [c_void_p, c_int, c_void_p, c_int, c_size_t, c_void_p]),
    "mstk_bn_apply": (c_int, [C.POINTER(BnArgs), c_void_p]),
    "mstk_time_embed": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "mstk_attention_fwd": (c_int, [C.POINTER(AttnArgs), c_void_p]),
    "mstk_split3": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_int, c_void_p]),
    "mstk_xattn_ctot": (c_int, [c_int, c_int]),
    "mstk_xattn_compose": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "mstk_xattn_fwd": (c_int, [C.POINTER(XattnArgs), c_void_p]),
    "mstk_q_sample": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "mstk_predict_x0": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "mstk_ddim_step": (c_int, [C.POINTER(DdimArgs), c_void_p, c_size_t, c_void_p]),
    "mstk_mse": (c_int, [c_void_p, c_void_p, c_int, c_size_t, c_void_p, c_void_p, c_void_p]),
    "mstk_kl": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    "mstk_feature_std_mse": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, c_float, c_void_p, c_void_p]),
    "mstk_sample_stats": (c_int, [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "mstk_feature_mse3": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                  c_void_p, c_void_p, c_void_p]),
    "mstk_feature_fused3": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p,
                                    c_void_p]),
    "mstk_feature_first3": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p,
                                    c_void_p, c_void_p]),
    "mstk_maxpool2x2": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_pack_conv_weight_dgrad": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mstk_pack_blocks": (c_int, [c_int, c_int, c_int]),
    "mstk_peer_buffer_bytes": (c_size_t, []),
    "mstk_peer_alloc": (c_int, [C.POINTER(c_void_p), c_void_p]),
    "mstk_peer_open": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "mstk_peer_close": (c_int, [c_void_p]),
    "mstk_peer_free": (c_int, [c_void_p]),
    "mstk_peer_allreduce_f64": (c_int, [c_void_p, c_int, c_int, C.POINTER(PeerCtx), c_void_p]),
    "mstk_pack_multi": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "mstk_conv2d_wgrad": (c_int, [C.POINTER(WgradArgs), c_void_p]),
    "mstk_act_bwd": (c_int, [C.POINTER(ActBwdArgs), c_void_p]),
    "mstk_bn_bwd_reduce": (c_int, [C.POINTER(BnBwdArgs), c_void_p]),
    "mstk_bn_bwd_apply": (c_int, [C.POINTER(BnBwdArgs), c_void_p]),
    "mstk_attention_bwd": (c_int, [C.POINTER(AttnBwdArgs), c_void_p]),
    "mstk_time_embed_train": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "mstk_gelu_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "mstk_sum_hw": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int64, c_int, c_void_p, c_void_p]),
    "mstk_add_bcast": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_void_p]),
    "mstk_scale_dev": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "mstk_mse_grad": (c_int, [c_void_p, c_void_p, c_float, c_void_p, c_int, c_size_t, c_void_p]),
    "mstk_eps_grad": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int, c_int, c_int64, c_void_p]),
    "mstk_adam_step_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float, c_float, c_float, c_void_p, c_float, c_void_p]),
    "mstk_adamw_step_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float, c_float, c_float, c_float, c_void_p, c_float, c_void_p]),
    "mstk_kl_grad": (c_int, [c_void_p, c_void_p, c_int, c_float, c_void_p, c_size_t, c_void_p]),
    "mstk_adam_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float, c_float, c_float, c_int64, c_float, c_void_p]),
    "mstk_mel_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mstk_mel_band_ranges": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "mstk_mel_frontend": (c_int, [C.POINTER(MelArgs), c_void_p]),
    "mstk_mel_static_scratch_bytes": (c_size_t, []),
    "mstk_mel_frontend_static": (c_int, [C.POINTER(MelArgs), c_void_p]),
    "mstk_mel_to_stft": (c_int, [C.POINTER(MelToStftArgs), c_void_p]),
    "mstk_db_image_to_mag": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p]),
    "mstk_griffinlim_workspace_bytes": (c_size_t, [c_int, c_int]),
    "mstk_griffinlim": (c_int, [C.POINTER(GriffinLimArgs), c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load libmstk.so and bind every declared symbol.  Raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise MstkError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for the kernels.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().mstk_last_error()
        raise MstkError(f"{what or 'mstk call'} failed (code {rc}): {msg.decode() if msg else '?'}")
