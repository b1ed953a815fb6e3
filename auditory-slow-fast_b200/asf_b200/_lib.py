"""ctypes binding of libasf_b200.so (include/asf_b200.h).

The product path has NO fallback: if the shared library is missing or a call fails, a RuntimeError is
raised.  Build it with ``auditory-slow-fast_b200/build.sh`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libasf_b200.so")

F32, BF16 = 0, 1
MAX_TAPS = 8


class Epilogue(C.Structure):
    _fields_ = [
        ("scale", C.c_void_p), ("shift", C.c_void_p),
        ("res", C.c_void_p), ("ld_res", C.c_int64),
        ("mask", C.c_void_p), ("ld_mask", C.c_int64),
        ("stats", C.c_void_p), ("stats_stride", C.c_int32),
        ("relu", C.c_int32), ("accumulate", C.c_int32),
        ("stats_ws", C.c_void_p), ("stats_ticket", C.c_void_p),
    ]


STATS_TICKETS = 128


def stats_ws_floats(channels):
    """ASF_STATS_WS_FLOATS(C): floats of per-CTA partial-sum workspace a statistics producer may use."""
    return 640 * max(int(channels), 256)


class ConvGeom(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("Tin", C.c_int32), ("Fin", C.c_int32), ("Cin", C.c_int32),
        ("Tout", C.c_int32), ("Fout", C.c_int32), ("Cout", C.c_int32),
        ("kt", C.c_int32), ("kf", C.c_int32), ("st", C.c_int32), ("sf", C.c_int32),
        ("pt", C.c_int32), ("pf", C.c_int32),
        ("ld_in", C.c_int64), ("ld_out", C.c_int64),
    ]


class TapGemm(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("nd", C.c_int32),
        ("a_dim", C.c_int64 * 4), ("a_stride", C.c_int64 * 4),
        ("K", C.c_int32),
        ("m_ext", C.c_int32 * 3), ("m_box", C.c_int32 * 3), ("m_to_a", C.c_int32 * 3),
        ("ntaps", C.c_int32),
        ("a_off", (C.c_int32 * 4) * MAX_TAPS), ("b_tap", C.c_int32 * MAX_TAPS),
        ("bw", C.c_void_p), ("N", C.c_int32), ("ntaps_w", C.c_int32),
        ("out", C.c_void_p), ("out_dtype", C.c_int32), ("ld_out", C.c_int64), ("o_stride", C.c_int64 * 3),
        ("zero_invalid", C.c_int32),
        ("epi", Epilogue),
        ("y", C.c_void_p), ("y_dim", C.c_int64 * 4), ("y_stride", C.c_int64 * 4),
        ("y_nd", C.c_int32), ("y_m_to_a", C.c_int32 * 3),
        ("dwp", C.c_void_p), ("ksplit", C.c_int32), ("block_n", C.c_int32), ("max_ctas", C.c_int32),
    ]


class UnpackDesc(C.Structure):
    _fields_ = [("dwp", C.c_void_p), ("dw", C.c_void_p), ("rscale", C.c_void_p),
                ("O", C.c_int32), ("I", C.c_int32), ("taps", C.c_int32), ("ksplit", C.c_int32),
                ("accumulate", C.c_int32), ("pad_", C.c_int32), ("chunk0", C.c_int64)]


class PackDesc(C.Structure):
    _fields_ = [("w", C.c_void_p), ("packed", C.c_void_p), ("packed_t", C.c_void_p), ("rscale", C.c_void_p),
                ("O", C.c_int32), ("I", C.c_int32), ("taps", C.c_int32), ("dtype", C.c_int32), ("chunk0", C.c_int64)]


PACK_CHUNK = 2048

SPECAUG_MAX_MASKS = 2


class SpecAugParams(C.Structure):
    _fields_ = [("coef", C.c_float * 3), ("fmask", (C.c_int32 * 2) * SPECAUG_MAX_MASKS),
                ("tmask", (C.c_int32 * 2) * SPECAUG_MAX_MASKS), ("pad_", C.c_int32)]


class SgdDesc(C.Structure):
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("n", C.c_int64), ("wd", C.c_float),
                ("group", C.c_int32), ("chunk0", C.c_int64)]


SGD_CHUNK = 4096

_lib = None


def lib():
    """Loads the shared library once; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "asf_b200: CUDA extension %s is missing -- build it with auditory-slow-fast_b200/build.sh "
                "(there is no CPU / PyTorch fallback)" % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _lib.asf_last_error.restype = C.c_char_p
        _declare(_lib)
    return _lib


_P, _I, _L, _F, _D = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double

# name -> argtypes (every function returns int)
SIGNATURES = {
    "asf_version": [],
    "asf_device_is_sm100": [],
    "asf_prof_start": [_I],
    "asf_prof_stop": [_P, _P, _P],
    "asf_prof_record": [_I, _P],
    "asf_logmel_tables": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "asf_logmel_fwd": [_P, _L, _I, _I, _I, _I, _I, _I, _F, _P, _P, _P, _I, _P, _I, _P],
    "asf_logmel_tc_supported": [_I, _I, _I],
    "asf_logmel_tc_btab_floats": [_I],
    "asf_logmel_tc_tables": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "asf_logmel_fwd_tc": [_P, _L, _I, _I, _I, _I, _I, _I, _F, _P, _P, _P, _P, _I, _P],
    "asf_specaug_params": [_P, _I, _I, _I, C.c_ulonglong, _P, _I, _I, _I, _I, _I, _P, _P],
    "asf_specaug_apply": [_P, _P, _P, _I, _I, _I, _P],
    "asf_peer_allreduce_f64": [_P, _I, _P, _I, _I, C.c_longlong, C.c_longlong, _P, _P],
    "asf_gather_clips": [_P, _P, _P, _I, _I, _P, _L, _P, _P],
    "asf_topk_counts": [_P, _I, _L, _P, _P, _I, _L, _P, _I, _I, _I, _P, _P],
    "asf_ensemble_scatter": [_P, _I, _L, _P, _P, _I, _I, _I, _P, _P, _P, _P],
    "asf_ensemble_reduce": [_P, _I, _I, _I, _I, _P, _P],
    "asf_conv_fwd_simt": [_P, _I, _P, _I, _P, _I, C.POINTER(ConvGeom), C.POINTER(Epilogue), _P],
    "asf_conv_dgrad_simt": [_P, _I, _P, _I, _P, _I, C.POINTER(ConvGeom), C.POINTER(Epilogue), _P],
    "asf_conv_wgrad_simt": [_P, _I, _P, _I, _P, C.POINTER(ConvGeom), _P, _I, _P, _P],
    "asf_tapgemm": [C.POINTER(TapGemm), _I, _P],
    "asf_tapgemm_probe": [_P],
    "asf_pack_weight": [_P, _P, _I, _I, _I, _I, _P, _P],
    "asf_pack_weight_t": [_P, _P, _I, _I, _I, _I, _P, _P],
    "asf_pack_weights_multi": [_P, _I, C.c_longlong, _P],
    "asf_sgd_multi": [_P, _I, C.c_longlong, _P, _F, _F, _I, _I, _P],
    "asf_unpack_wgrad": [_P, _P, _I, _I, _I, _P, _I, _I, _P],
    "asf_unpack_wgrad_multi": [_P, _I, C.c_longlong, _P],
    "asf_tapgemm_wgrad_ksplit": [C.POINTER(TapGemm)],
    "asf_channel_stats": [_P, _I, _L, _I, _L, _P, _P, _P, _P],
    "asf_channel_stats_noalloc": [_P, _I, _L, _I, _L, _P, _P],
    "asf_bn_finalize": [_P, _I, _D, _P, _P, _F, _F, _P, _P, _P, _P, _P, _P, _P, _P],
    "asf_bn_fold": [_P, _P, _P, _P, _F, _I, _P, _P, _P],
    "asf_bn_train_fwd": [_P, _L, _P, _L, _P, _L, _I, _L, _I, _P, _D, _P, _P, _F, _F, _P, _P, _P, _P, _P, _P, _P, _I, _P],
    "asf_bn_apply": [_P, _L, _P, _L, _P, _L, _I, _L, _I, _P, _P, _I, _P],
    "asf_bn_bwd_reduce": [_P, _L, _P, _L, _I, _L, _I, _P, _P, _P, _P, _P, _P],
    "asf_bn_bwd_apply": [_P, _L, _P, _L, _P, _L, _I, _L, _I, _P, _P, _P, _P, _D, _P, _P, _I, _P],
    "asf_relu_bwd": [_P, _L, _P, _L, _P, _L, _P, _L, _I, _L, _I, _P],
    "asf_bn_relu_maxpool_fwd": [_P, _P, _L, _I, _I, _I, _I, _I, _P, _P, _P, _P],
    "asf_bn_relu_maxpool_bwd": [_P, _P, _L, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P],
    "asf_avgpool_fwd": [_P, _I, _I, _I, _I, _P, _L, _I, _P],
    "asf_avgpool_bwd": [_P, _L, _I, _P, _P, _I, _I, _I, _I, _P],
    "asf_sgemm": [_P, _P, _P, _I, _I, _I, _I, _I, _L, _L, _L, _F, _F, _P, _P],
    "asf_softmax_rows": [_P, _I, _I, _L, _P],
    "asf_scale_mask": [_P, _P, _P, _L, _F, _P],
    "asf_dropout_apply": [_P, _P, _P, _P, _L, _F, _P],
    "asf_head_fwd": [_P, _L, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "asf_softmax_ce": [_I, _P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P],
    "asf_head_bwd": [_I, _P, _P, _P, _P, _P, _L, _P, _F, _I, _I, _P, _P, _P, _P, _P, _L, _P],
    "asf_gather_rows": [_P, _P, _P, _I, _I, _I, _I, _P],
    "asf_stem_wgrad": [_P, _P, _I, _P, C.POINTER(ConvGeom), _P],
    "asf_stem_wgrad_fused": [_P, _P, _I, _P, _P, _I, _P, C.POINTER(ConvGeom), _P],
}


def _declare(l):
    for name, args in SIGNATURES.items():
        fn = getattr(l, name)
        fn.argtypes = args
        fn.restype = C.c_int
    l.asf_launch_count.argtypes = []
    l.asf_launch_count.restype = C.c_longlong
    l.asf_narrow_conv_launches.argtypes = []
    l.asf_narrow_conv_launches.restype = C.c_longlong
    # queries returning a count (>= 1) or a negative error code
    l.asf_stem_wgrad_ksplit.argtypes = [C.POINTER(ConvGeom), _I]
    l.asf_stem_wgrad_ksplit.restype = C.c_int
    l.asf_stem_wgrad_fused_rows.argtypes = [C.POINTER(ConvGeom), _I]
    l.asf_stem_wgrad_fused_rows.restype = C.c_int
    l.asf_conv_wgrad_simt_ksplit.argtypes = [C.POINTER(ConvGeom)]
    l.asf_conv_wgrad_simt_ksplit.restype = C.c_int


def call(name, *args):
    """Calls an asf_* entry point; raises RuntimeError with the library's message on failure."""
    l = lib()
    rc = getattr(l, name)(*args)
    if rc != 0:
        msg = l.asf_last_error()
        raise RuntimeError("%s failed (code %d): %s" % (name, rc, msg.decode() if msg else ""))
    return rc


_SM100_OK = set()


def require_cuda_sm100():
    if not torch.cuda.is_available():
        raise RuntimeError("asf_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    dev = torch.cuda.current_device()
    if dev in _SM100_OK:
        return
    l = lib()
    ok = l.asf_device_is_sm100()
    if ok != 1:
        raise RuntimeError("asf_b200 kernels are built for sm_100a only; current device is not sm_100")
    _SM100_OK.add(dev)


def dtype_code(dt):
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise ValueError("unsupported dtype %s (float32 / bfloat16 only)" % dt)


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def ptr_array(tensors):
    """HOST array of device pointers (NULL for None) for the per-head arguments of the asf_head_* entry points."""
    return (C.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])


def int_array(vals):
    return (C.c_int * len(vals))(*[int(v) for v in vals])


def i64_array(vals):
    return (C.c_int64 * len(vals))(*[int(v) for v in vals])


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
