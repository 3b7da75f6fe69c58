"""ctypes binding of libmibmi.so (include/mibmi.h).  No torch types cross this boundary.

The library is built in-tree by mibminet_b200/_build.py (nvcc, sm_100a).  There is no CPU fallback:
if the library cannot be built or loaded, importing the compute entry points raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Mapping

import numpy as np

from . import _build

OK = 0
ERR_NULL, ERR_SHAPE, ERR_PARAM, ERR_ALIGN, ERR_CUDA, ERR_UNSUPPORTED = -1, -2, -3, -4, -5, -6
LAYOUT_BCT, LAYOUT_BTC = 0, 1
PATH_AUTO, PATH_GENERIC, PATH_FUSED, PATH_STREAM, PATH_FUSED_PIPE = 0, 1, 2, 4, 5
STREAM_AUTO, STREAM_FUSED, STREAM_GATHER = 0, 1, 2
OPT_FORCE_PATH, OPT_HOST_CHUNK, OPT_STREAM_MODE = 1, 3, 4

# every symbol include/mibmi.h declares (tests/test_abi.py checks the header against this list)
SYMBOLS = [
    "mibmi_version", "mibmi_strerror", "mibmi_last_cuda_error", "mibmi_create", "mibmi_destroy",
    "mibmi_set_option", "mibmi_last_path", "mibmi_last_kernel", "mibmi_fused_supported", "mibmi_launch_count",
    "mibmi_forward_int8", "mibmi_forward_int8_taps", "mibmi_forward_int8_host",
    "mibmi_quantize_input_f32", "mibmi_dequantize_logits_f32", "mibmi_debug_magic", "mibmi_debug_magic32",
    "mibmi_debug_pack_pipe", "mibmi_stream_windows", "mibmi_stream_supported", "mibmi_forward_stream_int8", "mibmi_forward_stream_int8_noalloc",
    "mibmi_forward_stream_int8_taps", "mibmi_forward_int8_gather",
    "mibmi_eegnet_create", "mibmi_eegnet_destroy", "mibmi_eegnet_forward_int8", "mibmi_eegnet_forward_int8_taps",
    "mibmi_eegnet_launch_count", "mibmi_eegnet_last_kernel",
]


class MibmiError(RuntimeError):
    def __init__(self, status: int, what: str = ""):
        self.status = status
        msg = lib().mibmi_strerror(status).decode() if _lib is not None else str(status)
        super().__init__(f"{what}: {msg} (status {status})" if what else f"{msg} (status {status})")


class Params(ctypes.Structure):
    """mibmi_params_t"""
    _fields_ = [
        ("C", ctypes.c_int32), ("T", ctypes.c_int32), ("F1", ctypes.c_int32), ("D", ctypes.c_int32),
        ("F2", ctypes.c_int32), ("N", ctypes.c_int32), ("Nf", ctypes.c_int32), ("reorder_bn", ctypes.c_int32),
        ("l1_factor", ctypes.c_void_p), ("l1_offset", ctypes.c_void_p), ("l1_weight_align", ctypes.c_void_p),
        ("l2_factor", ctypes.c_void_p), ("l2_offset", ctypes.c_void_p), ("l2_weight_reverse", ctypes.c_void_p),
        ("l3_factor", ctypes.c_int32), ("l3_weight", ctypes.c_void_p),
        ("l4_factor", ctypes.c_void_p), ("l4_offset", ctypes.c_void_p), ("l4_weight", ctypes.c_void_p),
        ("l5_factor", ctypes.c_int32), ("l5_bias", ctypes.c_void_p), ("l5_weight", ctypes.c_void_p),
    ]


class EEGNetParamsC(ctypes.Structure):
    """mibmi_eegnet_params_t"""
    _fields_ = [
        ("C", ctypes.c_int32), ("T", ctypes.c_int32), ("F1", ctypes.c_int32), ("D", ctypes.c_int32),
        ("F2", ctypes.c_int32), ("N", ctypes.c_int32), ("Nf", ctypes.c_int32),
        ("l1_weight", ctypes.c_void_p), ("l1_factor", ctypes.c_void_p), ("l1_offset", ctypes.c_void_p),
        ("l2_weight", ctypes.c_void_p), ("l2_factor", ctypes.c_void_p), ("l2_offset", ctypes.c_void_p),
        ("l3_weight", ctypes.c_void_p), ("l3_factor", ctypes.c_int32),
        ("l4_weight", ctypes.c_void_p), ("l4_factor", ctypes.c_void_p), ("l4_offset", ctypes.c_void_p),
        ("l5_weight", ctypes.c_void_p), ("l5_bias", ctypes.c_void_p), ("l5_factor", ctypes.c_int32),
    ]


_lib = None


def lib() -> ctypes.CDLL:
    """Build (if stale) and load libmibmi.so."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if os.environ.get("MIBMI_LIB_PATH"):          # experiment builds (tools/build_variant.py)
        path = os.environ["MIBMI_LIB_PATH"]
    elif _build.is_stale():
        if os.environ.get("MIBMI_NO_BUILD") and os.path.exists(path):
            pass
        else:
            path = _build.build_library()
    L = ctypes.CDLL(path)
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    L.mibmi_version.restype = i32
    L.mibmi_strerror.restype = ctypes.c_char_p
    L.mibmi_strerror.argtypes = [i32]
    L.mibmi_last_cuda_error.argtypes = [vp]
    L.mibmi_create.argtypes = [ctypes.POINTER(Params), i32, ctypes.POINTER(vp)]
    L.mibmi_destroy.argtypes = [vp]
    L.mibmi_set_option.argtypes = [vp, i32, i64]
    L.mibmi_last_path.argtypes = [vp]
    L.mibmi_last_kernel.argtypes = [vp]
    L.mibmi_last_kernel.restype = ctypes.c_char_p
    L.mibmi_fused_supported.argtypes = [vp, ctypes.POINTER(ctypes.c_char_p)]
    L.mibmi_launch_count.argtypes = [vp]
    L.mibmi_launch_count.restype = i64
    L.mibmi_forward_int8.argtypes = [vp, vp, i64, i32, vp, vp]
    L.mibmi_forward_int8_taps.argtypes = [vp, vp, i64, i32, vp, vp, vp, vp, vp, vp]
    L.mibmi_forward_int8_host.argtypes = [vp, vp, i64, i32, vp]
    L.mibmi_forward_int8_gather.argtypes = [vp, vp, i64, i32, ctypes.POINTER(vp), ctypes.c_int32, i64, vp]
    L.mibmi_stream_windows.argtypes = [vp, i64, ctypes.c_int32]
    L.mibmi_stream_windows.restype = i64
    L.mibmi_stream_supported.argtypes = [vp, ctypes.POINTER(ctypes.c_char_p)]
    L.mibmi_forward_stream_int8.argtypes = [vp, vp, i64, i64, ctypes.c_int32, vp, vp]
    L.mibmi_forward_stream_int8_noalloc.argtypes = [vp, vp, i64, i64, ctypes.c_int32, vp, vp]
    L.mibmi_forward_stream_int8_taps.argtypes = [vp, vp, i64, i64, ctypes.c_int32, vp, vp, vp, vp, vp, vp]
    L.mibmi_quantize_input_f32.argtypes = [vp, i64, ctypes.c_float, vp, vp]
    L.mibmi_dequantize_logits_f32.argtypes = [vp, i64, ctypes.c_float, vp, vp]
    L.mibmi_debug_magic.argtypes = [ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
    L.mibmi_debug_magic32.argtypes = [ctypes.c_int32, i64, ctypes.POINTER(ctypes.c_int32)]
    L.mibmi_debug_pack_pipe.argtypes = [ctypes.POINTER(Params), vp, i64, ctypes.POINTER(ctypes.c_char_p)]
    L.mibmi_debug_pack_pipe.restype = i64
    L.mibmi_eegnet_create.argtypes = [ctypes.POINTER(EEGNetParamsC), i32, ctypes.POINTER(vp)]
    L.mibmi_eegnet_destroy.argtypes = [vp]
    L.mibmi_eegnet_forward_int8.argtypes = [vp, vp, i64, vp, vp]
    L.mibmi_eegnet_forward_int8_taps.argtypes = [vp, vp, i64, vp, vp, vp, vp, vp]
    L.mibmi_eegnet_launch_count.argtypes = [vp]
    L.mibmi_eegnet_launch_count.restype = i64
    L.mibmi_eegnet_last_kernel.argtypes = [vp]
    L.mibmi_eegnet_last_kernel.restype = ctypes.c_char_p
    _lib = L
    return L


def make_params(arrays: Mapping[str, np.ndarray], C: int, T: int, F1: int = 16, D: int = 1, N: int = 4,
                reorder_bn: bool = True):
    """Exported arrays (gen_net_header.py names without 'net_') -> (Params, keepalive list)."""
    keep = []

    def ptr(name, dt):
        a = np.ascontiguousarray(np.asarray(arrays[name]).ravel(), dtype=dt)
        keep.append(a)
        return a.ctypes.data

    p = Params(C, T, F1, D, F1 * D, N, 64, int(bool(reorder_bn)),
               ptr("l1_factor", np.int32), ptr("l1_offset", np.int32), ptr("l1_weight_align", np.int8),
               ptr("l2_factor", np.int32), ptr("l2_offset", np.int32), ptr("l2_weight_reverse", np.int8),
               int(arrays["l3_factor"]), ptr("l3_weight", np.int8),
               ptr("l4_factor", np.int32), ptr("l4_offset", np.int32), ptr("l4_weight", np.int8),
               int(arrays["l5_factor"]), ptr("l5_bias", np.int8), ptr("l5_weight", np.int8))
    return p, keep


def check(status: int, what: str = ""):
    if status != OK:
        raise MibmiError(int(status), what)


def debug_magic(d: int):
    m, s = ctypes.c_int32(), ctypes.c_int32()
    st = lib().mibmi_debug_magic(int(d), ctypes.byref(m), ctypes.byref(s))
    return st, m.value, s.value


def debug_magic32(d: int, xmax: int):
    """-> (status, magic) of the shift-free layer-1 division (mibmi_debug_magic32)."""
    m = ctypes.c_int32()
    st = lib().mibmi_debug_magic32(int(d), int(xmax), ctypes.byref(m))
    return st, m.value


def _debug_pack(fn, arrays, C, T, reorder_bn):
    p, keep = make_params(arrays, C, T, reorder_bn=reorder_bn)
    L = lib()
    size = fn(ctypes.byref(p), None, 0, None)
    if size < 0:
        return None, L.mibmi_strerror(int(size)).decode()
    buf = (ctypes.c_uint8 * size)()
    why = ctypes.c_char_p()
    r = fn(ctypes.byref(p), buf, size, ctypes.byref(why))
    if r < 0:
        return None, (why.value or b"").decode()
    return bytes(buf), ""


def debug_pack_pipe(arrays, C, T, reorder_bn=True):
    """-> (bytes of the packed PipeNet of the pipeline kernel | None, reason)."""
    return _debug_pack(lib().mibmi_debug_pack_pipe, arrays, C, T, reorder_bn)
