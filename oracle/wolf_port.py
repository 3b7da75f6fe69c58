"""ORACLE (test infrastructure, not product code): ctypes loader for oracle/wolf_port.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this.  `build()` compiles the C restatement with gcc (-O2 -fopenmp) into oracle/_build/.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "wolf_port.c")
LIB = os.path.join(HERE, "_build", "libwolfport.so")


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.run(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", LIB, SRC], check=True)
    return LIB


class _Net(ctypes.Structure):
    _fields_ = [("C", ctypes.c_int), ("T", ctypes.c_int), ("F2", ctypes.c_int), ("N", ctypes.c_int),
                ("l1_factor", ctypes.c_void_p), ("l1_offset", ctypes.c_void_p), ("l1_weight_align", ctypes.c_void_p),
                ("l2_factor", ctypes.c_void_p), ("l2_offset", ctypes.c_void_p), ("l2_weight_reverse", ctypes.c_void_p),
                ("l3_factor", ctypes.c_int32), ("l3_weight", ctypes.c_void_p),
                ("l4_factor", ctypes.c_void_p), ("l4_offset", ctypes.c_void_p), ("l4_weight", ctypes.c_void_p),
                ("l5_factor", ctypes.c_int32), ("l5_bias", ctypes.c_void_p), ("l5_weight", ctypes.c_void_p)]


_lib = None


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.wolf_port_forward.restype = ctypes.c_int
        _lib.wolf_port_forward.argtypes = [ctypes.POINTER(_Net), ctypes.c_void_p, ctypes.c_long, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    return _lib


def forward(arrays, C, T, x, F2=16, N=4, reorder_bn=True, threads=0, return_layers=False):
    """arrays: exported layout (see oracle/golden_np.py); x int8 [B, C, T] -> int8 [B, N]."""
    lib = _load()
    x = np.ascontiguousarray(x, dtype=np.int8)
    assert x.ndim == 3 and x.shape[1:] == (C, T), x.shape
    B = x.shape[0]
    keep = {}

    def ptr(name, dt):
        a = np.ascontiguousarray(np.asarray(arrays[name]).ravel(), dtype=dt)
        keep[name] = a
        return a.ctypes.data

    net = _Net(C, T, F2, N,
               ptr("l1_factor", np.int32), ptr("l1_offset", np.int32), ptr("l1_weight_align", np.int8),
               ptr("l2_factor", np.int32), ptr("l2_offset", np.int32), ptr("l2_weight_reverse", np.int8),
               int(arrays["l3_factor"]), ptr("l3_weight", np.int8),
               ptr("l4_factor", np.int32), ptr("l4_offset", np.int32), ptr("l4_weight", np.int8),
               int(arrays["l5_factor"]), ptr("l5_bias", np.int8), ptr("l5_weight", np.int8))
    out = np.empty((B, N), np.int8)
    T8, T64 = T // 8, T // 64
    taps = np.empty((B, F2 * (T + 2 * T8 + T64)), np.int8) if return_layers else None
    used = lib.wolf_port_forward(ctypes.byref(net), x.ctypes.data, B, int(bool(reorder_bn)), out.ctypes.data,
                                 int(threads), taps.ctypes.data if return_layers else None)
    forward.last_threads = used
    if not return_layers:
        return out
    o = 0
    layers = {}
    for name, n in (("y1", T), ("y2", T8), ("y3", T8), ("y4", T64)):
        layers[name] = taps[:, o:o + F2 * n].reshape(B, F2, n)
        o += F2 * n
    return out, layers


def stream_windows(x, T, hop=16):
    """x int8 [S][C][L] -> every window materialised, [S * W][C][T] with W = (L - T) // hop + 1
    (what a caller of net_model_compute would have to build: one trial per window)."""
    x = np.ascontiguousarray(x, dtype=np.int8)
    S, C, L = x.shape
    W = (L - T) // hop + 1
    v = np.lib.stride_tricks.sliding_window_view(x, T, axis=2)[:, :, ::hop][:, :, :W]   # [S][C][W][T]
    return np.ascontiguousarray(v.transpose(0, 2, 1, 3)).reshape(S * W, C, T), W


def forward_stream(arrays, C, T, x, hop=16, return_layers=False, threads=0):
    """Sliding-window oracle: the integer network on every materialised window.
    -> logits [S][W][N]  (and the per-window layer outputs [S][W][...] if return_layers)."""
    xs, W = stream_windows(x, T, hop)
    S = x.shape[0]
    r = forward(arrays, C, T, xs, threads=threads, return_layers=return_layers)
    if not return_layers:
        return r.reshape(S, W, -1)
    y, layers = r
    return y.reshape(S, W, -1), {k: v.reshape((S, W) + v.shape[1:]) for k, v in layers.items()}
