"""Host side of the int8 edgeEEGNet / MI-BMInet path: one `Int8Net` = one mibmi handle.

PyTorch is used for device memory and streams only; all compute goes through the C-ABI
(include/mibmi.h) into the hand-written sm_100a kernels.  No CPU fallback exists.

The Python-level counterpart in the reference is `GoldenModel.__call__`
(edge-eegnet_wolf/python_utils/golden_model.py:104-107: one int [C, T] trial -> int [N] logits); the
C-level one is `net_model_compute` (src/cl/net/model.c:42).  `Int8Net.forward` is the batched form.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Mapping, Optional, Union

import numpy as np
import torch

from . import _lib
from .convert import WolfParams, align4, params_from_export, params_from_state_dict

_LAYOUTS = {"bct": _lib.LAYOUT_BCT, "btc": _lib.LAYOUT_BTC}
_PATHS = {"auto": _lib.PATH_AUTO, "generic": _lib.PATH_GENERIC, "fused": _lib.PATH_FUSED,
          "fused_pipe": _lib.PATH_FUSED_PIPE}
_PATH_NAMES = {v: k for k, v in _PATHS.items()}
_PATH_NAMES[_lib.PATH_STREAM] = "stream"
_STREAM_MODES = {"auto": _lib.STREAM_AUTO, "fused": _lib.STREAM_FUSED, "gather": _lib.STREAM_GATHER}


def _device_index(device) -> int:
    d = torch.device(device if device is not None else "cuda")
    if d.type != "cuda":
        raise ValueError("mibminet_b200 runs on CUDA devices only (there is no CPU path)")
    return d.index if d.index is not None else torch.cuda.current_device()


class Int8Net:
    """Integer network resident on one GPU."""

    def __init__(self, params: Union[WolfParams, Mapping[str, np.ndarray]], C: Optional[int] = None,
                 T: Optional[int] = None, device=None, reorder_bn: bool = True, F1: int = 16, D: int = 1, N: int = 4):
        if isinstance(params, WolfParams):
            self.params: Optional[WolfParams] = params
            arrays = params.to_wolf_arrays()
            C, T, F1, D, N = params.C, params.T, params.F1, params.D, params.N
        else:
            self.params = None
            arrays = dict(params)
            if C is None or T is None:
                raise ValueError("C and T are required with raw exported arrays")
        self.arrays: Dict[str, np.ndarray] = arrays
        self.C, self.T, self.F1, self.D, self.F2, self.N = int(C), int(T), int(F1), int(D), int(F1 * D), int(N)
        self.T8, self.T64 = self.T // 8, (self.T // 8) // 8
        self.reorder_bn = bool(reorder_bn)
        self.device_index = _device_index(device)
        self.device = torch.device("cuda", self.device_index)
        self._L = _lib.lib()
        p, keep = _lib.make_params(arrays, self.C, self.T, F1=self.F1, D=self.D, N=self.N, reorder_bn=self.reorder_bn)
        h = ctypes.c_void_p()
        _lib.check(self._L.mibmi_create(ctypes.byref(p), self.device_index, ctypes.byref(h)), "mibmi_create")
        del keep
        self._h = h

    # ---- constructors --------------------------------------------------------------------------
    @classmethod
    def from_state_dict(cls, sd: Mapping, C: int, T: int, F1: int = 16, D: int = 1, N: int = 4, **kw) -> "Int8Net":
        """QuantLab edgeEEGNet state_dict / net.npz mapping (convert_torch_format.py maths)."""
        return cls(params_from_state_dict(sd, C=C, T=T, F1=F1, D=D, N=N), **kw)

    @classmethod
    def from_export(cls, net_npz: str, config_json: str, **kw) -> "Int8Net":
        """QuantLab/export_net_data.py output: net.npz + config.json."""
        return cls(params_from_export(net_npz, config_json), **kw)

    # ---- lifetime --------------------------------------------------------------------------------
    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._L.mibmi_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- options / introspection ---------------------------------------------------------------
    def set_path(self, path: str):
        _lib.check(self._L.mibmi_set_option(self._h, _lib.OPT_FORCE_PATH, _PATHS[path]), "set_path")

    def set_host_chunk(self, n: int):
        _lib.check(self._L.mibmi_set_option(self._h, _lib.OPT_HOST_CHUNK, int(n)), "set_host_chunk")

    @property
    def last_path(self) -> str:
        return _PATH_NAMES.get(self._L.mibmi_last_path(self._h), "none")

    @property
    def last_kernel(self) -> str:
        """Name of the __global__ function the most recent forward launched."""
        return (self._L.mibmi_last_kernel(self._h) or b"").decode()

    @property
    def launch_count(self) -> int:
        return int(self._L.mibmi_launch_count(self._h))

    def fused_supported(self):
        why = ctypes.c_char_p()
        ok = self._L.mibmi_fused_supported(self._h, ctypes.byref(why))
        return bool(ok), (why.value or b"").decode()

    def trial_bytes(self, layout: str = "bct") -> int:
        return self.C * self.T if layout == "bct" else self.T * align4(self.C)

    # ---- compute -----------------------------------------------------------------------------------
    def _check_x(self, x: torch.Tensor, layout: str):
        if x.dtype != torch.int8:
            raise TypeError("x must be int8 (use quantize_input for float trials)")
        if not x.is_cuda or x.device.index != self.device_index:
            raise ValueError(f"x must live on {self.device}")
        want = (self.C, self.T) if layout == "bct" else (self.T, align4(self.C))
        if x.dim() == 4 and x.shape[1] == 1:      # module order [B, 1, C, T]
            x = x[:, 0]
        if x.dim() != 3 or tuple(x.shape[1:]) != want:
            raise ValueError(f"x must be [B, {want[0]}, {want[1]}] for layout '{layout}', got {tuple(x.shape)}")
        return x.contiguous()

    def forward(self, x: torch.Tensor, layout: str = "bct", out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """int8 trials [B, C, T] ('bct') or [B, T, C_ALIGN] ('btc', Wolf order) -> int8 logits [B, N].
        Asynchronous on the current torch CUDA stream."""
        x = self._check_x(x, layout)
        B = x.shape[0]
        if out is None:
            out = torch.empty((B, self.N), dtype=torch.int8, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_forward_int8(self._h, x.data_ptr(), B, _LAYOUTS[layout], out.data_ptr(), stream),
                   "mibmi_forward_int8")
        return out

    __call__ = forward

    def forward_gather(self, x: torch.Tensor, out_ptrs, row0: int, layout: str = "bct"):
        """Forward fused with the all-gather of a sharded batch: the logits of the local trials go to rows
        [row0, row0 + B) of every buffer in `out_ptrs` (device pointers; own buffer first, then the peer-mapped
        buffers of the other ranks).  See mibmi_forward_int8_gather."""
        x = self._check_x(x, layout)
        arr = (ctypes.c_void_p * len(out_ptrs))(*[int(p) for p in out_ptrs])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_forward_int8_gather(self._h, x.data_ptr(), x.shape[0], _LAYOUTS[layout], arr,
                                                     len(out_ptrs), int(row0), stream), "mibmi_forward_int8_gather")

    def forward_taps(self, x: torch.Tensor, layout: str = "bct"):
        """-> (logits, {'y1','y2','y3','y4'}) : inter-layer int8 activations (generic path)."""
        x = self._check_x(x, layout)
        B = x.shape[0]
        mk = lambda n: torch.empty((B, self.F2, n), dtype=torch.int8, device=self.device)
        out = torch.empty((B, self.N), dtype=torch.int8, device=self.device)
        taps = {"y1": mk(self.T), "y2": mk(self.T8), "y3": mk(self.T8), "y4": mk(self.T64)}
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_forward_int8_taps(self._h, x.data_ptr(), B, _LAYOUTS[layout], out.data_ptr(),
                                                   taps["y1"].data_ptr(), taps["y2"].data_ptr(),
                                                   taps["y3"].data_ptr(), taps["y4"].data_ptr(), stream),
                   "mibmi_forward_int8_taps")
        return out, taps

    # ---- sliding-window scoring of continuous recordings (BASELINE config 5) -----------------------
    def stream_supported(self):
        why = ctypes.c_char_p()
        ok = self._L.mibmi_stream_supported(self._h, ctypes.byref(why))
        return bool(ok), (why.value or b"").decode()

    def set_stream_mode(self, mode: str):
        _lib.check(self._L.mibmi_set_option(self._h, _lib.OPT_STREAM_MODE, _STREAM_MODES[mode]), "set_stream_mode")

    def stream_windows(self, L: int, hop: int = 16) -> int:
        return (int(L) - self.T) // int(hop) + 1

    def _check_stream(self, x: torch.Tensor):
        if x.dtype != torch.int8:
            raise TypeError("x must be int8")
        if not x.is_cuda or x.device.index != self.device_index:
            raise ValueError(f"x must live on {self.device}")
        if x.dim() != 3 or x.shape[1] != self.C or x.shape[2] < self.T:
            raise ValueError(f"x must be [S, {self.C}, L >= {self.T}], got {tuple(x.shape)}")
        return x.contiguous()

    def forward_stream(self, x: torch.Tensor, hop: int = 16, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """int8 streams [S, C, L] -> int8 logits [S, W, N] of every window x[s, :, w*hop : w*hop + T].
        Asynchronous on the current torch CUDA stream."""
        x = self._check_stream(x)
        S, L = x.shape[0], x.shape[2]
        W = self.stream_windows(L, hop)
        if out is None:
            out = torch.empty((S, W, self.N), dtype=torch.int8, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_forward_stream_int8(self._h, x.data_ptr(), S, L, int(hop), out.data_ptr(), stream),
                   "mibmi_forward_stream_int8")
        return out

    def forward_stream_taps(self, x: torch.Tensor, hop: int = 16):
        """-> (logits, {'y2_stream','y2_edges','y3','y4'}) from the fused streaming kernel (see mibmi.h)."""
        x = self._check_stream(x)
        S, L = x.shape[0], x.shape[2]
        W = self.stream_windows(L, hop)
        z = lambda *shape: torch.zeros(shape, dtype=torch.int8, device=self.device)
        out = z(S, W, self.N)
        taps = {"y2_stream": z(S, self.F2, 2 * (W - 1) + self.T8), "y2_edges": z(S, W, self.F2, 8),
                "y3": z(S, W, self.F2, self.T8), "y4": z(S, W, self.F2, self.T64)}
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_forward_stream_int8_taps(
            self._h, x.data_ptr(), S, L, int(hop), out.data_ptr(), taps["y2_stream"].data_ptr(),
            taps["y2_edges"].data_ptr(), taps["y3"].data_ptr(), taps["y4"].data_ptr(), stream),
            "mibmi_forward_stream_int8_taps")
        return out, taps

    def forward_host(self, x: Union[torch.Tensor, np.ndarray], layout: str = "bct",
                     out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """HOST int8 trials -> HOST int8 logits through mibmi_forward_int8_host (chunked, pipelined
        H2D / compute / D2H; pinned host tensors are copied from directly).  Synchronous."""
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(np.ascontiguousarray(x))
        if x.is_cuda or x.dtype != torch.int8:
            raise TypeError("forward_host takes a CPU int8 tensor")
        x = x.contiguous()
        B = x.shape[0]
        if x.numel() != B * self.trial_bytes(layout):
            raise ValueError("bad trial shape")
        if out is None:
            out = torch.empty((B, self.N), dtype=torch.int8, pin_memory=True)
        _lib.check(self._L.mibmi_forward_int8_host(self._h, x.data_ptr(), B, _LAYOUTS[layout], out.data_ptr()),
                   "mibmi_forward_int8_host")
        return out

    def quantize_input(self, x: torch.Tensor, abs_max: Optional[float] = None) -> torch.Tensor:
        """float32 trials -> int8 (gen_input_header.py:73; STEActivation floorToZero)."""
        if abs_max is None:
            if self.params is None:
                raise ValueError("abs_max required")
            abs_max = self.params.input_scale
        x = x.to(self.device, torch.float32).contiguous()
        q = torch.empty(x.shape, dtype=torch.int8, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_quantize_input_f32(x.data_ptr(), x.numel(), float(abs_max), q.data_ptr(), stream),
                   "mibmi_quantize_input_f32")
        return q

    def dequantize_logits(self, q: torch.Tensor, abs_max: Optional[float] = None) -> torch.Tensor:
        """int8 logits -> float32 (functional.py:338-353)."""
        if abs_max is None:
            if self.params is None:
                raise ValueError("abs_max required")
            abs_max = self.params.output_scale
        q = q.contiguous()
        y = torch.empty(q.shape, dtype=torch.float32, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_dequantize_logits_f32(q.data_ptr(), q.numel(), float(abs_max), y.data_ptr(), stream),
                   "mibmi_dequantize_logits_f32")
        return y


class EEGNetInt8:
    """Integer network of the EEGNet / Q-EEGNet layer order (temporal conv first) on one GPU:
    one mibmi_eegnet handle.  Integer counterpart of QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:137-168."""

    def __init__(self, params, device=None):
        from .convert import EEGNetParams
        if not isinstance(params, EEGNetParams):
            raise TypeError("EEGNetInt8 takes a convert.EEGNetParams")
        self.params = params.validate()
        self.C, self.T, self.F1, self.D, self.F2, self.N, self.Nf = (params.C, params.T, params.F1, params.D, params.F2,
                                                                   params.N, params.Nf)
        self.K2, self.T8, self.T64 = params.F1 * params.D, params.T // 8, (params.T // 8) // 8
        self.device_index = _device_index(device)
        self.device = torch.device("cuda", self.device_index)
        self._L = _lib.lib()
        keep = []

        def ptr(a, dt):
            a = np.ascontiguousarray(np.asarray(a).ravel(), dtype=dt)
            keep.append(a)
            return a.ctypes.data

        p = _lib.EEGNetParamsC(self.C, self.T, self.F1, self.D, self.F2, self.N, self.Nf,
                               ptr(params.w1, np.int8), ptr(params.fac1, np.int32), ptr(params.off1, np.int32),
                               ptr(params.w2, np.int8), ptr(params.fac2, np.int32), ptr(params.off2, np.int32),
                               ptr(params.w3, np.int8), int(params.fac3),
                               ptr(params.w4, np.int8), ptr(params.fac4, np.int32), ptr(params.off4, np.int32),
                               ptr(params.w5, np.int8), ptr(params.b5, np.int8), int(params.fac5))
        h = ctypes.c_void_p()
        _lib.check(self._L.mibmi_eegnet_create(ctypes.byref(p), self.device_index, ctypes.byref(h)), "mibmi_eegnet_create")
        del keep
        self._h = h

    @classmethod
    def from_state_dict(cls, sd: Mapping, C: int, T: int, F1: int = 8, D: int = 2, F2: Optional[int] = None, N: int = 4,
                        device=None) -> "EEGNetInt8":
        from .convert import eegnet_params_from_state_dict
        return cls(eegnet_params_from_state_dict(sd, C=C, T=T, F1=F1, D=D, F2=F2, N=N), device=device)

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._L.mibmi_eegnet_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self) -> int:
        return int(self._L.mibmi_eegnet_launch_count(self._h))

    @property
    def last_kernel(self) -> str:
        """Name of the __global__ function the most recent forward launched."""
        return (self._L.mibmi_eegnet_last_kernel(self._h) or b"").decode()

    def _check_x(self, x: torch.Tensor) -> torch.Tensor:
        if x.dtype != torch.int8 or not x.is_cuda or x.device.index != self.device_index:
            raise ValueError(f"x must be an int8 tensor on {self.device}")
        if x.dim() == 4 and x.shape[1] == 1:
            x = x[:, 0]
        if x.dim() != 3 or tuple(x.shape[1:]) != (self.C, self.T):
            raise ValueError(f"x must be [B, {self.C}, {self.T}], got {tuple(x.shape)}")
        return x.contiguous()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """int8 trials [B, C, T] -> int8 logits [B, N]; asynchronous on the current torch CUDA stream."""
        x = self._check_x(x)
        out = torch.empty((x.shape[0], self.N), dtype=torch.int8, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_eegnet_forward_int8(self._h, x.data_ptr(), x.shape[0], out.data_ptr(), stream),
                   "mibmi_eegnet_forward_int8")
        return out

    __call__ = forward

    def forward_taps(self, x: torch.Tensor):
        x = self._check_x(x)
        B = x.shape[0]
        out = torch.empty((B, self.N), dtype=torch.int8, device=self.device)
        taps = {"y2": torch.empty((B, self.K2, self.T8), dtype=torch.int8, device=self.device),
                "y3": torch.empty((B, self.K2, self.T8), dtype=torch.int8, device=self.device),
                "y4": torch.empty((B, self.F2, self.T64), dtype=torch.int8, device=self.device)}
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self._L.mibmi_eegnet_forward_int8_taps(self._h, x.data_ptr(), B, out.data_ptr(), taps["y2"].data_ptr(),
                                                          taps["y3"].data_ptr(), taps["y4"].data_ptr(), stream),
                   "mibmi_eegnet_forward_int8_taps")
        return out, taps

    quantize_input = Int8Net.quantize_input
    dequantize_logits = Int8Net.dequantize_logits
