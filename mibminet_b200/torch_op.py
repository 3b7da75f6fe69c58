"""`torch.ops.mibminet.forward_int8` -- the torch-operator face of the C-ABI (SURVEY 8(b), row "torch op").

    torch.ops.mibminet.forward_int8(Tensor x_int8, Tensor blob, int[] dims, bool reorder_bn) -> Tensor int8[B, N]

* ``x_int8``   int8 CUDA tensor ``[B, C, T]`` (or ``[B, 1, C, T]``), module order
* ``blob``     uint8 CPU tensor: the exported arrays of gen_net_header.py (edge-eegnet_wolf/data/gen_net_header.py:109-224) in
               the fixed order of :data:`BLOB_FIELDS`, written by :func:`pack_blob`
* ``dims``     ``[C, T, F1, D, N]``
* CUDA-only dispatch key: calling it with a CPU ``x_int8`` raises (there is no CPU path).

The operator is plumbing around ``mibmi_create`` / ``mibmi_forward_int8`` (include/mibmi.h): the handle built from a blob is
cached per (blob bytes, dims, reorder_bn, device).  Importing this module registers the operator.
"""
from __future__ import annotations

import hashlib
from typing import Dict, List, Mapping, Tuple

import numpy as np
import torch

from .convert import align4
from .engine import Int8Net

# (name, dtype) in blob order; sizes follow from dims (F2 = F1 * D, C_ALIGN / T64_ALIGN = round up to 4)
BLOB_FIELDS = (("l1_factor", np.int32), ("l1_offset", np.int32), ("l2_factor", np.int32), ("l2_offset", np.int32),
               ("l3_factor", np.int32), ("l4_factor", np.int32), ("l4_offset", np.int32), ("l5_factor", np.int32),
               ("l1_weight_align", np.int8), ("l2_weight_reverse", np.int8), ("l3_weight", np.int8), ("l4_weight", np.int8),
               ("l5_bias", np.int8), ("l5_weight", np.int8))


def _sizes(dims) -> Dict[str, int]:
    C, T, F1, D, N = (int(v) for v in dims)
    F2, T64A = F1 * D, align4((T // 8) // 8)
    return {"l1_factor": F2, "l1_offset": F2, "l2_factor": F2, "l2_offset": F2, "l3_factor": 1, "l4_factor": F2,
            "l4_offset": F2, "l5_factor": 1, "l1_weight_align": F2 * align4(C), "l2_weight_reverse": F2 * 64,
            "l3_weight": F2 * 16, "l4_weight": F2 * F2, "l5_bias": N, "l5_weight": N * F2 * T64A}


def pack_blob(arrays: Mapping[str, np.ndarray], dims) -> torch.Tensor:
    """Exported arrays (names of gen_net_header.py without ``net_``) -> uint8 CPU tensor."""
    sz = _sizes(dims)
    parts = []
    for name, dt in BLOB_FIELDS:
        a = np.ascontiguousarray(np.asarray(arrays[name]).ravel(), dtype=dt)
        if a.size != sz[name]:
            raise ValueError(f"{name}: expected {sz[name]} elements for dims {list(dims)}, got {a.size}")
        parts.append(a.view(np.uint8))
    return torch.from_numpy(np.concatenate(parts).copy())


def unpack_blob(blob: torch.Tensor, dims) -> Dict[str, np.ndarray]:
    raw = blob.detach().cpu().contiguous().numpy().view(np.uint8).ravel()
    sz, out, pos = _sizes(dims), {}, 0
    for name, dt in BLOB_FIELDS:
        n = sz[name] * np.dtype(dt).itemsize
        if pos + n > raw.size:
            raise ValueError("blob is shorter than dims require")
        out[name] = raw[pos:pos + n].view(dt).copy()
        pos += n
    if pos != raw.size:
        raise ValueError(f"blob has {raw.size} bytes, dims require {pos}")
    out["l3_factor"], out["l5_factor"] = np.int32(out["l3_factor"][0]), np.int32(out["l5_factor"][0])
    return out


_cache: Dict[Tuple, Int8Net] = {}


def _net_for(blob: torch.Tensor, dims: List[int], reorder_bn: bool, device: torch.device) -> Int8Net:
    if len(dims) != 5:
        raise ValueError("dims must be [C, T, F1, D, N]")
    raw = blob.detach().cpu().contiguous().numpy().tobytes()
    key = (hashlib.sha1(raw).hexdigest(), tuple(int(v) for v in dims), bool(reorder_bn), device.index)
    net = _cache.get(key)
    if net is None:
        C, T, F1, D, N = (int(v) for v in dims)
        net = Int8Net(unpack_blob(blob, dims), C=C, T=T, F1=F1, D=D, N=N, reorder_bn=bool(reorder_bn), device=device)
        if len(_cache) >= 64:
            _cache.pop(next(iter(_cache)))
        _cache[key] = net
    return net


def _forward_int8_cuda(x: torch.Tensor, blob: torch.Tensor, dims: List[int], reorder_bn: bool) -> torch.Tensor:
    return _net_for(blob, dims, reorder_bn, x.device).forward(x)


def _forward_int8_meta(x: torch.Tensor, blob: torch.Tensor, dims: List[int], reorder_bn: bool) -> torch.Tensor:
    return x.new_empty((x.shape[0], int(dims[4])), dtype=torch.int8)


_lib_def = torch.library.Library("mibminet", "DEF")
_lib_def.define("forward_int8(Tensor x_int8, Tensor blob, int[] dims, bool reorder_bn) -> Tensor")
_lib_def.impl("forward_int8", _forward_int8_cuda, "CUDA")
_lib_def.impl("forward_int8", _forward_int8_meta, "Meta")

forward_int8 = torch.ops.mibminet.forward_int8
