"""`verification.npz` of a QuantLab export: per-layer check of the GPU integer path.

`QuantLab/export_net_data.py:112-167` writes the float-simulated tensor after every module of one sample into
`verification.npz`.  The reference checks its integer golden model against the five activation taps of that file
(`edge-eegnet_wolf/test/python/golden_model/testcase.py:67-93`; error measure `:128-153`): quantise the float tap with the
layer's output scale, compare with the integer layer output, accept a maximal difference of TOLERANCE steps.

:func:`verify_export` runs the same check with the GPU kernels in the place of the golden model (chained, as
`test_model` does: the input is quantised once, every later layer sees the integer output of the layer before).
:func:`integer_verification` goes the other way: the integer taps of the GPU path under the reference's tap names and
shapes, scaled back to floats, so that a file written by :func:`write_verification` can stand in for `verification.npz`
in the reference's layer tests (`test/cl/net/layer2/testcase.py:56`, `layer3/testcase.py:56`).
"""
from __future__ import annotations

from typing import Dict, Mapping, Optional, Union

import numpy as np
import torch

from .convert import WolfParams, params_from_export, quantize_to_int
from .engine import Int8Net

TOLERANCE = 10  # quantisation steps; golden_model/testcase.py TOLERANCE and SURVEY 8c O3

# tap of verification.npz  ->  (key of Int8Net.forward_taps, scale = quantN.absMaxValue the tap was quantised with)
ACTIV_TAPS = (("layer1_activ", "y1", "quant2"), ("layer2_activ", "y2", "quant3"), ("layer3_activ", "y3", "quant4"),
              ("layer4_activ", "y4", "quant5"), ("output_quant", "out", "quant6"))


def _compare(y_hat: np.ndarray, y_exp: np.ndarray, tolerance: float, epsilon: float = 1e-4) -> Dict[str, object]:
    diff = np.abs(y_hat.astype(np.int64) - y_exp.astype(np.int64))
    return {"result": bool(diff.max() <= tolerance + epsilon), "max error": int(diff.max()),
            "avg error": float(diff.mean())}


def gpu_taps(net: Int8Net, x_int8: np.ndarray) -> Dict[str, np.ndarray]:
    """One trial ``int8[C, T]`` -> the five integer activations, trimmed to their logical lengths."""
    p = net.params
    x = torch.from_numpy(np.ascontiguousarray(x_int8, dtype=np.int8).reshape(1, p.C, p.T)).to(net.device)
    out, taps = net.forward_taps(x)
    torch.cuda.synchronize(net.device)
    T8, T64 = p.T // 8, (p.T // 8) // 8
    res = {"y1": taps["y1"][0, :, :p.T], "y2": taps["y2"][0, :, :T8], "y3": taps["y3"][0, :, :T8],
           "y4": taps["y4"][0, :, :T64], "out": out[0]}
    return {k: v.cpu().numpy().astype(np.int8) for k, v in res.items()}


def expected_taps(verification: Mapping[str, np.ndarray], params: WolfParams) -> Dict[str, np.ndarray]:
    """The float taps of verification.npz in integer form (what `test_model` compares with)."""
    out = {}
    for name, key, q in ACTIV_TAPS:
        a = np.asarray(verification[name], dtype=np.float32)
        a = a.reshape(params.N) if key == "out" else a.reshape(params.F2, -1)
        out[key] = quantize_to_int(a, params.scales[q]).astype(np.int64)
    return out


def verify_export(net_npz: str, config_json: str, verification_npz: str, tolerance: float = TOLERANCE,
                  device: Optional[Union[int, str, torch.device]] = None,
                  net: Optional[Int8Net] = None) -> Dict[str, Dict[str, object]]:
    """Per-layer report ``{"1".."5": {"result", "max error", "avg error"}}`` (keys of the reference's report)."""
    params = params_from_export(net_npz, config_json)
    own = net is None
    if own:
        net = Int8Net(params, device=device)
    try:
        data = np.load(verification_npz)
        x = quantize_to_int(np.asarray(data["input"], np.float32).reshape(params.C, params.T), params.input_scale)
        got, exp = gpu_taps(net, x.astype(np.int8)), expected_taps(data, params)
        return {str(i + 1): _compare(got[key], exp[key], tolerance) for i, (_n, key, _q) in enumerate(ACTIV_TAPS)}
    finally:
        if own:
            net.close()


def _dequantize_exact(a: np.ndarray, scale: float) -> np.ndarray:
    """int8 -> float32 on the grid ``a * s / 127``, nudged by single ulps away from zero wherever float32 rounding would
    make `quantize_to_int` (truncation) land one step short, so that quantising the written tap returns ``a`` exactly."""
    a = np.asarray(a).astype(np.int64)
    f = (a.astype(np.float64) * float(scale) / 127.0).astype(np.float32)
    for _ in range(4):
        bad = quantize_to_int(f, scale) != a
        if not bad.any():
            return f
        f = np.where(bad, np.nextafter(f, np.where(a < 0, -np.inf, np.inf).astype(np.float32)), f).astype(np.float32)
    raise ArithmeticError("scale too close to the float32 grid to round-trip")


def integer_verification(net: Int8Net, x_float: np.ndarray) -> Dict[str, np.ndarray]:
    """Float trial ``[C, T]`` -> {tap name: float32 array in the shape export_net_data.py stores}, computed by the GPU
    integer path: ``input``, ``input_quant`` and the five activation taps (the names the reference's tests load)."""
    p = net.params
    x_float = np.asarray(x_float, dtype=np.float32).reshape(p.C, p.T)
    x_int = quantize_to_int(x_float, p.input_scale).astype(np.int8)
    got = gpu_taps(net, x_int)
    deq = lambda a, q: _dequantize_exact(a, p.scales[q])
    # 'input' keeps the leading dimension of the dataset sample export_net_data.py:109,116 stores (the reference's tests
    # index ["input"][0, :, :], test/cl/net/model/testcase.py:55)
    out = {"input": x_float.reshape(1, p.C, p.T), "input_quant": deq(x_int, "quant1").reshape(1, 1, p.C, p.T)}
    for name, key, q in ACTIV_TAPS:
        a = deq(got[key], q)
        out[name] = a.reshape(1, p.N) if key == "out" else a.reshape(1, p.F2, 1, -1)
    return out


def write_verification(path: str, net: Int8Net, x_float: np.ndarray) -> None:
    np.savez(path, **integer_verification(net, x_float))
