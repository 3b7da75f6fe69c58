"""Float state_dict -> integer (Wolf) parameters for the int8 edgeEEGNet / MI-BMInet forward.

Host-side, run once per network.  Mirrors the reference's export maths:

* weight quantisation      edge-eegnet_wolf/python_utils/convert_torch_format.py:36-85
                           edge-eegnet_wolf/python_utils/functional.py:308-335
* BN folding / requant     convert_torch_format.py:88-113, :133-205   (R = 127, eps = 1e-3)
* input quantisation       edge-eegnet_wolf/data/gen_input_header.py:73-76
* exported array layout    edge-eegnet_wolf/data/gen_net_header.py:49-228 (net.h / net.c)

The reference runs these formulas on ``np.float32`` scalars/arrays loaded from ``net.npz``; the
rounding of ``factor``/``offset`` therefore depends on float32 intermediate precision.  Every
intermediate below is pinned to the dtype the reference code produces under NumPy >= 2 (NEP 50
promotion: python scalars are weak, np.float32 stays float32), which is the environment the parity
fixtures in tests/golden/ were generated in (see oracle/gen_golden.py).
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Dict, Mapping, Optional

import numpy as np

NUM_LEVELS = 255
R = (NUM_LEVELS - 1) / 2  # 127.0
NF = 64  # temporal filter length hard-wired in the Wolf tool-chain (golden_model.py:216-219)
L3_TAPS = 16
BN_EPS = 0.001


def align4(d: int) -> int:
    """header_file.py:255-265 (align_array_size, n=4)."""
    return (int(d) + 3) // 4 * 4


def _f32(x):
    return np.asarray(x, dtype=np.float32)


def quantize_to_int(x, scale_factor, num_levels: int = NUM_LEVELS) -> np.ndarray:
    """trunc(clip(x / s, -1, 1) * (L-1) / 2)   -- functional.py:308-335 (float32 in, float32 maths)."""
    assert num_levels % 2
    x = _f32(x) / np.float32(scale_factor)
    x = np.clip(x, -1, 1)
    x = x * np.float32(num_levels - 1)
    x = x / np.float32(2)
    return x.astype(np.int64)  # astype(int) truncates toward zero


def dequantize(x, scale_factor, num_levels: int = NUM_LEVELS) -> np.ndarray:
    """functional.py:338-353."""
    return np.asarray(x) / ((num_levels - 1) / 2) * scale_factor


def _arr(net: Mapping, key: str) -> np.ndarray:
    v = net[key]
    if hasattr(v, "detach"):  # torch tensor
        v = v.detach().cpu().numpy()
    return np.asarray(v)


def _inq_conv2d(net, name, store_reversed=False):
    """convert_torch_format.py:36-60.  store_reversed=True keeps torch (cross-correlation) order."""
    w = _arr(net, f"{name}.weightFrozen")
    s = _f32(_arr(net, f"{name}.sParam")).ravel()[0]
    if not store_reversed:
        w = np.flip(w, (-2, -1))
    return quantize_to_int(w, s), s


def _inq_linear(net, name):
    """convert_torch_format.py:63-85.  NB: the bias is quantised with the *weight* scale (:83)."""
    s = _f32(_arr(net, f"{name}.sParam")).ravel()[0]
    w = quantize_to_int(_arr(net, f"{name}.weightFrozen"), s)
    b = quantize_to_int(_arr(net, f"{name}.bias"), s)
    return w, b, s


def _batch_norm(net, name, eps=BN_EPS):
    """convert_torch_format.py:88-113."""
    mean = _f32(_arr(net, f"{name}.running_mean"))
    var = _f32(_arr(net, f"{name}.running_var"))
    gamma = _f32(_arr(net, f"{name}.weight"))
    beta = _f32(_arr(net, f"{name}.bias"))
    root = np.sqrt(var + np.float32(eps))
    scale = gamma / root
    offset = beta - mean * gamma / root
    return scale, offset


def _ste_quant(net, name):
    """convert_torch_format.py:116-130."""
    try:
        return _f32(_arr(net, f"{name}.absMaxValue")).ravel()[0]
    except KeyError:
        return _f32(_arr(net, f"{name}.quant.absMaxValue")).ravel()[0]


def _div_factor(s_in, s_w, s_out, pool=1) -> int:
    """convert_torch_format.py:133-163 (float32 scalars, python-float R is weak)."""
    f = np.float32(s_out) * np.float32(R) / (np.float32(s_in) * np.float32(s_w))
    f = f * np.float32(pool)
    return int(round(float(f)))


def _div_factor_bn(s_in, s_w, s_out, bn_scale, bn_offset, pool=1):
    """convert_torch_format.py:166-205."""
    r = np.float32(R)
    den = bn_scale * np.float32(s_in) * np.float32(s_w)
    factor = np.float32(s_out) * r / den
    bias = bn_offset * r * r / den
    factor = factor * np.float32(pool)
    bias = bias * np.float32(pool)
    return factor.round().astype(np.int64), bias.round().astype(np.int64)


@dataclass
class WolfParams:
    """Integer network in *logical* form (all weights in torch / cross-correlation order).

    The exported Wolf arrays (net.h) are produced by :meth:`to_wolf_arrays`.
    """

    C: int
    T: int
    F1: int
    D: int
    F2: int
    N: int
    w1: np.ndarray  # [F2, C]     int8, dot-product order        (net_l1_weight)
    fac1: np.ndarray  # [F2] int32
    off1: np.ndarray  # [F2] int32
    w2: np.ndarray  # [F2, 64]    torch order                    (net_l2_weight_reverse)
    fac2: np.ndarray
    off2: np.ndarray
    w3: np.ndarray  # [F2, 16]    torch order (net_l3_weight is this, flipped)
    fac3: int
    w4: np.ndarray  # [F2out, F2in]
    fac4: np.ndarray
    off4: np.ndarray
    w5: np.ndarray  # [N, F2 * T64]  flatten order (k, q)
    b5: np.ndarray  # [N] int8
    fac5: int
    input_scale: float = 1.0  # quant1.absMaxValue
    output_scale: float = 1.0  # quant6.absMaxValue
    scales: Dict[str, float] = field(default_factory=dict)

    # ---- derived sizes -------------------------------------------------------------------
    @property
    def T8(self):
        return self.T // 8

    @property
    def T64(self):
        return (self.T // 8) // 8

    @property
    def C_ALIGN(self):
        return align4(self.C)

    @property
    def T64_ALIGN(self):
        return align4(self.T64)

    def validate(self):
        """Constraints of the Wolf tool-chain (golden_model.py:61-83, gen_net_header.py:61-73)."""
        if self.F2 != self.F1 * self.D:
            raise ValueError("F2 must equal F1*D")
        if self.D != 1:
            raise ValueError("the integer pipeline only exists for D == 1 (golden_model.py:169)")
        if self.F2 % 4 != 0:
            raise ValueError("F2 must be a multiple of 4 (gen_net_header.py:63)")
        if self.w2.shape != (self.F2, NF):
            raise ValueError("temporal filter length must be 64")
        for name in ("fac1", "fac2", "fac4"):
            if np.any(getattr(self, name) == 0):
                raise ValueError(f"{name} contains a zero divisor")
        if self.fac3 == 0 or self.fac5 == 0:
            raise ValueError("zero scalar divisor")
        return self

    # ---- exported (net.h) layout ---------------------------------------------------------
    def to_wolf_arrays(self) -> Dict[str, np.ndarray]:
        """Arrays exactly as gen_net_header.py writes them into net.c (names without 'net_')."""
        F2, C, CA = self.F2, self.C, self.C_ALIGN
        T64, T64A = self.T64, self.T64_ALIGN
        w1a = np.zeros((F2, CA), np.int8)
        w1a[:, :C] = self.w1
        w5a = np.zeros((self.N, F2 * T64A), np.int8)
        for i in range(F2):
            w5a[:, i * T64A:i * T64A + T64] = self.w5[:, i * T64:(i + 1) * T64]
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        i8 = lambda a: np.ascontiguousarray(a, dtype=np.int8)
        return {
            "l1_factor": i32(self.fac1), "l1_offset": i32(self.off1), "l1_weight_align": i8(w1a).ravel(),
            "l2_factor": i32(self.fac2), "l2_offset": i32(self.off2),
            "l2_weight_reverse": i8(self.w2).ravel(), "l2_weight": i8(self.w2[:, ::-1]).ravel(),
            "l3_factor": np.int32(self.fac3), "l3_weight": i8(self.w3[:, ::-1]).ravel(),
            "l4_factor": i32(self.fac4), "l4_offset": i32(self.off4), "l4_weight": i8(self.w4).ravel(),
            "l5_factor": np.int32(self.fac5), "l5_bias": i8(self.b5), "l5_weight": i8(w5a).ravel(),
        }

    @classmethod
    def from_wolf_arrays(cls, a: Mapping[str, np.ndarray], C, T, F1=16, D=1, N=4, **kw) -> "WolfParams":
        F2 = F1 * D
        CA, T64 = align4(C), (T // 8) // 8
        T64A = align4(T64)
        w5a = np.asarray(a["l5_weight"]).reshape(N, F2, T64A)
        return cls(
            C=C, T=T, F1=F1, D=D, F2=F2, N=N,
            w1=np.asarray(a["l1_weight_align"]).reshape(F2, CA)[:, :C].astype(np.int8),
            fac1=np.asarray(a["l1_factor"], np.int32), off1=np.asarray(a["l1_offset"], np.int32),
            w2=np.asarray(a["l2_weight_reverse"]).reshape(F2, NF).astype(np.int8),
            fac2=np.asarray(a["l2_factor"], np.int32), off2=np.asarray(a["l2_offset"], np.int32),
            w3=np.asarray(a["l3_weight"]).reshape(F2, L3_TAPS)[:, ::-1].astype(np.int8),
            fac3=int(a["l3_factor"]),
            w4=np.asarray(a["l4_weight"]).reshape(F2, F2).astype(np.int8),
            fac4=np.asarray(a["l4_factor"], np.int32), off4=np.asarray(a["l4_offset"], np.int32),
            w5=np.ascontiguousarray(w5a[:, :, :T64]).reshape(N, F2 * T64).astype(np.int8),
            b5=np.asarray(a["l5_bias"], np.int8), fac5=int(a["l5_factor"]), **kw).validate()


def params_from_state_dict(net: Mapping, C: int, T: int, F1: int = 16, D: int = 1,
                           F2: Optional[int] = None, N: int = 4) -> WolfParams:
    """QuantLab edgeEEGNet state_dict (or net.npz mapping) -> WolfParams.

    Reads only the keys the Wolf tool-chain reads: ``<layer>.weightFrozen``, ``<layer>.sParam``,
    ``fc.bias``, BN statistics and ``quantN.absMaxValue``.
    """
    if F2 is None:
        F2 = F1 * D
    T64 = (T // 8) // 8
    s = {f"quant{i}": _ste_quant(net, f"quant{i}") for i in range(1, 7)}

    w1, sw1 = _inq_conv2d(net, "conv1", store_reversed=True)
    assert w1.shape == (F2, 1, C, 1), w1.shape
    bn_s, bn_o = _batch_norm(net, "batch_norm1")
    fac1, off1 = _div_factor_bn(s["quant1"], sw1, s["quant2"], bn_s, bn_o)

    w2, sw2 = _inq_conv2d(net, "conv2", store_reversed=True)
    assert w2.shape == (F2, 1, 1, NF), w2.shape
    bn_s, bn_o = _batch_norm(net, "batch_norm2")
    fac2, off2 = _div_factor_bn(s["quant2"], sw2, s["quant3"], bn_s, bn_o, pool=8)

    w3, sw3 = _inq_conv2d(net, "sep_conv1", store_reversed=True)
    assert w3.shape == (F2, 1, 1, L3_TAPS), w3.shape
    fac3 = _div_factor(s["quant3"], sw3, s["quant4"])

    w4, sw4 = _inq_conv2d(net, "sep_conv2", store_reversed=True)
    assert w4.shape == (F2, F2, 1, 1), w4.shape
    bn_s, bn_o = _batch_norm(net, "batch_norm3")
    fac4, off4 = _div_factor_bn(s["quant4"], sw4, s["quant5"], bn_s, bn_o, pool=8)

    w5, b5, sw5 = _inq_linear(net, "fc")
    assert w5.shape == (N, F2 * T64), w5.shape
    fac5 = _div_factor(s["quant5"], sw5, s["quant6"])

    i32 = lambda a: np.asarray(a, dtype=np.int64).astype(np.int32)
    for name, arr in (("l1", fac1), ("l1o", off1), ("l2", fac2), ("l2o", off2), ("l4", fac4), ("l4o", off4)):
        if np.any(np.abs(arr) > 2**31 - 1):
            raise OverflowError(f"{name}: requant parameter does not fit int32")
    return WolfParams(
        C=C, T=T, F1=F1, D=D, F2=F2, N=N,
        w1=w1.reshape(F2, C).astype(np.int8), fac1=i32(fac1), off1=i32(off1),
        w2=w2.reshape(F2, NF).astype(np.int8), fac2=i32(fac2), off2=i32(off2),
        w3=w3.reshape(F2, L3_TAPS).astype(np.int8), fac3=fac3,
        w4=w4.reshape(F2, F2).astype(np.int8), fac4=i32(fac4), off4=i32(off4),
        w5=w5.astype(np.int8), b5=b5.astype(np.int8), fac5=fac5,
        input_scale=float(s["quant1"]), output_scale=float(s["quant6"]),
        scales={k: float(v) for k, v in s.items()}).validate()


def params_from_export(net_npz: str, config_json: str) -> WolfParams:
    """Load QuantLab's export (QuantLab/export_net_data.py:85-87,170): net.npz + config.json."""
    with open(config_json, "r") as f:
        p = json.load(f)["indiv"]["net"]["params"]
    if p.get("weightInqNumLevels") != 255 or p.get("actSTENumLevels") != 255:
        raise ValueError("only 255-level nets are supported (golden_model.py:61-62)")
    if not p.get("floorToZero"):
        raise ValueError("only floorToZero nets are supported (golden_model.py:65)")
    net = np.load(net_npz)
    return params_from_state_dict(net, C=p["C"], T=p["T"], F1=p["F1"], D=p["D"], F2=p.get("F2"), N=p["N"])


def quantize_input(x: np.ndarray, input_scale: float) -> np.ndarray:
    """gen_input_header.py:73 -- float [.., C, T] -> int8 in [-127, 127] (truncation toward zero)."""
    return quantize_to_int(x, input_scale).astype(np.int8)


def to_wolf_input(x_int8: np.ndarray) -> np.ndarray:
    """[..., C, T] -> Wolf order [..., T, C_ALIGN] (gen_input_header.py:75-76)."""
    xt = np.swapaxes(x_int8, -1, -2)
    C = xt.shape[-1]
    out = np.zeros((*xt.shape[:-1], align4(C)), np.int8)
    out[..., :C] = xt
    return out


# =====================================================================================================
# EEGNet / Q-EEGNet layer order: temporal conv (1, Nf) first, then the depthwise spatial
# conv -- QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:100-106 (modules), :137-168 (forward).
# The reference tree holds NO integer model of this topology (the Q-EEGNet Wolf C code is not vendored);
# the integer pipeline below applies the same export maths (a11 / a12: convert_torch_format.py:36-205) to the
# module's layers, REORDER_BN form for the two pooled layers.  Parity: unpinned by the reference.
# =====================================================================================================
@dataclass
class EEGNetParams:
    """Integer EEGNet-order network, all weights in torch (cross-correlation) order.

    L1  v[f][c][t] = sum_j xpad[c][t+j] w1[f][j]        (pad Nf/2-1 | Nf/2, eegnet.py:101-102)
        y1 = clip8(tdiv(v + off1[f], fac1[f]))           (batch_norm1 folded, quant2)
    L2  a[k][t]   = sum_c y1[k // D][c][t] w2[k][c]      (depthwise spatial, groups = F1, eegnet.py:105)
        s = sum_8 max(a, -(off2[k] >> 3)); y2 = clip8(tdiv(s + off2[k], fac2[k]))   (batch_norm2, ReLU, pool 8, quant3)
    L3..L5 as in the edgeEEGNet pipeline (sep_conv1 / sep_conv2 + batch_norm3 + pool / fc).
    """
    C: int
    T: int
    F1: int
    D: int
    F2: int
    N: int
    Nf: int
    w1: np.ndarray    # [F1, Nf]
    fac1: np.ndarray  # [F1]
    off1: np.ndarray
    w2: np.ndarray    # [F1*D, C]
    fac2: np.ndarray  # [F1*D]
    off2: np.ndarray
    w3: np.ndarray    # [F1*D, 16]
    fac3: int
    w4: np.ndarray    # [F2, F1*D]
    fac4: np.ndarray  # [F2]
    off4: np.ndarray
    w5: np.ndarray    # [N, F2 * T64]
    b5: np.ndarray
    fac5: int
    input_scale: float = 1.0
    output_scale: float = 1.0
    scales: Dict[str, float] = field(default_factory=dict)

    @property
    def K2(self):
        return self.F1 * self.D

    def validate(self):
        if self.Nf % 4 or not 4 <= self.Nf <= 128:
            raise ValueError("temporal filter length must be a multiple of 4 in [4, 128]")
        if self.w1.shape != (self.F1, self.Nf) or self.w2.shape != (self.K2, self.C):
            raise ValueError("weight shapes do not match (F1, Nf) / (F1*D, C)")
        if self.w4.shape != (self.F2, self.K2) or self.w5.shape != (self.N, self.F2 * ((self.T // 8) // 8)):
            raise ValueError("sep_conv2 / fc shapes do not match")
        for name in ("fac1", "fac2", "fac4"):
            if np.any(getattr(self, name) == 0):
                raise ValueError(f"{name} contains a zero divisor")
        if self.fac3 == 0 or self.fac5 == 0:
            raise ValueError("zero scalar divisor")
        return self


def eegnet_params_from_state_dict(net: Mapping, C: int, T: int, F1: int = 8, D: int = 2, F2: Optional[int] = None,
                                  N: int = 4) -> EEGNetParams:
    """QuantLab EEGNet state_dict (eegnet.py:100-128 key names) -> EEGNetParams, convert_torch_format.py maths."""
    K2 = F1 * D
    if F2 is None:
        F2 = K2
    T64 = (T // 8) // 8
    s = {f"quant{i}": _ste_quant(net, f"quant{i}") for i in range(1, 7)}
    w1, sw1 = _inq_conv2d(net, "conv1", store_reversed=True)
    assert w1.ndim == 4 and w1.shape[:3] == (F1, 1, 1), w1.shape
    Nf = w1.shape[3]
    bn_s, bn_o = _batch_norm(net, "batch_norm1")
    fac1, off1 = _div_factor_bn(s["quant1"], sw1, s["quant2"], bn_s, bn_o)
    w2, sw2 = _inq_conv2d(net, "conv2", store_reversed=True)
    assert w2.shape == (K2, 1, C, 1), w2.shape
    bn_s, bn_o = _batch_norm(net, "batch_norm2")
    fac2, off2 = _div_factor_bn(s["quant2"], sw2, s["quant3"], bn_s, bn_o, pool=8)
    w3, sw3 = _inq_conv2d(net, "sep_conv1", store_reversed=True)
    assert w3.shape == (K2, 1, 1, L3_TAPS), w3.shape
    fac3 = _div_factor(s["quant3"], sw3, s["quant4"])
    w4, sw4 = _inq_conv2d(net, "sep_conv2", store_reversed=True)
    assert w4.shape == (F2, K2, 1, 1), w4.shape
    bn_s, bn_o = _batch_norm(net, "batch_norm3")
    fac4, off4 = _div_factor_bn(s["quant4"], sw4, s["quant5"], bn_s, bn_o, pool=8)
    w5, b5, sw5 = _inq_linear(net, "fc")
    assert w5.shape == (N, F2 * T64), w5.shape
    fac5 = _div_factor(s["quant5"], sw5, s["quant6"])
    i32 = lambda a: np.asarray(a, dtype=np.int64).astype(np.int32)
    for arr in (fac1, off1, fac2, off2, fac4, off4):
        if np.any(np.abs(arr) > 2**31 - 1):
            raise OverflowError("requant parameter does not fit int32")
    return EEGNetParams(
        C=C, T=T, F1=F1, D=D, F2=F2, N=N, Nf=Nf,
        w1=w1.reshape(F1, Nf).astype(np.int8), fac1=i32(fac1), off1=i32(off1),
        w2=w2.reshape(K2, C).astype(np.int8), fac2=i32(fac2), off2=i32(off2),
        w3=w3.reshape(K2, L3_TAPS).astype(np.int8), fac3=fac3,
        w4=w4.reshape(F2, K2).astype(np.int8), fac4=i32(fac4), off4=i32(off4),
        w5=w5.astype(np.int8), b5=b5.astype(np.int8), fac5=fac5,
        input_scale=float(s["quant1"]), output_scale=float(s["quant6"]),
        scales={k: float(v) for k, v in s.items()}).validate()
