"""Synthetic random-init edgeEEGNet state_dicts and int8 trials (there are no trained nets or
datasets in the reference tree).

The state_dict follows the schema QuantLab's exporter writes to net.npz
(QuantLab/export_net_data.py:85-87; keys probed from
QuantLab/quantlab/PhysionetMMMI/edgeEEGNet/edgeEEGnet.py:100-128):
  quant{1..6}.absMaxValue[1]; {conv1,conv2,sep_conv1,sep_conv2,fc}.{weight,weightFrozen,sParam[1]};
  fc.bias; batch_norm{1,2,3}.{weight,bias,running_mean,running_var,num_batches_tracked}.
``weightFrozen`` is snapped to the 255-level grid linspace(-s, s, 255) (inq_ops.py:152-158,212).
"""
from __future__ import annotations

import json
from typing import Dict

import numpy as np

NF = 64


def _snap(w: np.ndarray, s: np.float32) -> np.ndarray:
    q = np.round(w.astype(np.float64) / float(s) * 127.0)
    return (q / 127.0 * float(s)).astype(np.float32)


def _fakeq(x, s):
    """STEActivation with floorToZero (ste_ops.py:88-102), float64 numpy."""
    y = np.clip(x / s, -1.0, 1.0) * 127.0
    return np.trunc(y) / 127.0 * s


def _corr_same(x, w, pad):
    # x [B,K,T], w [K,taps] -> [B,K,T]
    xp = np.pad(x, ((0, 0), (0, 0), pad))
    T = x.shape[-1]
    out = np.zeros_like(x)
    for j in range(w.shape[1]):
        out += xp[..., j:j + T] * w[None, :, j:j + 1]
    return out


def make_state_dict(C=64, T=480, F1=16, D=1, N=4, seed=0, negative_gamma=True,
                    calib_trials=6, sat_quantile=0.995) -> Dict[str, np.ndarray]:
    """Random-init float net whose inter-layer activations exercise the int8 range (some saturation).

    BN gamma takes both signs when ``negative_gamma`` (=> negative requant factors).
    ``absMaxValue`` of quant2..6 is calibrated like STEActivation's monitor epoch, on random trials.
    """
    rng = np.random.default_rng(seed)
    F2 = F1 * D
    T8, T64 = T // 8, (T // 8) // 8
    sd: Dict[str, np.ndarray] = {}

    def add_w(name, shape):
        w = rng.uniform(-1.0, 1.0, size=shape).astype(np.float32)
        s = np.float32(np.abs(w).max())
        sd[f"{name}.weight"] = w
        sd[f"{name}.weightFrozen"] = _snap(w, s)
        sd[f"{name}.sParam"] = np.array([s], np.float32)

    add_w("conv1", (F2, 1, C, 1))
    add_w("conv2", (F2, 1, 1, NF))
    add_w("sep_conv1", (F2, 1, 1, 16))
    add_w("sep_conv2", (F2, F2, 1, 1))
    add_w("fc", (N, F2 * T64))
    sd["fc.bias"] = rng.uniform(-0.1, 0.1, size=(N,)).astype(np.float32)
    for i in (1, 2, 3):
        g = rng.uniform(0.5, 1.5, size=(F2,))
        if negative_gamma:
            g *= rng.choice([-1.0, 1.0], size=(F2,), p=[0.3, 0.7])
        sd[f"batch_norm{i}.weight"] = g.astype(np.float32)
        sd[f"batch_norm{i}.bias"] = rng.uniform(-0.5, 0.5, size=(F2,)).astype(np.float32)
        sd[f"batch_norm{i}.running_mean"] = rng.uniform(-0.5, 0.5, size=(F2,)).astype(np.float32)
        sd[f"batch_norm{i}.running_var"] = rng.uniform(0.5, 2.0, size=(F2,)).astype(np.float32)
        sd[f"batch_norm{i}.num_batches_tracked"] = np.array(1, np.int64)

    # ---- calibrate activation ranges with a float fake-quant forward (edgeEEGnet.py:137-168) ----
    def bn(x, i):
        g = sd[f"batch_norm{i}.weight"].astype(np.float64)
        b = sd[f"batch_norm{i}.bias"].astype(np.float64)
        m = sd[f"batch_norm{i}.running_mean"].astype(np.float64)
        v = sd[f"batch_norm{i}.running_var"].astype(np.float64)
        sc = g / np.sqrt(v + 1e-3)
        return x * sc[None, :, None] + (b - m * sc)[None, :, None]

    def absmax(x):
        return np.float32(max(np.quantile(np.abs(x), sat_quantile), 1e-3))

    s1 = np.float32(1.0)
    x = rng.integers(-127, 128, size=(calib_trials, C, T)).astype(np.float64) / 127.0 * float(s1)
    wf = lambda n: sd[f"{n}.weightFrozen"].astype(np.float64)
    h = bn(np.einsum("kc,bct->bkt", wf("conv1").reshape(F2, C), x), 1)
    s2 = absmax(h); h = _fakeq(h, float(s2))
    h = np.maximum(bn(_corr_same(h, wf("conv2").reshape(F2, NF), (31, 32)), 2), 0.0)
    h = h[..., :T8 * 8].reshape(calib_trials, F2, T8, 8).mean(-1)
    s3 = absmax(h); h = _fakeq(h, float(s3))
    h = _corr_same(h, wf("sep_conv1").reshape(F2, 16), (7, 8))
    s4 = absmax(h); h = _fakeq(h, float(s4))
    h = np.maximum(bn(np.einsum("ki,bit->bkt", wf("sep_conv2").reshape(F2, F2), h), 3), 0.0)
    h = h[..., :T64 * 8].reshape(calib_trials, F2, T64, 8).mean(-1)
    s5 = absmax(h); h = _fakeq(h, float(s5))
    h = h.reshape(calib_trials, -1) @ wf("fc").T + sd["fc.bias"].astype(np.float64)[None]
    s6 = np.float32(max(np.abs(h).max() * 0.8, 1e-3))
    for i, s in enumerate((s1, s2, s3, s4, s5, s6), start=1):
        sd[f"quant{i}.absMaxValue"] = np.array([s], np.float32)
    return sd


def make_config(C=64, T=480, F1=16, D=1, N=4) -> dict:
    """Minimal QuantLab config.json the Wolf tool-chain accepts (golden_model.py:55-83)."""
    return {"indiv": {"net": {"params": {
        "F1": F1, "D": D, "F2": F1 * D, "C": C, "T": T, "N": N,
        "weightInqNumLevels": 255, "actSTENumLevels": 255, "floorToZero": True}}}}


def save_export(prefix: str, sd: Dict[str, np.ndarray], cfg: dict):
    np.savez(prefix + "net.npz", **sd)
    with open(prefix + "config.json", "w") as f:
        json.dump(cfg, f)


def make_trials(B: int, C: int, T: int, seed: int = 0, dist: str = "uniform") -> np.ndarray:
    """int8 [B, C, T] in [-127, 127] (what quantize_to_int can emit, functional.py:308-335)."""
    rng = np.random.default_rng(seed)
    if dist == "uniform":
        return rng.integers(-127, 128, size=(B, C, T), dtype=np.int8)
    if dist == "gauss":
        return np.clip(np.round(rng.normal(0.0, 40.0, size=(B, C, T))), -127, 127).astype(np.int8)
    raise ValueError(dist)


def make_eegnet_state_dict(C=22, T=1125, F1=8, D=2, N=4, Nf=128, seed=0, negative_gamma=True, calib_trials=3,
                           sat_quantile=0.995) -> Dict[str, np.ndarray]:
    """Random-init float net of the EEGNet / Q-EEGNet order (QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:100-128):
    conv1 (F1,1,1,Nf), batch_norm1 (F1), conv2 (F1*D,1,C,1), batch_norm2, sep_conv1 (F1*D,1,1,16),
    sep_conv2 (F2,F1*D,1,1), batch_norm3 (F2), fc (N, F2*T64); quant1..6 calibrated on random trials."""
    rng = np.random.default_rng(seed)
    K2 = F2 = F1 * D
    T8, T64 = T // 8, (T // 8) // 8
    sd: Dict[str, np.ndarray] = {}

    def add_w(name, shape):
        w = rng.uniform(-1.0, 1.0, size=shape).astype(np.float32)
        s = np.float32(np.abs(w).max())
        sd[f"{name}.weight"] = w
        sd[f"{name}.weightFrozen"] = _snap(w, s)
        sd[f"{name}.sParam"] = np.array([s], np.float32)

    add_w("conv1", (F1, 1, 1, Nf))
    add_w("conv2", (K2, 1, C, 1))
    add_w("sep_conv1", (K2, 1, 1, 16))
    add_w("sep_conv2", (F2, K2, 1, 1))
    add_w("fc", (N, F2 * T64))
    sd["fc.bias"] = rng.uniform(-0.1, 0.1, size=(N,)).astype(np.float32)
    for i, n in ((1, F1), (2, K2), (3, F2)):
        g = rng.uniform(0.5, 1.5, size=(n,))
        if negative_gamma:
            g *= rng.choice([-1.0, 1.0], size=(n,), p=[0.3, 0.7])
        sd[f"batch_norm{i}.weight"] = g.astype(np.float32)
        sd[f"batch_norm{i}.bias"] = rng.uniform(-0.5, 0.5, size=(n,)).astype(np.float32)
        sd[f"batch_norm{i}.running_mean"] = rng.uniform(-0.5, 0.5, size=(n,)).astype(np.float32)
        sd[f"batch_norm{i}.running_var"] = rng.uniform(0.5, 2.0, size=(n,)).astype(np.float32)
        sd[f"batch_norm{i}.num_batches_tracked"] = np.array(1, np.int64)

    def bn(x, i, axis):
        g, b, m, v = (sd[f"batch_norm{i}.{k}"].astype(np.float64) for k in ("weight", "bias", "running_mean", "running_var"))
        sc = g / np.sqrt(v + 1e-3)
        sh = [1] * x.ndim
        sh[axis] = -1
        return x * sc.reshape(sh) + (b - m * sc).reshape(sh)

    def absmax(x):
        return np.float32(max(np.quantile(np.abs(x), sat_quantile), 1e-3))

    wf = lambda n: sd[f"{n}.weightFrozen"].astype(np.float64)
    s1 = np.float32(1.0)
    x = rng.integers(-127, 128, size=(calib_trials, C, T)).astype(np.float64) / 127.0
    xp = np.pad(x, ((0, 0), (0, 0), (Nf // 2 - 1, Nf // 2)))
    win = np.lib.stride_tricks.sliding_window_view(xp, Nf, axis=-1)            # [B, C, T, Nf]
    h = bn(win @ wf("conv1").reshape(F1, Nf).T, 1, 3)                          # [B, C, T, F1]
    s2 = absmax(h); h = _fakeq(h, float(s2))
    w2 = wf("conv2").reshape(K2, C)
    h = np.stack([np.einsum("bct,c->bt", h[..., k // D], w2[k]) for k in range(K2)], 1)   # [B, K2, T]
    h = np.maximum(bn(h, 2, 1), 0.0)[..., :T8 * 8].reshape(calib_trials, K2, T8, 8).mean(-1)
    s3 = absmax(h); h = _fakeq(h, float(s3))
    h = _corr_same(h, wf("sep_conv1").reshape(K2, 16), (7, 8))
    s4 = absmax(h); h = _fakeq(h, float(s4))
    h = np.maximum(bn(np.einsum("ki,bit->bkt", wf("sep_conv2").reshape(F2, K2), h), 3, 1), 0.0)
    h = h[..., :T64 * 8].reshape(calib_trials, F2, T64, 8).mean(-1)
    s5 = absmax(h); h = _fakeq(h, float(s5))
    h = h.reshape(calib_trials, -1) @ wf("fc").T + sd["fc.bias"].astype(np.float64)[None]
    s6 = np.float32(max(np.abs(h).max() * 0.8, 1e-3))
    for i, s in enumerate((s1, s2, s3, s4, s5, s6), start=1):
        sd[f"quant{i}.absMaxValue"] = np.array([s], np.float32)
    return sd
