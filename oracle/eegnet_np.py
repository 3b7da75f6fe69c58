"""ORACLE (test infrastructure, not product code): NumPy restatement of the integer EEGNet / Q-EEGNet layer order.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this file.

PARITY UNPINNED.  The reference tree holds no integer model of this topology: QuantLab's EEGNet module
(QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:100-106 modules, :137-168 forward) exists only as a float-simulated
quantised forward, and the Q-EEGNet Wolf C code is not vendored (q-eegnet_wolf/ holds one micro-benchmark).  This file applies
the arithmetic of the edgeEEGNet golden model (edge-eegnet_wolf/python_utils/golden_model.py:153-379,
functional.py:55-287: truncating division, clip to [-128, 127], REORDER_BN pooling) to the EEGNet layer order; what anchors
it to the reference is tests/test_eegnet_order.py::test_integer_forward_tracks_float_sim, which compares every layer with a
float64 re-statement of the module's forward (fake-quantised activations, eegnet.py:137-168) within the tolerance the
reference accepts for its own integer model (10 LSB, edge-eegnet_wolf/test/.../testcase.py).
"""
import numpy as np

from golden_np import clip8, tdiv, xcorr_valid


def forward(p, x, return_layers=False):
    """p: mibminet_b200.convert.EEGNetParams; x: int [B, C, T] -> int8 [B, N]."""
    x = np.asarray(x, dtype=np.int64)
    B, C, T = x.shape
    assert (C, T) == (p.C, p.T)
    Nf, F1, D, K2, F2 = p.Nf, p.F1, p.D, p.F1 * p.D, p.F2
    T8, T64 = T // 8, (T // 8) // 8
    col = lambda v: np.asarray(v, np.int64)[None, :, None]
    # L1: temporal conv per EEG channel (eegnet.py:101-104), float64 matmul is exact here (|v| <= 128 * 127 * 128 < 2^53)
    xp = np.pad(x, ((0, 0), (0, 0), (Nf // 2 - 1, Nf // 2)))
    win = np.lib.stride_tricks.sliding_window_view(xp, Nf, axis=-1)                                    # [B, C, T, Nf]
    v = np.rint(win.astype(np.float64) @ p.w1.astype(np.float64).T).astype(np.int64)                   # [B, C, T, F1]
    y1 = clip8(tdiv(v + np.asarray(p.off1, np.int64), np.asarray(p.fac1, np.int64)))                   # [B, C, T, F1]
    # L2: depthwise spatial conv (eegnet.py:105-109), REORDER_BN pooling (golden_model.py:244-247)
    w2 = p.w2.astype(np.int64)
    a = np.stack([np.einsum("bct,c->bt", y1[..., k // D], w2[k]) for k in range(K2)], 1)               # [B, K2, T]
    off, fac = np.asarray(p.off2, np.int64), np.asarray(p.fac2, np.int64)
    s = np.maximum(a, col(-(off >> 3)))[..., :T8 * 8].reshape(B, K2, T8, 8).sum(-1)
    y2 = clip8(tdiv(s + col(off), col(fac)))
    # L3 (golden_model.py:285-289)
    w = xcorr_valid(np.pad(y2, ((0, 0), (0, 0), (7, 8))), p.w3.astype(np.int64)[None])
    y3 = clip8(tdiv(w, int(p.fac3)))
    # L4 (golden_model.py:330-342)
    e = np.einsum("ki,bit->bkt", p.w4.astype(np.int64), y3)
    off, fac = np.asarray(p.off4, np.int64), np.asarray(p.fac4, np.int64)
    s = np.maximum(e, col(-(off >> 3)))[..., :T64 * 8].reshape(B, F2, T64, 8).sum(-1)
    y4 = clip8(tdiv(s + col(off), col(fac)))
    # L5 (golden_model.py:374-379)
    z = y4.reshape(B, -1) @ p.w5.astype(np.int64).T + p.b5.astype(np.int64)[None]
    out = clip8(tdiv(z, int(p.fac5))).astype(np.int8)
    if return_layers:
        return out, {"y1": y1.astype(np.int8), "y2": y2.astype(np.int8), "y3": y3.astype(np.int8), "y4": y4.astype(np.int8)}
    return out


def float_sim(sd, x_int, C, T, F1, D, N=4):
    """float64 restatement of EEGNet.forward (eegnet.py:137-168) with floorToZero STE activations (ste_ops.py:88-102) and
    the frozen weights; x_int: int [B, C, T] (already quantised input, quant1 scale).  Returns per-layer activations in
    integer units (value / scale * 127) for comparison with the integer pipeline."""
    g = lambda k: np.asarray(sd[k], np.float64)
    s = [None] + [float(g(f"quant{i}.absMaxValue").ravel()[0]) for i in range(1, 7)]
    fq = lambda h, sc: np.trunc(np.clip(h / sc, -1.0, 1.0) * 127.0) / 127.0 * sc
    K2 = F1 * D
    Nf = sd["conv1.weightFrozen"].shape[-1]
    T8, T64 = T // 8, (T // 8) // 8
    B = x_int.shape[0]

    def bn(h, i, axis):
        sc = g(f"batch_norm{i}.weight") / np.sqrt(g(f"batch_norm{i}.running_var") + 1e-3)
        sh = [1] * h.ndim
        sh[axis] = -1
        return h * sc.reshape(sh) + (g(f"batch_norm{i}.bias") - g(f"batch_norm{i}.running_mean") * sc).reshape(sh)

    x = np.asarray(x_int, np.float64) / 127.0 * s[1]
    xp = np.pad(x, ((0, 0), (0, 0), (Nf // 2 - 1, Nf // 2)))
    win = np.lib.stride_tricks.sliding_window_view(xp, Nf, axis=-1)
    h1 = fq(bn(win @ g("conv1.weightFrozen").reshape(F1, Nf).T, 1, 3), s[2])                          # [B, C, T, F1]
    w2 = g("conv2.weightFrozen").reshape(K2, C)
    h = np.stack([np.einsum("bct,c->bt", h1[..., k // D], w2[k]) for k in range(K2)], 1)
    h = np.maximum(bn(h, 2, 1), 0.0)[..., :T8 * 8].reshape(B, K2, T8, 8).mean(-1)
    h2 = fq(h, s[3])
    xp2 = np.pad(h2, ((0, 0), (0, 0), (7, 8)))
    w3 = g("sep_conv1.weightFrozen").reshape(K2, 16)
    h = sum(xp2[..., j:j + T8] * w3[None, :, j:j + 1] for j in range(16))
    h3 = fq(h, s[4])
    h = np.maximum(bn(np.einsum("ki,bit->bkt", g("sep_conv2.weightFrozen").reshape(-1, K2), h3), 3, 1), 0.0)
    h = h[..., :T64 * 8].reshape(B, -1, T64, 8).mean(-1)
    h4 = fq(h, s[5])
    h = h4.reshape(B, -1) @ g("fc.weightFrozen").T + g("fc.bias")[None]
    h5 = fq(h, s[6])
    u = lambda h, sc: h / sc * 127.0
    return {"y1": u(h1, s[2]), "y2": u(h2, s[3]), "y3": u(h3, s[4]), "y4": u(h4, s[5]), "out": u(h5, s[6])}
