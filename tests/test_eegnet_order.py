"""EEGNet / Q-EEGNet layer order: temporal conv first, then depthwise spatial conv.

The reference holds no integer model of this order (QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py is float-simulated).
CPU tests: the integer restatement (oracle/eegnet_np.py) tracks (a) the committed outputs of the reference's own module
(tests/golden/eegnet_order_*.npz, written by oracle/gen_golden_eegnet.py) and (b) a float64 restatement of that forward, layer by
layer, within the 10 LSB the reference accepts for its integer edgeEEGNet model (test/python/golden_model/testcase.py:42).
GPU tests: mibmi_eegnet_forward_int8 is bit-exact against the integer restatement.
"""
import ctypes
import os

import numpy as np
import pytest

import eegnet_np
from conftest import GOLDEN_DIR
from mibminet_b200 import convert, synth

TOL_LSB = 10      # tolerance of the reference's own integer-vs-float test


@pytest.mark.parametrize("tag", ["c22_t1125", "c64_t480"])
def test_integer_forward_tracks_the_reference_module(tag):
    g = np.load(os.path.join(GOLDEN_DIR, f"eegnet_order_{tag}.npz"))
    C, T, F1, D = int(g["C"]), int(g["T"]), int(g["F1"]), int(g["D"])
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, seed=int(g["seed"]), negative_gamma=False)
    p = convert.eegnet_params_from_state_dict(sd, C=C, T=T, F1=F1, D=D)
    out = eegnet_np.forward(p, g["x_int"]).astype(np.float64)
    ref = g["logits_units"]                      # the reference module's logits in int8 units (y / s6 * 127)
    assert np.abs(out - np.clip(ref, -128, 127)).max() <= TOL_LSB
    sim = eegnet_np.float_sim(sd, g["x_int"], C, T, F1, D)["out"]
    assert np.abs(sim - ref).max() <= 2.0        # the float64 restatement vs torch float32


@pytest.mark.parametrize("C,T,F1,D,Nf", [(22, 1125, 8, 2, 128), (64, 480, 8, 2, 128), (22, 1125, 16, 1, 64), (8, 480, 4, 2, 32)])
def test_integer_forward_tracks_float_sim(C, T, F1, D, Nf):
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, Nf=Nf, seed=3, negative_gamma=False)
    p = convert.eegnet_params_from_state_dict(sd, C=C, T=T, F1=F1, D=D)
    x = synth.make_trials(4, C, T, seed=9)
    out, layers = eegnet_np.forward(p, x, return_layers=True)
    sim = eegnet_np.float_sim(sd, x, C, T, F1, D)
    for k in ("y1", "y2", "y3", "y4"):
        assert np.abs(layers[k].astype(np.float64) - sim[k]).max() <= TOL_LSB, k
    assert np.abs(out - sim["out"]).max() <= TOL_LSB


def test_layers_one_and_two_against_plain_loops():
    """the vectorised restatement against scalar Python loops on a tiny net (both BN-gamma signs)"""
    C, T, F1, D, Nf = 3, 64, 4, 2, 8
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, Nf=Nf, seed=5)
    p = convert.eegnet_params_from_state_dict(sd, C=C, T=T, F1=F1, D=D)
    x = synth.make_trials(1, C, T, seed=2)
    _, layers = eegnet_np.forward(p, x, return_layers=True)
    tdiv = lambda a, b: int(abs(a) // abs(b)) * (-1 if (a < 0) != (b < 0) else 1)
    clip = lambda v: max(-128, min(127, v))
    xp = np.pad(x[0].astype(np.int64), ((0, 0), (Nf // 2 - 1, Nf // 2)))
    y1 = np.zeros((F1, C, T), np.int64)
    for f in range(F1):
        for c in range(C):
            for t in range(T):
                v = int((xp[c, t:t + Nf] * p.w1[f].astype(np.int64)).sum())
                y1[f, c, t] = clip(tdiv(v + int(p.off1[f]), int(p.fac1[f])))
    assert np.array_equal(np.moveaxis(layers["y1"][0], -1, 0), y1)
    for k in range(F1 * D):
        for q in range(T // 8):
            s = 0
            for e in range(8):
                a = int((y1[k // D, :, 8 * q + e] * p.w2[k].astype(np.int64)).sum())
                s += max(a, -(int(p.off2[k]) >> 3))
            assert layers["y2"][0, k, q] == clip(tdiv(s + int(p.off2[k]), int(p.fac2[k])))


def test_create_validates_without_a_gpu():
    from mibminet_b200 import _lib
    L = _lib.lib()
    h = ctypes.c_void_p()
    assert L.mibmi_eegnet_create(None, 0, ctypes.byref(h)) == _lib.ERR_NULL
    p = _lib.EEGNetParamsC()
    assert L.mibmi_eegnet_create(ctypes.byref(p), 0, ctypes.byref(h)) == _lib.ERR_NULL
    assert L.mibmi_eegnet_forward_int8(None, None, 1, None, None) == _lib.ERR_NULL
    assert L.mibmi_eegnet_destroy(None) == _lib.OK


# ---------------------------------------------------------------------------------------------------------
def _gpu_net(C, T, F1, D, Nf, seed):
    import torch
    from mibminet_b200.engine import EEGNetInt8
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, Nf=Nf, seed=seed)
    p = convert.eegnet_params_from_state_dict(sd, C=C, T=T, F1=F1, D=D)
    return sd, p, EEGNetInt8(p, device="cuda:0"), torch


@pytest.mark.gpu
@pytest.mark.parametrize("C,T,F1,D,Nf,B", [(22, 1125, 8, 2, 128, 96), (64, 480, 8, 2, 128, 96), (22, 1125, 16, 1, 64, 40),
                                           (8, 480, 4, 2, 32, 301), (19, 484, 8, 1, 128, 33), (3, 64, 4, 2, 8, 7)])
def test_gpu_bit_exact_with_taps(C, T, F1, D, Nf, B):
    sd, p, net, torch = _gpu_net(C, T, F1, D, Nf, seed=C + T)
    x = synth.make_trials(B, C, T, seed=4)
    ref, layers = eegnet_np.forward(p, x, return_layers=True)
    xd = torch.from_numpy(x).cuda()
    y = net.forward(xd).cpu().numpy()
    assert np.array_equal(y, ref)
    y2, taps = net.forward_taps(xd)
    assert np.array_equal(y2.cpu().numpy(), ref)
    for k in ("y2", "y3", "y4"):
        assert np.array_equal(taps[k].cpu().numpy(), layers[k]), k
    assert net.launch_count == 2
    # F1 = 8 filters of 128 taps (the QuantLab EEGNet configuration): both contractions on tcgen05.mma
    assert net.last_kernel == ("eegnet_tc_kernel" if (F1 == 8 and Nf == 128) else "eegnet_mma_kernel" if Nf % 32 == 0 else "eegnet_forward_kernel")


@pytest.mark.gpu
@pytest.mark.parametrize("C,T,D,B", [(22, 1125, 2, 1500), (64, 480, 2, 1000), (19, 484, 1, 777), (5, 128, 2, 300)])
def test_gpu_tcgen05_kernel_over_many_trials(C, T, D, B, monkeypatch):
    """eegnet_tc_kernel with several trials per persistent CTA (input stages, accumulator buffers and y2 buffers wrap around):
    equal to the oracle on a sample and to the mma.sync kernel on every trial, logits and taps."""
    sd, p, net, torch = _gpu_net(C, T, 8, D, 128, seed=7)
    x = synth.make_trials(B, C, T, seed=8, dist="gauss")
    xd = torch.from_numpy(x).cuda()
    y, taps = net.forward_taps(xd)
    assert net.last_kernel == "eegnet_tc_kernel"
    y1 = net.forward(xd[1:])                       # another alignment of the trials (C * T odd shifts the rows)
    monkeypatch.setenv("MIBMI_EEGNET_TC", "0")
    yr, tr = net.forward_taps(xd)
    assert net.last_kernel == "eegnet_mma_kernel"
    for k in ("y2", "y3", "y4"):
        assert torch.equal(taps[k], tr[k]), k
    assert torch.equal(y, yr) and torch.equal(y1, yr[1:])
    assert np.array_equal(y[:48].cpu().numpy(), eegnet_np.forward(p, x[:48]))


@pytest.mark.gpu
def test_gpu_tcgen05_kernel_with_factors_outside_the_magic_range():
    """|factor| < 4 in layers 2 - 5 is outside the proven range of the magic division: eegnet_tc_kernel then divides in hardware there
    (layer 1 keeps its own magic) and must stay exact; negative factors go through the sign path."""
    import dataclasses
    from mibminet_b200.engine import EEGNetInt8
    sd, p, _, torch = _gpu_net(22, 1125, 8, 2, 128, seed=12)
    x = synth.make_trials(200, 22, 1125, seed=13)
    fac2 = np.asarray(p.fac2).copy(); fac2[::3] *= -1
    fac4 = np.asarray(p.fac4).copy(); fac4[1::2] *= -1
    for kw in ({"fac3": -int(p.fac3), "fac2": fac2.astype(np.int32), "fac4": fac4.astype(np.int32), "fac5": -int(p.fac5)},
               {"fac3": 3}, {"fac5": -2}):
        p2 = dataclasses.replace(p, **kw)
        net = EEGNetInt8(p2, device="cuda:0")
        y = net.forward(torch.from_numpy(x).cuda()).cpu().numpy()
        assert net.last_kernel == "eegnet_tc_kernel"
        assert np.array_equal(y, eegnet_np.forward(p2, x)), kw.keys()


@pytest.mark.gpu
def test_gpu_forward_captured_in_a_cuda_graph():
    """The EEGNet-order forward only launches (no allocation, no synchronisation): captured once, replayed on new buffer contents."""
    sd, p, net, torch = _gpu_net(22, 1125, 8, 2, 128, seed=14)
    B = 150                                          # more trials than SMs: the persistent CTAs of eegnet_tc_kernel loop
    xs = [synth.make_trials(B, 22, 1125, seed=140 + i) for i in range(2)]
    xd = torch.from_numpy(xs[0]).cuda()
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        net.forward(xd)                              # first call outside the capture (function attributes)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        y = net.forward(xd)
    assert net.last_kernel == "eegnet_tc_kernel"
    for x in xs:
        ref = net.forward(torch.from_numpy(x).cuda()).cpu().numpy()       # a plain call on other buffers
        xd.copy_(torch.from_numpy(x).cuda())
        y.fill_(-77)
        graph.replay()
        torch.cuda.synchronize()
        assert np.array_equal(y.cpu().numpy(), ref)
        assert np.array_equal(ref[:24], eegnet_np.forward(p, x[:24]))


@pytest.mark.gpu
def test_gpu_saturating_inputs_and_empty_batch():
    sd, p, net, torch = _gpu_net(22, 1125, 8, 2, 128, seed=1)
    x = np.stack([np.full((22, 1125), 127, np.int8), np.full((22, 1125), -128, np.int8),
                  np.tile(np.array([127, -128], np.int8), (22, 563))[:, :1125]])
    assert np.array_equal(net.forward(torch.from_numpy(x).cuda()).cpu().numpy(), eegnet_np.forward(p, x))
    assert net.forward(torch.empty((0, 22, 1125), dtype=torch.int8, device="cuda")).shape == (0, 4)


@pytest.mark.gpu
def test_gpu_small_factors_use_the_hardware_division():
    """|l1_factor| < 4 is outside the proven range of the magic division: the kernel instantiation with the hardware
    division must give the same exact results."""
    import dataclasses
    from mibminet_b200.engine import EEGNetInt8
    sd, p, _, torch = _gpu_net(22, 1125, 8, 2, 128, seed=2)
    fac1 = np.asarray(p.fac1).copy()
    fac1[::2] = [3, -2, 1, -1]
    p2 = dataclasses.replace(p, fac1=fac1.astype(np.int32))
    x = synth.make_trials(24, 22, 1125, seed=6)
    net = EEGNetInt8(p2, device="cuda:0")
    assert np.array_equal(net.forward(torch.from_numpy(x).cuda()).cpu().numpy(), eegnet_np.forward(p2, x))


@pytest.mark.gpu
def test_gpu_module_drop_in_float_in_float_out():
    import torch
    from mibminet_b200.modules import EEGNet
    C, T, F1, D = 22, 1125, 8, 2
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, seed=11, negative_gamma=False)
    m = EEGNet(F1=F1, D=D, C=C, T=T, N=4, weightInqSchedule={"1": 1.0}, floorToZero=True)
    m.load_state_dict(sd)
    m = m.cuda()
    g = np.load(os.path.join(GOLDEN_DIR, "eegnet_order_c22_t1125.npz"))
    x = torch.from_numpy(g["x_int"].astype(np.float32) / 127.0 * float(sd["quant1.absMaxValue"][0]))[:, None].cuda()
    y = m(x).cpu().numpy()
    s6 = float(sd["quant6.absMaxValue"][0])
    assert np.abs(y / s6 * 127.0 - np.clip(g["logits_units"], -128, 127)).max() <= TOL_LSB    # vs the reference's own module
