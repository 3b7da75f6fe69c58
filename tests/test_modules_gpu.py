"""The drop-in nn.Modules (mibminet_b200/modules.py) -- the operator surface north_star names first -- on the GPU:
the reference's constructor keywords, `load_state_dict(strict=True)` of the reference's own checkpoints
(tests/golden/export_*/net.npz were written by the reference's QuantLab edgeEEGNet module, oracle/gen_verification.py),
`forward(x[B, 1, C, T] float)`.

  reference: QuantLab/quantlab/{PhysionetMMMI,BCI-CompIV-2a}/edgeEEGNet/edgeEEGnet.py:16-21 (constructor), :137-178 (forward,
             with_stats), :180 (forward_with_tensor_stats); edge-eegnet_torch/edgeEEGnet_quant.py:14-107.
  tolerance: the reference accepts its integer model within 10 quantisation steps of the float-simulated module
             (edge-eegnet_wolf/test/python/golden_model/testcase.py:42,129-153); integer results are bit-exact.
"""
import glob
import json
import os

import numpy as np
import pytest
import torch

import wolf_port
from mibminet_b200 import convert, modules, synth
from mibminet_b200.engine import Int8Net

EXPORTS = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "export_*")))
TOLERANCE = 10      # golden_model/testcase.py:42


def _load(d):
    with open(os.path.join(d, "config.json")) as f:
        kw = json.load(f)["indiv"]["net"]["params"]
    sd = dict(np.load(os.path.join(d, "net.npz")))
    return kw, sd, np.load(os.path.join(d, "verification.npz"))


def _module(kw, sd):
    # the reference builds its net as edgeEEGNet(**config["indiv"]["net"]["params"]) (QuantLab/main.py); Nf is not part of the
    # exported config: the Wolf tool-chain fixes it at 64 (golden_model.py:216-219)
    m = modules.edgeEEGNet(Nf=64, **kw)
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    return m


@pytest.mark.parametrize("d", EXPORTS, ids=os.path.basename)
def test_module_state_dict_keys_match_the_reference_checkpoint(d):
    """CPU: every key of the reference checkpoint, nothing else, same shapes (strict load)."""
    kw, sd, _ = _load(d)
    m = _module(kw, sd)
    own = m.state_dict()
    assert sorted(own) == sorted(sd)
    for k, v in sd.items():
        assert tuple(own[k].shape) == tuple(np.asarray(v).shape), k
    with pytest.raises(RuntimeError):
        m.forward(torch.zeros(1, 1, kw["C"], kw["T"]))           # no CPU path


def test_module_refuses_what_the_integer_toolchain_cannot_run():
    """golden_model.py:61-65,83,216-219: 255 levels, floorToZero, F2 == F1 * D, Nf == 64."""
    ok = dict(F1=16, D=1, C=64, T=480, N=4, Nf=64, floorToZero=True)
    modules.edgeEEGNet(**ok)
    for bad in (dict(floorToZero=False), dict(Nf=128), dict(weightInqNumLevels=127), dict(actSTENumLevels=15), dict(F2=32),
                dict(D=2)):
        with pytest.raises(ValueError):
            modules.edgeEEGNet(**{**ok, **bad})
    with pytest.raises(ValueError):
        modules.edgeEEGNet()                                        # the reference's own defaults: floorToZero=False, Nf=128
    with pytest.raises(ValueError):
        modules.EEGNet(floorToZero=False)


@pytest.mark.gpu
@pytest.mark.parametrize("d", EXPORTS, ids=os.path.basename)
def test_module_forward_against_oracle_and_reference_taps(d):
    kw, sd, ver = _load(d)
    C, T = kw["C"], kw["T"]
    m = _module(kw, sd).cuda()
    p = convert.params_from_state_dict(sd, C=C, T=T)
    arrays = p.to_wolf_arrays()

    # (i) forward_int8 == the integer oracle, bit for bit, in both input ranks
    x8 = synth.make_trials(257, C, T, seed=5)
    ref = wolf_port.forward(arrays, C, T, x8)
    xd = torch.from_numpy(x8).cuda()
    assert np.array_equal(m.forward_int8(xd).cpu().numpy(), ref)
    assert np.array_equal(m.forward_int8(xd[:, None]).cpu().numpy(), ref)
    assert np.array_equal(Int8Net(p, device="cuda:0").forward(xd).cpu().numpy(), ref)

    # (ii) float in, float out: the reference module's own sample within its tolerance of the float-simulated logits,
    #      and exactly the de-quantised integer logits of the oracle on the quantised sample
    xf = torch.from_numpy(np.asarray(ver["input"], np.float32).reshape(1, 1, C, T)).cuda()
    out = m(xf)
    assert out.shape == (1, kw["N"]) and out.dtype == torch.float32 and out.is_cuda
    step = p.output_scale / 127.0
    err = np.abs(out.cpu().numpy() - ver["output_quant"].reshape(1, -1)).max() / step
    assert err <= TOLERANCE + 1e-3, err
    xq = convert.quantize_input(np.asarray(ver["input"], np.float32).reshape(1, C, T), p.input_scale)
    yq = wolf_port.forward(arrays, C, T, xq)
    assert np.allclose(out.cpu().numpy(), convert.dequantize(yq, p.output_scale), rtol=1e-6, atol=0)
    # a float batch [B, 1, C, T]: quant1 on the GPU == the reference's quantiser
    rng = np.random.default_rng(3)
    xb = (rng.standard_normal((33, 1, C, T)) * 0.4 * p.input_scale).astype(np.float32)
    yb = m(torch.from_numpy(xb).cuda()).cpu().numpy()
    refb = wolf_port.forward(arrays, C, T, convert.quantize_input(xb[:, 0], p.input_scale))
    assert np.allclose(yb, convert.dequantize(refb, p.output_scale), rtol=1e-6, atol=0)

    # (iii) with_stats / forward_with_tensor_stats: the reference's (stats, output) tuple, edgeEEGnet.py:169-178
    stats, y2 = m(xf, with_stats=True)
    assert [n for n, _ in stats] == ["conv1_w", "conv2_w", "sep_conv1_w", "sep_conv2_w", "fc_w", "fc_b"]
    assert all(torch.is_tensor(t) for _, t in stats) and torch.equal(y2, out)
    stats2, y3 = m.forward_with_tensor_stats(xf)
    assert len(stats2) == 6 and torch.equal(y3, out)


@pytest.mark.gpu
def test_module_cache_follows_in_place_edits():
    """The packed integer network is rebuilt when a buffer changes in place (quant6 rescales the output)."""
    kw, sd, ver = _load(EXPORTS[0])
    m = _module(kw, sd).cuda()
    xf = torch.from_numpy(np.asarray(ver["input"], np.float32).reshape(1, 1, kw["C"], kw["T"])).cuda()
    y0 = m(xf).clone()
    e0 = m.integer_network()
    assert m.integer_network() is e0                     # unchanged state: cached
    m.quant1.absMaxValue.mul_(0.5)                       # input quantiser twice as fine: different integers
    sd2 = {k: v.detach().cpu().numpy() for k, v in m.state_dict().items()}
    p2 = convert.params_from_state_dict(sd2, C=kw["C"], T=kw["T"])
    y1 = m(xf)
    assert m.integer_network() is not e0
    xq = convert.quantize_input(np.asarray(ver["input"], np.float32).reshape(1, kw["C"], kw["T"]), p2.input_scale)
    ref = convert.dequantize(wolf_port.forward(p2.to_wolf_arrays(), kw["C"], kw["T"], xq), p2.output_scale)
    assert np.allclose(y1.cpu().numpy(), ref, rtol=1e-6, atol=0)
    assert y0.shape == y1.shape


@pytest.mark.gpu
def test_edgeeegnetquant_signature_shell():
    """edgeEEGNetQuant keeps the constructor / forward signature of edge-eegnet_torch/edgeEEGnet_quant.py:14,81 (x[B, C, T]) and
    computes with the Wolf integer scheme (its qnnpack arithmetic is third-party and unpinned by the reference)."""
    C, T = 22, 1125
    sd = synth.make_state_dict(C=C, T=T, seed=4)
    p = convert.params_from_state_dict(sd, C=C, T=T)
    q = modules.edgeEEGNetQuant(qconfig=None, F1=16, D=1, C=C, T=T, N=4)
    with pytest.raises(RuntimeError):
        q(torch.zeros(1, C, T, device="cuda"))
    for src in (p, sd):
        q.load_integer_params(src)
        x8 = synth.make_trials(65, C, T, seed=6)
        y = q(torch.from_numpy(x8).cuda())
        ref = convert.dequantize(wolf_port.forward(p.to_wolf_arrays(), C, T, x8), p.output_scale)
        assert y.shape == (65, 4) and np.allclose(y.cpu().numpy(), ref, rtol=1e-6, atol=0)
    xf = (np.random.default_rng(0).standard_normal((5, C, T)) * 0.3 * p.input_scale).astype(np.float32)
    yf = q(torch.from_numpy(xf).cuda()).cpu().numpy()
    reff = convert.dequantize(wolf_port.forward(p.to_wolf_arrays(), C, T, convert.quantize_input(xf, p.input_scale)), p.output_scale)
    assert np.allclose(yf, reff, rtol=1e-6, atol=0)
    with pytest.raises(RuntimeError):
        q(torch.zeros(1, C, T))                              # CPU tensor: no CPU path
