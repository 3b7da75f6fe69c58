"""`verification.npz` of a QuantLab export.  Fixtures tests/golden/export_*/ were written by
oracle/gen_verification.py from the REFERENCE's own QuantLab edgeEEGNet module (float-simulated forward); the check is the
reference's edge-eegnet_wolf/test/python/golden_model/testcase.py:67-93 (integer model within TOLERANCE steps of the taps)."""
import glob
import os

import numpy as np
import pytest

import golden_np
from mibminet_b200 import convert

EXPORTS = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "export_*")))
NAMES = (("layer1_activ", "y1", "quant2"), ("layer2_activ", "y2", "quant3"), ("layer3_activ", "y3", "quant4"),
         ("layer4_activ", "y4", "quant5"), ("output_quant", "out", "quant6"))


def _files(d):
    return os.path.join(d, "net.npz"), os.path.join(d, "config.json"), os.path.join(d, "verification.npz")


def test_exports_present():
    assert len(EXPORTS) >= 2


@pytest.mark.parametrize("d", EXPORTS, ids=os.path.basename)
def test_oracle_within_tolerance_of_reference_float_taps(d):
    """Pins converter + integer oracle to the reference's QuantLab module, layer by layer (chained like test_model)."""
    net, cfg, ver = _files(d)
    p = convert.params_from_export(net, cfg)
    data = np.load(ver)
    # quant1 of the module emits grid values k * s / 127; truncating them again may land one step short in float32
    assert np.abs(convert.quantize_to_int(data["input"], p.input_scale)
                  - convert.quantize_to_int(data["input_quant"][0, 0], p.input_scale)).max() <= 1
    x = convert.quantize_input(data["input"], p.input_scale)[None]
    y, taps = golden_np.forward(p.to_wolf_arrays(), p.C, p.T, x, reorder_bn=True, return_layers=True)
    taps = dict(taps, out=y)
    for name, key, q in NAMES:
        exp = convert.quantize_to_int(data[name].reshape(taps[key][0].shape), p.scales[q])
        err = np.abs(taps[key][0].astype(np.int64) - exp).max()
        assert err <= 10, (name, err)              # the reference's TOLERANCE
        assert err <= (6 if key == "out" else 3), (name, err)   # measured on these fixtures (errors add up layer by layer)


@pytest.mark.gpu
@pytest.mark.parametrize("d", EXPORTS, ids=os.path.basename)
def test_verify_export_on_gpu(d, tmp_path):
    from mibminet_b200 import verify
    from mibminet_b200.engine import Int8Net
    net_npz, cfg, ver = _files(d)
    rep = verify.verify_export(net_npz, cfg, ver)
    assert sorted(rep) == ["1", "2", "3", "4", "5"]
    assert all(r["result"] for r in rep.values()), rep

    # GPU taps are bit-exact against the oracle, and a verification file written from them reads back to the same integers
    p = convert.params_from_export(net_npz, cfg)
    data = np.load(ver)
    x = convert.quantize_input(data["input"], p.input_scale)
    y, taps = golden_np.forward(p.to_wolf_arrays(), p.C, p.T, x[None], reorder_bn=True, return_layers=True)
    taps = dict(taps, out=y)
    net = Int8Net(p)
    got = verify.gpu_taps(net, x)
    for _n, key, _q in NAMES:
        assert np.array_equal(got[key], taps[key][0].reshape(got[key].shape)), key
    path = str(tmp_path / "verification.npz")
    verify.write_verification(path, net, data["input"])
    back = np.load(path)
    assert back["layer1_activ"].shape == data["layer1_activ"].shape
    assert back["layer2_activ"][0, :, 0, :].shape == (p.F2, p.T // 8)          # the slicing of layer3/testcase.py:56
    for name, key, q in NAMES:
        assert back[name].shape == data[name].shape
        assert np.array_equal(convert.quantize_to_int(back[name], p.scales[q]).reshape(got[key].shape), got[key]), name
    rep2 = verify.verify_export(net_npz, cfg, path, tolerance=0, net=net)
    assert all(r["result"] and r["max error"] == 0 for r in rep2.values()), rep2
    net.close()


@pytest.mark.parametrize("scale", [1.0, 0.1, 3.3333333, 11.781072616577148, 71.22114562988281, 1e-3, 12345.678])
def test_written_taps_quantise_back_exactly(scale):
    """verify._dequantize_exact: float32 grid values that `quantize_to_int` (truncation) maps back to every int8 code."""
    from mibminet_b200 import verify
    a = np.arange(-127, 128, dtype=np.int64)
    f = verify._dequantize_exact(a, scale)
    assert f.dtype == np.float32
    assert np.array_equal(convert.quantize_to_int(f, scale), a)
    assert np.all(np.abs(f.astype(np.float64) - a * float(scale) / 127.0) <= 4 * np.spacing(np.abs(f)).astype(np.float64))


def test_expected_taps_shapes():
    from mibminet_b200 import verify
    d = EXPORTS[0]
    net, cfg, ver = _files(d)
    p = convert.params_from_export(net, cfg)
    exp = verify.expected_taps(np.load(ver), p)
    assert exp["y1"].shape == (p.F2, p.T) and exp["y2"].shape == (p.F2, p.T // 8) and exp["y4"].shape == (p.F2, p.T // 64)
    assert exp["out"].shape == (p.N,) and all(np.abs(v).max() <= 127 for v in exp.values())
