"""CPU tests: the oracle (oracle/golden_np.py, oracle/wolf_port.c) against the reference's outputs
(tests/golden/*.npz, produced by oracle/gen_golden.py from the reference's GoldenModel and C model),
plus seeded re-expressions of the reference's primitive known-answer tests
(edge-eegnet_wolf/test/cl/func/{dotp,conv,xcorr,conv_scale,xcorr_scale,transform}/testcase.py)."""
import numpy as np
import pytest

import golden_np
import wolf_port
from conftest import golden_files


def test_fixtures_present():
    assert len(golden_files()) >= 8


@pytest.mark.parametrize("reorder", [True, False])
def test_numpy_oracle_matches_reference_golden(golden, reorder):
    fx = golden
    sfx = "reorder" if reorder else "plain"
    # c_layer4_noclip=False follows the Python golden (golden_model.py:338-339)
    y, taps = golden_np.forward(fx["arrays"], fx["C"], fx["T"], fx["x"], reorder_bn=reorder,
                                c_layer4_noclip=False, return_layers=True)
    assert np.array_equal(y, fx[f"golden_{sfx}"])
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name], fx[f"golden_{sfx}_{name}"]), name


def test_numpy_oracle_matches_reference_c(golden):
    fx = golden
    y = golden_np.forward(fx["arrays"], fx["C"], fx["T"], fx["x"], reorder_bn=True)
    assert np.array_equal(y, fx["c_default"])
    assert np.array_equal(y, fx["c_nosimd"])
    # plain variant follows the C (layer4.c:115-117: no per-element clip)
    yp = golden_np.forward(fx["arrays"], fx["C"], fx["T"], fx["x"], reorder_bn=False, c_layer4_noclip=True)
    assert np.array_equal(yp, fx["c_plain"])


@pytest.mark.parametrize("reorder", [True, False])
def test_c_port_matches_reference(golden, reorder):
    fx = golden
    y, taps = wolf_port.forward(fx["arrays"], fx["C"], fx["T"], fx["x"], reorder_bn=reorder, return_layers=True)
    if reorder:
        assert np.array_equal(y, fx["golden_reorder"])
        assert np.array_equal(y, fx["c_default"])
        for name in ("y1", "y2", "y3", "y4"):
            assert np.array_equal(taps[name], fx[f"golden_reorder_{name}"]), name
    else:
        assert np.array_equal(y, fx["c_plain"])
        for name in ("y1", "y2", "y3"):   # layers 1-3 agree between C and golden in plain mode
            assert np.array_equal(taps[name], fx[f"golden_plain_{name}"]), name


def test_c_port_equals_numpy_oracle_random_batch():
    from mibminet_b200 import convert, synth
    for C, T, seed in ((64, 480, 11), (22, 1125, 12), (8, 480, 13)):
        arrays = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=seed), C=C, T=T).to_wolf_arrays()
        x = synth.make_trials(24, C, T, seed=seed, dist="gauss" if seed % 2 else "uniform")
        for rb in (True, False):
            assert np.array_equal(wolf_port.forward(arrays, C, T, x, reorder_bn=rb),
                                  golden_np.forward(arrays, C, T, x, reorder_bn=rb))


# ---- primitive known-answer tests (seeded; the reference's are unseeded) ------------------------------

@pytest.mark.parametrize("n", [22, 24, 1024, 1025, 1026, 1027, 1031])
def test_dotp(n):  # test/cl/func/dotp/testcase.py:41-48,66
    rng = np.random.default_rng(n)
    a, b = rng.integers(-128, 128, n), rng.integers(-128, 128, n)
    got = golden_np.xcorr_valid(a[None], b[None])
    assert got.shape == (1, 1) and int(got[0, 0]) == int(np.dot(a, b))


@pytest.mark.parametrize("la,lb", [(155, 16), (1188, 64), (4096, 128)])
def test_xcorr_and_conv(la, lb):  # test/cl/func/{xcorr,conv}/testcase.py
    rng = np.random.default_rng(la)
    a, b = rng.integers(-128, 128, la), rng.integers(-128, 128, lb)
    assert np.array_equal(golden_np.xcorr_valid(a[None], b[None])[0], np.correlate(a, b, mode="valid"))
    assert np.array_equal(golden_np.xcorr_valid(a[None], b[::-1][None])[0], np.convolve(a, b, mode="valid"))


@pytest.mark.parametrize("la,lb", [(155, 16), (1188, 64)])
def test_conv_scale(la, lb):  # test/cl/func/conv_scale/testcase.py:41-51,62-66
    rng = np.random.default_rng(la + 1)
    a, b = rng.integers(-128, 128, la), rng.integers(-128, 128, lb)
    div_factor = 128 * lb // 8
    offset = 10 * div_factor
    r = np.convolve(a, b, mode="valid")
    exp = np.clip(((r + offset) / div_factor).astype(int), -128, 127)
    got = golden_np.clip8(golden_np.tdiv(golden_np.xcorr_valid(a[None], b[::-1][None])[0] + offset, div_factor))
    assert np.array_equal(got, exp)


@pytest.mark.parametrize("n", [1024, 1025, 1026, 1027])
def test_transform(n):  # test/cl/func/transform/testcase.py:44-51 (apply_factor_offset, clip_balanced=False)
    rng = np.random.default_rng(n)
    x = rng.integers(-(2 ** 20), 2 ** 20, n)
    for div, bias in ((37, 0), (-91, 1234), (1, -5), (100000, 77)):
        exp = np.clip(((x + bias) / div).astype(int), -128, 127)   # functional.py:81-91
        assert np.array_equal(golden_np.clip8(golden_np.tdiv(x + bias, div)), exp)


def test_tdiv_truncates_toward_zero():
    a = np.array([7, -7, 7, -7, 0, -1, 1, 8388607, -8388608])
    b = np.array([2, 2, -2, -2, 5, 5, -5, -852, 3])
    assert golden_np.tdiv(a, b).tolist() == [3, -3, -3, 3, 0, 0, 0, -9845, -2796202]
