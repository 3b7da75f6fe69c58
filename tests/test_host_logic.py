"""CPU tests of the host-side logic of libmibmi.so: ABI surface, validation, magic-number division,
the fused-kernel parameter packing (through a lane-level emulation of the kernel), converters."""
import ctypes
import os
import re

import numpy as np
import pytest

import golden_np
from conftest import ROOT, golden_files, load_golden
from fused_emulator import forward_trial, parse, tdiv_magic
from mibminet_b200 import _lib, convert, synth


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "mibmi.h")).read()
    declared = sorted(set(re.findall(r"MIBMI_API\s+[\w\s\*]+?\b(mibmi_\w+)\s*\(", hdr)))
    assert declared == sorted(_lib.SYMBOLS)
    L = _lib.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert L.mibmi_version() == 100
    assert L.mibmi_strerror(0) == b"ok" and b"CUDA" in L.mibmi_strerror(_lib.ERR_CUDA)


def test_validation_errors_need_no_gpu():
    fx = load_golden(golden_files()[0])
    L = _lib.lib()
    h = ctypes.c_void_p()
    p, keep = _lib.make_params(fx["arrays"], fx["C"], fx["T"])
    assert L.mibmi_create(None, 0, ctypes.byref(h)) == _lib.ERR_NULL
    p.Nf = 128
    assert L.mibmi_create(ctypes.byref(p), 0, ctypes.byref(h)) == _lib.ERR_SHAPE     # golden_model.py:216-219
    p.Nf = 64
    p.D, p.F2 = 2, 32
    assert L.mibmi_create(ctypes.byref(p), 0, ctypes.byref(h)) == _lib.ERR_SHAPE     # golden_model.py:169
    p.D, p.F2 = 1, 16
    bad = dict(fx["arrays"])
    bad["l2_factor"] = bad["l2_factor"].copy()
    bad["l2_factor"][3] = 0
    p2, keep2 = _lib.make_params(bad, fx["C"], fx["T"])
    assert L.mibmi_create(ctypes.byref(p2), 0, ctypes.byref(h)) == _lib.ERR_PARAM
    assert L.mibmi_forward_int8(None, None, 1, 0, None, None) == _lib.ERR_NULL
    assert L.mibmi_forward_int8_host(None, None, 0, 0, None) == _lib.ERR_NULL


@pytest.mark.parametrize("d", [4, 5, 7, 8, 9, 63, 64, 65, 277, 852, 1145, 4096, 5771, 65535, 65536, 1 << 20, 16777259,
                               (1 << 29) + 3, (1 << 31) - 1])
def test_magic_division_exact(d):
    st, m, s = _lib.debug_magic(d)
    assert st == 0 and 0 < m < 2 ** 31 and s >= 0
    rng = np.random.default_rng(d)
    lim = 2 ** 29 - 1
    xs = np.concatenate([np.arange(-70000, 70000), rng.integers(-lim, lim + 1, 400000),
                         np.array([lim, -lim, lim - 1, -lim + 1]),
                         (np.arange(-300, 300)[:, None] * d + np.array([-1, 0, 1])[None]).ravel()])
    xs = xs[np.abs(xs) <= lim]
    assert np.array_equal(tdiv_magic(xs, m, s), golden_np.tdiv(xs, d))


def test_magic_division_exhaustive_small_divisor():
    for d in (4, 6, 277):
        st, m, s = _lib.debug_magic(d)
        xs = np.arange(-(1 << 23), 1 << 23)   # every accumulator value layer 2 can produce
        assert np.array_equal(tdiv_magic(xs, m, s), golden_np.tdiv(xs, d))


def test_magic_rejects_small_divisors():
    for d in (0, 1, 2, 3, -5):
        assert _lib.debug_magic(d)[0] == _lib.ERR_PARAM


def test_fused_packing_emulation_matches_reference(golden):
    fx = golden
    blob, why = _lib.debug_pack_fused(fx["arrays"], fx["C"], fx["T"])
    if fx["T"] % 16:
        assert blob is None and "instantiation" in why
        return
    assert blob is not None, why
    net = parse(blob)
    n = 2 if fx["C"] > 32 else 1
    for i in range(n):
        out, taps = forward_trial(net, fx["C"], fx["T"], fx["x"][i])
        for name in ("y1", "y2", "y3", "y4"):
            assert np.array_equal(taps[name], fx[f"golden_reorder_{name}"][i]), name
        assert np.array_equal(out, fx["golden_reorder"][i])


def test_fused_packing_refuses_what_it_cannot_run():
    fx = load_golden(golden_files()[0])
    assert _lib.debug_pack_fused(fx["arrays"], fx["C"], fx["T"], reorder_bn=False)[0] is None
    a = dict(fx["arrays"])
    a["l1_factor"] = a["l1_factor"].copy()
    a["l1_factor"][0] = 3
    blob, why = _lib.debug_pack_fused(a, fx["C"], fx["T"])
    assert blob is None and "l1_factor" in why
    a = dict(fx["arrays"])
    a["l1_offset"] = a["l1_offset"].copy()
    a["l1_offset"][5] = 2 ** 30
    blob, why = _lib.debug_pack_fused(a, fx["C"], fx["T"])
    assert blob is None and "range" in why


def test_converter_reproduces_reference_export(golden):
    """mibminet_b200.convert (ours) == the arrays the reference's gen_net_header.py wrote."""
    fx = golden
    p = convert.params_from_state_dict(fx["sd"], C=fx["C"], T=fx["T"])
    ours = p.to_wolf_arrays()
    for k, v in fx["arrays"].items():
        assert np.array_equal(np.asarray(ours[k]).ravel(), np.asarray(v).ravel()), k
    rt = convert.WolfParams.from_wolf_arrays(ours, C=fx["C"], T=fx["T"]).to_wolf_arrays()
    for k in ours:
        assert np.array_equal(np.asarray(rt[k]), np.asarray(ours[k])), k


def test_input_quantisation_and_wolf_layout():
    rng = np.random.default_rng(0)
    x = rng.normal(0, 0.6, size=(3, 22, 40)).astype(np.float32)
    q = convert.quantize_input(x, 1.3)
    exp = np.trunc(np.clip(x / np.float32(1.3), -1, 1) * np.float32(254) / np.float32(2)).astype(np.int8)
    assert np.array_equal(q, exp) and q.min() >= -127
    w = convert.to_wolf_input(q)
    assert w.shape == (3, 40, 24) and np.array_equal(w[..., :22], np.swapaxes(q, 1, 2)) and not w[..., 22:].any()


def test_synthetic_nets_exercise_negative_factors_and_saturation():
    sd = synth.make_state_dict(C=64, T=480, seed=0)
    p = convert.params_from_state_dict(sd, C=64, T=480)
    assert (p.fac1 < 0).any() and (p.fac2 < 0).any()
    x = synth.make_trials(8, 64, 480, seed=1)
    y, taps = golden_np.forward(p.to_wolf_arrays(), 64, 480, x, return_layers=True)
    assert (np.abs(taps["y1"].astype(int)) >= 127).any() and len(np.unique(y)) > 8


def test_torch_op_is_registered_and_cuda_only(golden):
    """torch.ops.mibminet.forward_int8: schema, blob round trip, no CPU kernel."""
    import torch
    from mibminet_b200 import torch_op
    fx = golden
    dims = [fx["C"], fx["T"], 16, 1, 4]
    blob = torch_op.pack_blob(fx["arrays"], dims)
    back = torch_op.unpack_blob(blob, dims)
    for name, _ in torch_op.BLOB_FIELDS:
        assert np.array_equal(np.asarray(back[name]).ravel(), np.asarray(fx["arrays"][name]).ravel()), name
    assert "forward_int8(Tensor x_int8, Tensor blob, int[] dims, bool reorder_bn) -> Tensor" in str(
        torch.ops.mibminet.forward_int8.default._schema)
    x = torch.from_numpy(fx["x"][:2])
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.mibminet.forward_int8(x, blob, dims, True)          # CPU tensor: no kernel registered
    y = torch.ops.mibminet.forward_int8(x.to("meta"), blob, dims, True)
    assert y.shape == (2, 4) and y.dtype == torch.int8
    with pytest.raises(ValueError):
        torch_op.unpack_blob(blob[:-1], dims)
