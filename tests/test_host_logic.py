"""CPU tests of the host-side logic of libmibmi.so: ABI surface, validation, magic-number division,
the fused-kernel parameter packing (layout of the tcgen05 operands), converters."""
import ctypes
import os
import re

import numpy as np
import pytest

import golden_np
from conftest import ROOT, golden_files, load_golden
from mibminet_b200 import _lib, convert, synth


def tdiv_magic(x, m, s):
    """(mulhi32(x, m) >> s) + (x < 0) on int64 arrays: the device-side division of the fused kernels."""
    x = np.asarray(x, dtype=np.int64)
    return ((x * int(m)) >> 32 >> int(s)) + (x < 0)


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


def _clip8(v):
    return np.clip(v, -128, 127)


FAST1_DIVISORS = [4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 129, 255, 256, 257, 1023, 1025, 4095, 4097,
                  65535, 65536, 65537, (1 << 20) - 1, (1 << 20) + 1, (1 << 24) + 1]


@pytest.mark.parametrize("d", FAST1_DIVISORS)
def test_fast1_division_adversarial(d):
    """The shift-free layer-1 form (FAST1): clip8(mulhi32(x, m) + (x < 0)) == clip8(trunc(x / d)) on the WHOLE proven range
    |x| <= xmax, for every xmax the host accepts: every value of the +-131 d window where exactness is needed (sampled with
    all residues near the multiples of d for large d), and the saturated region up to xmax."""
    rng = np.random.default_rng(d)
    m0 = (1 << 32) // d + 1
    e = m0 * d - (1 << 32)
    assert 0 < e <= d
    edge = ((1 << 32) - 1) // e                 # the largest |x| with x e < 2^32
    bounds = [129 * d + 1, 130 * d, 130 * d + d // 2 + 1, 1000 * d + 7, 128 * 127 * 64 + 4000, (1 << 29) - 1, edge, edge + 1]
    accepted = 0
    for xmax in bounds:
        if xmax >= (1 << 29):
            continue
        st, m = _lib.debug_magic32(d, xmax)
        if st != 0:
            assert st == _lib.ERR_UNSUPPORTED
            continue
        accepted += 1
        assert m == (1 << 32) // d + 1
        win = min(xmax, 131 * d + 5)
        if 2 * win + 1 <= 3_000_000:
            xs = np.arange(-win, win + 1)
        else:       # large d: every multiple of d in the window with its neighbours, plus random values
            k = np.arange(-131, 132)[:, None] * d + np.array([-2, -1, 0, 1, 2, d - 2, d - 1, d // 2])[None]
            xs = np.concatenate([k.ravel(), rng.integers(-win, win + 1, 200000)])
            xs = xs[np.abs(xs) <= win]
        far = np.concatenate([rng.integers(-xmax, xmax + 1, 200000), np.array([xmax, -xmax, xmax - 1, -xmax + 1])])
        xs = np.concatenate([xs, far])
        got = _clip8(((xs * int(m)) >> 32) + (xs < 0))
        assert np.array_equal(got, _clip8(golden_np.tdiv(xs, d))), (d, xmax)
        assert xmax <= edge or 130 * d <= edge, "accepted although exactness is not proven on the window"
    assert accepted >= 1 or edge >= (1 << 29)
    if edge + 1 < min(130 * d, 1 << 29):         # just beyond the proof: must be refused
        assert _lib.debug_magic32(d, edge + 1)[0] == _lib.ERR_UNSUPPORTED


def test_fast1_rejects_what_it_cannot_prove():
    """e = m d - 2^32 close to d and a wide accumulator range: the host must fall back to the shifted form."""
    assert _lib.debug_magic32(3, 100)[0] == _lib.ERR_UNSUPPORTED                # d < 4
    # d = 2^16 + 1: m = 65536, e = m d - 2^32 = 65536; exact only while |x| e < 2^32, i.e. |x| < 65536 << 130 d
    st, m = _lib.debug_magic32(65537, 10_000_000)
    assert st == _lib.ERR_UNSUPPORTED
    st, m = _lib.debug_magic32(65537, 65535)
    assert st == 0 and m == 65536
    # boundary of the strict inequality: x e == 2^32 must be refused.  d = 2^26: m = 65, e = 2^26; x = 64 gives x e == 2^32
    assert _lib.debug_magic32(1 << 26, 63)[0] == 0
    assert _lib.debug_magic32(1 << 26, 64)[0] == _lib.ERR_UNSUPPORTED


def _pipe_blob(arrays, C, T, **kw):
    blob, why = _lib.debug_pack_pipe(arrays, C, T, **kw)
    return (None if blob is None else np.frombuffer(blob, np.uint8)), why


# byte offsets inside PipeNet (mibmi_common.cuh): w1blk 4*2*32 words, off1 / m1 / sh1 16 ints each, fast1 + 3 pad, w1umma
_OFF_OFF1 = 4 * 2 * 32 * 4
_OFF_W1UMMA = _OFF_OFF1 + 3 * 64 + 16
_OFF_W2UMMA = _OFF_W1UMMA + 16 * 64
_OFF_C2 = _OFF_W2UMMA + 16 * 1536 + 16 * 24 * 32 * 4          # thr2, sig2, soff2, m2, sh2: 16 ints each


def test_fused_packing_tcgen05_operands(golden):
    """The operands the tcgen05 kernels read (PipeNet::w1umma / w2umma) against the exported arrays: layer 1 = W1 with the
    factor sign folded in, as K-major 8 x 16-byte core matrices; layer 2 = the Toeplitz matrix B[n][kk] = tap (kk - n - 1)."""
    fx = golden
    C, T = fx["C"], fx["T"]
    blob, why = _pipe_blob(fx["arrays"], C, T)
    if blob is None:
        assert "instantiation" in why, why
        return
    CA, kpad = convert.align4(C), (C + 31) // 32 * 32
    w1 = np.asarray(fx["arrays"]["l1_weight_align"]).reshape(16, CA).astype(np.int64)
    sg = np.where(np.asarray(fx["arrays"]["l1_factor"]) < 0, -1, 1)
    u1 = blob[_OFF_W1UMMA:_OFF_W1UMMA + 16 * 64].view(np.int8)
    for n in range(16):
        for c in range(min(C, 64)):
            assert u1[(n >> 3) * (kpad // 16) * 128 + (c >> 4) * 128 + (n & 7) * 16 + (c & 15)] == sg[n] * w1[n, c], (n, c)
    w2 = np.asarray(fx["arrays"]["l2_weight_reverse"]).reshape(16, 64)
    u2 = blob[_OFF_W2UMMA:_OFF_W2UMMA + 16 * 1536].view(np.int8).reshape(16, 1536)
    n, kk = np.meshgrid(np.arange(16), np.arange(96), indexing="ij")
    tap = kk - n - 1
    for k in range(16):
        exp = np.where((tap >= 0) & (tap < 64), w2[k][np.clip(tap, 0, 63)], 0)
        got = u2[k][(n >> 3) * 768 + (kk >> 4) * 128 + (n & 7) * 16 + (kk & 15)]
        assert np.array_equal(got, exp), k
    off1 = blob[_OFF_OFF1:_OFF_OFF1 + 64].view(np.int32)
    assert np.array_equal(off1, sg * np.asarray(fx["arrays"]["l1_offset"]))


_OFF_M3 = _OFF_C2 + 5 * 64 + 16 * 4 * 32 * 4                       # after thr2 .. sh2 and w3frag: m3, sh3


def test_layer3_magic_is_exact_after_the_clip():
    """PipeNet::m3 / sh3: the no-shift form (sh3 == 0, m3 = floor(2^32 / d) + 1) where it is proven, the shifted one otherwise; either
    way clip8(tdiv_magic(x)) == clip8(trunc(x / d)) for every layer-3 accumulator the weights allow (brute force over the range)."""
    import dataclasses
    C, T = 64, 480
    p0 = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=3), C=C, T=T)
    seen = set()
    for fac3 in (4, 5, 7, 64, 255, 257, 1000, 4099, 65535, 65537, 1 << 20, -9, -300):
        arrays = dict(p0.to_wolf_arrays())
        arrays["l3_factor"] = np.int32(fac3)
        blob, why = _pipe_blob(arrays, C, T)
        assert blob is not None, why
        m3, sh3 = (int(v) for v in blob[_OFF_M3:_OFF_M3 + 8].view(np.int32))
        d = abs(fac3)
        xmax = int(128 * np.abs(np.asarray(arrays["l3_weight"]).reshape(16, 16).astype(np.int64)).sum(axis=1).max())
        x = np.arange(-xmax, xmax + 1, dtype=np.int64)
        q = ((x * m3) >> 32 >> sh3) + (x < 0)
        ref = np.sign(x) * (np.abs(x) // d)
        assert np.array_equal(np.clip(q, -128, 127), np.clip(ref, -128, 127)), fac3
        if sh3 == 0:
            assert m3 == (1 << 32) // d + 1
        seen.add(sh3 == 0)
    assert seen == {True, False}          # both forms occur over these factors


@pytest.mark.parametrize("C", [38, 24, 19, 8])
def test_fused_packing_superrow_weights(C):
    """PipeNet::w1x, the block-structured layer-1 weights of the Wolf-order super-row form (rows of 40 / 24 / 20 / 8 bytes):
    W1x[16 p + f][C_ALIGN p + c] = sign(factor f) W1[f][c] for the S = 16 / gcd(C_ALIGN, 16) samples p of a super-row, zero
    elsewhere, as K-major 8 x 16-byte core matrices over KP = ceil(S C_ALIGN / 32) * 32 bytes."""
    import math
    T = 480
    arrays = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=5), C=C, T=T).to_wolf_arrays()
    blob, why = _pipe_blob(arrays, C, T)
    assert blob is not None, why
    CA = convert.align4(C)
    S = 16 // math.gcd(CA, 16)
    KP = (S * CA + 31) // 32 * 32
    w1x = blob[len(blob) - 64 * 96:].view(np.int8)                     # last member of PipeNet
    w1 = np.asarray(arrays["l1_weight_align"]).reshape(16, CA).astype(np.int64)
    sg = np.where(np.asarray(arrays["l1_factor"]) < 0, -1, 1)
    exp = np.zeros((16 * S, KP), np.int64)
    for p_ in range(S):
        exp[16 * p_:16 * p_ + 16, CA * p_:CA * p_ + C] = sg[:, None] * w1[:, :C]
    n, k = np.meshgrid(np.arange(16 * S), np.arange(KP), indexing="ij")
    got = w1x[(n >> 3) * (KP // 16) * 128 + (k >> 4) * 128 + (n & 7) * 16 + (k & 15)]
    assert np.array_equal(got, exp)


def test_fused_packing_plain_variant_constants():
    """REORDER_BN off: the pooled layers' constants describe fac >> 3 and off >> 3 (layer2.c:139-210): sign folded into sig / soff,
    magic pair exact for |fac >> 3| on the accumulator range; |fac >> 3| < 4 is refused (generic kernel)."""
    fx = load_golden(golden_files()[0])
    blob, why = _pipe_blob(fx["arrays"], fx["C"], fx["T"], reorder_bn=False)
    assert blob is not None, why
    assert blob[_OFF_OFF1 + 3 * 64:_OFF_OFF1 + 3 * 64 + 8].view(np.int32).tolist()[1] == 1          # PipeNet::plain
    c2 = blob[_OFF_C2:_OFF_C2 + 5 * 64].view(np.int32).reshape(5, 16).astype(np.int64)             # thr2, sig2, soff2, m2, sh2
    fac, off = np.asarray(fx["arrays"]["l2_factor"], np.int64) >> 3, np.asarray(fx["arrays"]["l2_offset"], np.int64) >> 3
    assert np.array_equal(c2[1], np.where(fac < 0, -1, 1)) and np.array_equal(c2[2], np.where(fac < 0, -off, off))
    rng = np.random.default_rng(0)
    for k in range(16):
        x = np.concatenate([rng.integers(-2 ** 28, 2 ** 28, 2000), np.arange(-300, 300) * abs(fac[k]), [0, 1, -1]]).astype(np.int64)
        q = ((x * c2[3][k]) >> 32 >> c2[4][k]) + (x < 0)
        assert np.array_equal(q, np.sign(x) * (np.abs(x) // abs(fac[k]))), k
    a = dict(fx["arrays"])
    a["l2_factor"] = a["l2_factor"].copy()
    a["l2_factor"][3] = 31                                             # 31 >> 3 == 3
    blob, why = _pipe_blob(a, fx["C"], fx["T"], reorder_bn=False)
    assert blob is None and ">> 3" in why
    assert _pipe_blob(a, fx["C"], fx["T"])[0] is not None              # fine with REORDER_BN


def test_fused_packing_refuses_what_it_cannot_run():
    fx = load_golden(golden_files()[0])
    a = dict(fx["arrays"])
    a["l1_factor"] = a["l1_factor"].copy()
    a["l1_factor"][0] = 3
    blob, why = _pipe_blob(a, fx["C"], fx["T"])
    assert blob is None and "l1_factor" in why
    a = dict(fx["arrays"])
    a["l1_offset"] = a["l1_offset"].copy()
    a["l1_offset"][5] = 2 ** 30
    blob, why = _pipe_blob(a, fx["C"], fx["T"])
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
