"""GPU parity tests (run on the B200 box: pytest -m gpu).  Every call goes through the C-ABI
(libmibmi.so); the oracle (oracle/golden_np.py, oracle/wolf_port.c) and the committed reference
outputs (tests/golden/) are the checkers.  Bar: bit-exact, every logit."""
import numpy as np
import pytest
import torch

import golden_np
import wolf_port
from conftest import golden_files, load_golden
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net

pytestmark = pytest.mark.gpu

FUSED_SHAPES = [(64, 480), (38, 480), (24, 480), (19, 480), (16, 480), (8, 480), (22, 1125)]


def _cuda(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def _net(C, T, seed, **kw):
    sd = synth.make_state_dict(C=C, T=T, seed=seed)
    p = convert.params_from_state_dict(sd, C=C, T=T)
    return Int8Net(p, device="cuda:0", **kw), p.to_wolf_arrays()


# ---- against the reference's own outputs ---------------------------------------------------------------
@pytest.mark.parametrize("path", ["generic", "fused"])
def test_reference_fixtures(golden, path):
    fx = golden
    net = Int8Net(fx["arrays"], C=fx["C"], T=fx["T"], device="cuda:0")
    ok, why = net.fused_supported()
    if path == "fused" and not ok:
        assert (fx["C"], fx["T"]) not in FUSED_SHAPES, why
        pytest.skip(why)
    net.set_path(path)
    y = net.forward(_cuda(fx["x"])).cpu().numpy()
    assert net.last_path == path
    assert np.array_equal(y, fx["golden_reorder"])      # Python golden model (golden_model.py:104)
    assert np.array_equal(y, fx["c_default"])           # C model, Makefile-default flags (model.c:42)


def test_reference_fixtures_taps_and_plain_variant(golden):
    fx = golden
    x = _cuda(fx["x"])
    net = Int8Net(fx["arrays"], C=fx["C"], T=fx["T"], device="cuda:0")
    y, taps = net.forward_taps(x)
    assert np.array_equal(y.cpu().numpy(), fx["golden_reorder"])
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name].cpu().numpy(), fx[f"golden_reorder_{name}"]), name
    plain = Int8Net(fx["arrays"], C=fx["C"], T=fx["T"], device="cuda:0", reorder_bn=False)
    yp, tp = plain.forward_taps(x)                            # AUTO with taps: the exact generic kernel
    assert plain.last_path == "generic"
    assert np.array_equal(yp.cpu().numpy(), fx["c_plain"])   # follows the C (layer4.c:115-117)
    for name in ("y1", "y2", "y3"):
        assert np.array_equal(tp[name].cpu().numpy(), fx[f"golden_plain_{name}"]), name


def test_plain_variant_on_the_fused_kernels(golden):
    """REORDER_BN off (layer2.c:139-210, layer4.c:99-140: every accumulator divided, ReLU, sum of 8, >> 3) on the tcgen05 kernels:
    logits equal to the reference C built without -DREORDER_BN, per-layer taps of the fused kernel equal to the oracle's;
    Wolf order and the forced pipeline kernel fall back / refuse."""
    fx = golden
    C, T = fx["C"], fx["T"]
    x = _cuda(fx["x"])
    plain = Int8Net(fx["arrays"], C=C, T=T, device="cuda:0", reorder_bn=False)
    ok, why = plain.fused_supported()
    assert ok, why
    y = plain.forward(x).cpu().numpy()
    assert plain.last_path == "fused" and plain.last_kernel in ("tc_forward_kernel", "tcr_forward_kernel")
    assert np.array_equal(y, fx["c_plain"])
    ref, lay = wolf_port.forward(fx["arrays"], C, T, fx["x"], reorder_bn=False, return_layers=True)
    assert np.array_equal(ref, fx["c_plain"])
    plain.set_path("fused")
    yt, taps = plain.forward_taps(x)
    assert np.array_equal(yt.cpu().numpy(), ref)
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name].cpu().numpy(), lay[name]), name
    plain.set_path("auto")
    xw = _cuda(convert.to_wolf_input(fx["x"]))
    assert np.array_equal(plain.forward(xw, layout="btc").cpu().numpy(), ref) and plain.last_path == "generic"
    with pytest.raises(Exception):
        plain.set_path("fused_pipe")


@pytest.mark.parametrize("C,T,B", [(64, 480, 65536), (22, 1125, 20000), (38, 480, 4099), (8, 480, 777)])
def test_plain_variant_fused_vs_oracle_random(C, T, B):
    arrays = _net(C, T, seed=41)[1]
    net = Int8Net(arrays, C=C, T=T, device="cuda:0", reorder_bn=False)
    x = synth.make_trials(B, C, T, seed=6)
    y = net.forward(_cuda(x)).cpu().numpy()
    assert net.last_path == "fused", net.fused_supported()
    ref = wolf_port.forward(arrays, C, T, x, reorder_bn=False)
    bad = np.nonzero((y != ref).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} trials differ, first {bad[:5]}"
    assert np.array_equal(ref[:300], golden_np.forward(arrays, C, T, x[:300], reorder_bn=False))


def test_reference_fixtures_taps_from_the_fused_kernel(golden):
    """The pipeline kernel's diagnostic instantiation stores y1..y4 from its registers: every layer
    of the fused path is pinned to the reference's per-layer outputs, not only the logits."""
    fx = golden
    net = Int8Net(fx["arrays"], C=fx["C"], T=fx["T"], device="cuda:0")
    if not net.fused_supported()[0]:
        pytest.skip(net.fused_supported()[1])
    net.set_path("fused")
    y, taps = net.forward_taps(_cuda(fx["x"]))
    assert net.last_path == "fused"
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name].cpu().numpy(), fx[f"golden_reorder_{name}"]), name
    assert np.array_equal(y.cpu().numpy(), fx["golden_reorder"])


def test_wolf_input_layout(golden):
    fx = golden
    net = Int8Net(fx["arrays"], C=fx["C"], T=fx["T"], device="cuda:0")
    xw = convert.to_wolf_input(fx["x"])   # [B][T][C_ALIGN]  (gen_input_header.py:75-76)
    y = net.forward(_cuda(xw), layout="btc").cpu().numpy()
    assert np.array_equal(y, fx["golden_reorder"])
    # rows of C_ALIGN bytes that are a multiple of 16 go through the tcgen05 kernel (K-major operand, 3-D tensor map), 22 x 1125
    # through the realign warps of tcr_forward_kernel; the others are transposed into a scratch chunk by chunk
    # (mibmi_reorder.cu) and run on the fused kernel of the shape
    ok, _why = net.fused_supported()
    assert net.last_path == ("fused" if ok else "generic")
    if ok:
        assert net.last_kernel == ("tcr_forward_kernel" if fx["T"] == 1125 else "tc_forward_kernel")
    if fx["C"] in (64, 16) and fx["T"] == 480:
        xr = synth.make_trials(1237, fx["C"], fx["T"], seed=9)
        ref = wolf_port.forward(fx["arrays"], fx["C"], fx["T"], xr)
        assert np.array_equal(net.forward(_cuda(convert.to_wolf_input(xr)), layout="btc").cpu().numpy(), ref)
        net.set_path("fused")                       # the diagnostic instantiation of the same kernel
        _, taps = net.forward_taps(_cuda(xw), layout="btc")
        assert net.last_kernel == "tc_forward_kernel"
        for name in ("y1", "y2", "y3", "y4"):
            assert np.array_equal(taps[name].cpu().numpy(), fx[f"golden_reorder_{name}"]), name


@pytest.mark.parametrize("C,T", [(38, 480), (24, 480), (19, 480), (8, 480)])
@pytest.mark.parametrize("chunk", [None, 101])
def test_wolf_order_superrows_inside_the_kernel(C, T, chunk, monkeypatch):
    """Wolf-order rows of 40 / 24 / 20 / 8 bytes at T = 480 (the layout net_model_compute receives, gen_input_header.py:75-76): ONE
    launch of tc_forward_kernel -- 2 or 4 samples form a super-row of lcm(C_ALIGN, 16) bytes that a tensor map can describe, the
    layer-1 weights are block structured (PipeNet::w1x), an MMA row yields 2 or 4 samples x 16 filters.  No scratch, no second pass
    over HBM.  Sub-batches, taps of the fused kernel, the host entry point; the mma.sync pipeline kernel (forced) still goes through
    the reorder scratch, chunk by chunk."""
    if chunk:
        monkeypatch.setenv("MIBMI_WOLF_CHUNK", str(chunk))
    net, arrays = _net(C, T, 31)
    B = 1237
    x = synth.make_trials(B, C, T, seed=77)
    ref, lay = wolf_port.forward(arrays, C, T, x, return_layers=True)
    xw = _cuda(convert.to_wolf_input(x))
    n0 = net.launch_count
    y = net.forward(xw, layout="btc").cpu().numpy()
    bad = np.nonzero((y != ref).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} trials differ, first {bad[:5]}"
    assert net.last_path == "fused" and net.last_kernel == "tc_forward_kernel" and net.launch_count - n0 == 1
    assert np.array_equal(net.forward(xw[:1], layout="btc").cpu().numpy(), ref[:1])
    assert np.array_equal(net.forward(xw[4:7], layout="btc").cpu().numpy(), ref[4:7])
    assert np.array_equal(net.forward(xw[1000:], layout="btc").cpu().numpy(), ref[1000:])
    net.set_path("fused")
    y, taps = net.forward_taps(xw[:300], layout="btc")
    assert np.array_equal(y.cpu().numpy(), ref[:300])
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name].cpu().numpy(), lay[name][:300]), name
    # host entry point (two staging streams)
    net.set_path("auto")
    assert np.array_equal(net.forward_host(convert.to_wolf_input(x), layout="btc"), ref)
    # the pipeline kernel has no Wolf-order loader: reorder into the scratch, then the kernel, per chunk
    net.set_path("fused_pipe")
    n0 = net.launch_count
    assert np.array_equal(net.forward(xw, layout="btc").cpu().numpy(), ref)
    assert net.last_kernel == "pipe_forward_kernel"
    if chunk:
        assert net.launch_count - n0 == 2 * -(-B // chunk)


def test_wolf_order_22x1125_inside_the_kernel():
    """BASELINE config 3 in the order net_model_compute receives it ([T][24] rows, gen_input_header.py:75-76): ONE launch of
    tcr_forward_kernel, the 24-byte rows are expanded to the K-major layer-1 operand in shared memory (no scratch, no second pass
    over HBM).  A Wolf-order trial is 27 000 bytes: trials alternate between 0 and 8 modulo 16, any sub-batch start is legal when
    16-byte aligned (even trial index)."""
    C, T = 22, 1125
    net, arrays = _net(C, T, 31)
    B = 1237
    x = synth.make_trials(B, C, T, seed=77)
    ref, lay = wolf_port.forward(arrays, C, T, x, return_layers=True)
    xw = _cuda(convert.to_wolf_input(x))
    n0 = net.launch_count
    y = net.forward(xw, layout="btc").cpu().numpy()
    bad = np.nonzero((y != ref).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} trials differ, first {bad[:5]}"
    assert net.last_path == "fused" and net.last_kernel == "tcr_forward_kernel" and net.launch_count - n0 == 1
    for lo, hi in ((0, 1), (4, 7), (2, 445), (1000, 1237)):
        assert np.array_equal(net.forward(xw[lo:hi], layout="btc").cpu().numpy(), ref[lo:hi]), (lo, hi)
    # odd start: 8 modulo 16 -> the exact generic kernel under AUTO
    assert np.array_equal(net.forward(xw[1:40], layout="btc").cpu().numpy(), ref[1:40])
    assert net.last_path == "generic"
    net.set_path("fused")
    yt, taps = net.forward_taps(xw[:300], layout="btc")
    assert net.last_kernel == "tcr_forward_kernel"
    assert np.array_equal(yt.cpu().numpy(), ref[:300])
    for name in ("y1", "y2", "y3", "y4"):
        assert np.array_equal(taps[name].cpu().numpy(), lay[name][:300]), name
    net.set_path("auto")
    assert np.array_equal(net.forward_host(convert.to_wolf_input(x), layout="btc"), ref)
    # padding bytes of the Wolf rows (channels 22, 23) are never trusted: garbage there must not change anything
    xg = convert.to_wolf_input(x[:64]).copy()
    xg[:, :, 22:] = np.random.default_rng(0).integers(-128, 128, xg[:, :, 22:].shape)
    assert np.array_equal(net.forward(_cuda(xg), layout="btc").cpu().numpy(), ref[:64])
    # saturating inputs
    xs = np.stack([np.full((C, T), v, np.int8) for v in (127, -127, -128, 0)])
    assert np.array_equal(net.forward(_cuda(convert.to_wolf_input(xs)), layout="btc").cpu().numpy(), wolf_port.forward(arrays, C, T, xs))


# ---- against the oracle on seeded random nets / trials ------------------------------------------------
@pytest.mark.parametrize("C,T", FUSED_SHAPES)
@pytest.mark.parametrize("seed", [21, 22, 23])
def test_fused_vs_oracle_random(C, T, seed):
    net, arrays = _net(C, T, seed)
    ok, why = net.fused_supported()
    assert ok, why
    B = 777   # not a multiple of anything: ragged tail over the persistent grid
    x = synth.make_trials(B, C, T, seed=seed, dist="gauss" if seed % 2 else "uniform")
    y = net.forward(_cuda(x)).cpu().numpy()
    assert net.last_path == "fused"
    ref = wolf_port.forward(arrays, C, T, x)
    bad = np.nonzero((y != ref).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} trials differ, first {bad[:5]}: {y[bad[:2]]} vs {ref[bad[:2]]}"
    net.set_path("generic")
    assert np.array_equal(net.forward(_cuda(x)).cpu().numpy(), ref)


@pytest.mark.parametrize("C,T", [(22, 1125), (64, 480), (5, 64), (64, 1024), (3, 200)])
def test_generic_vs_oracle_random(C, T):
    net, arrays = _net(C, T, seed=31)
    net.set_path("generic")
    x = synth.make_trials(65, C, T, seed=5)
    for rb in (True, False):
        n = Int8Net(arrays, C=C, T=T, device="cuda:0", reorder_bn=rb)
        n.set_path("generic")
        y = n.forward(_cuda(x)).cpu().numpy()
        assert np.array_equal(y, golden_np.forward(arrays, C, T, x, reorder_bn=rb)), rb
        assert np.array_equal(y, wolf_port.forward(arrays, C, T, x, reorder_bn=rb)), rb


def test_edge_batches_and_extreme_inputs():
    C, T = 64, 480
    net, arrays = _net(C, T, seed=41)
    for B in (1, 2, 3, 295, 296, 297, 593):    # around 1x / 2x the persistent grid (148 SMs x 2 CTAs)
        x = synth.make_trials(B, C, T, seed=B)
        assert np.array_equal(net.forward(_cuda(x)).cpu().numpy(), wolf_port.forward(arrays, C, T, x)), B
    # saturating inputs: all +127, all -127, all -128 (outside quantize_to_int's range but valid int8), zeros
    x = np.stack([np.full((C, T), v, np.int8) for v in (127, -127, -128, 0)])
    x = np.concatenate([x, np.where(np.arange(T)[None, :] % 2 == 0, 127, -128).astype(np.int8)[None].repeat(C, 1)
                        .reshape(1, C, T)])
    assert np.array_equal(net.forward(_cuda(x)).cpu().numpy(), wolf_port.forward(arrays, C, T, x))
    # empty batch
    y = net.forward(torch.empty((0, C, T), dtype=torch.int8, device="cuda:0"))
    assert y.shape == (0, 4)


@pytest.mark.parametrize("C,T,layout", [(64, 480, "bct"), (22, 1125, "bct"), (22, 1125, "btc"), (38, 480, "btc"), (5, 64, "bct")])
def test_forward_captured_in_a_cuda_graph(C, T, layout):
    """mibmi_forward_int8 under stream capture: the call only encodes a tensor map on the host and launches, so it can be
    captured once and replayed (a serving loop's small batches); every replay reads the buffers' current contents."""
    net, arrays = _net(C, T, seed=45)
    B = 37
    xs = [synth.make_trials(B, C, T, seed=450 + i, dist="gauss" if i else "uniform") for i in range(3)]
    wolf = lambda x: convert.to_wolf_input(x) if layout == "btc" else x
    xd = _cuda(wolf(xs[0])).clone()
    out = torch.zeros((B, 4), dtype=torch.int8, device="cuda:0")
    net.forward(xd, layout=layout, out=out)          # first call outside the capture (function attributes, module load)
    torch.cuda.synchronize()
    kernel = net.last_kernel
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        net.forward(xd, layout=layout, out=out)
    assert net.last_kernel == kernel
    for x in xs:
        xd.copy_(_cuda(wolf(x)))
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), wolf_port.forward(arrays, C, T, x))


def test_stream_forward_captured_in_a_cuda_graph():
    C, T, S, L = 64, 480, 3, 480 + 16 * 50
    net, arrays, x0 = _stream_case(C, T, S, L, 46)
    xd = _cuda(x0).clone()
    out = net.forward_stream(xd)
    torch.cuda.synchronize()
    assert net.last_path == "stream"
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        net.forward_stream(xd, out=out)
    for seed in (47, 48):
        x = synth.make_trials(S, C, L, seed=seed)
        xd.copy_(_cuda(x))
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), wolf_port.forward_stream(arrays, C, T, x))


def test_unaligned_rows_22x1125_every_misalignment():
    """BASELINE config 3 (BCI-IV-2a): rows of 1125 bytes are never 16-byte aligned; the byte shift of a
    row depends on (trial, channel).  A 16-byte-aligned sub-batch view may start at any multiple of 8 trials."""
    C, T = 22, 1125
    net, arrays = _net(C, T, seed=51)
    assert net.fused_supported()[0], net.fused_supported()[1]
    B = 1237
    x = synth.make_trials(B, C, T, seed=52)
    ref = wolf_port.forward(arrays, C, T, x)
    xd = _cuda(x)
    y = net.forward(xd).cpu().numpy()
    assert net.last_path == "fused"
    assert np.array_equal(y, ref)
    for lo in (8, 16, 1000):                      # views: lo * 24750 is a multiple of 16
        assert np.array_equal(net.forward(xd[lo:]).cpu().numpy(), ref[lo:]), lo
    assert net.last_kernel == "tcr_forward_kernel"     # tcgen05 kernel with realigning warps (mibmi_tcr.cu)
    for Bs in (1, 2, 3, 4, 7, 149, 443, 444, 445, 889):   # around 1x / 2x the persistent grid of 148 CTAs x groups of 3 trials
        assert np.array_equal(net.forward(xd[:Bs].clone()).cpu().numpy(), ref[:Bs]), Bs
    net.set_path("fused_pipe")                          # the mma.sync predecessor stays selectable and exact
    assert np.array_equal(net.forward(xd).cpu().numpy(), ref)
    assert net.last_kernel == "pipe_forward_kernel"
    net.set_path("auto")
    # saturating inputs
    xs = np.stack([np.full((C, T), v, np.int8) for v in (127, -127, -128, 0)])
    assert np.array_equal(net.forward(_cuda(xs)).cpu().numpy(), wolf_port.forward(arrays, C, T, xs))


def test_full_size_batch_22x1125_every_logit():
    """BASELINE config 3 at 65536 trials: every logit against the CPU integer path + permutation property."""
    C, T, B = 22, 1125, 65536
    net, arrays = _net(C, T, seed=2)
    x = synth.make_trials(B, C, T, seed=77)
    y = net.forward(_cuda(x)).cpu().numpy()
    assert net.last_path == "fused"
    ref = wolf_port.forward(arrays, C, T, x)
    assert np.array_equal(y, ref)
    perm = np.random.default_rng(2).permutation(B)
    assert np.array_equal(net.forward(_cuda(x[perm])).cpu().numpy(), y[perm])


@pytest.mark.parametrize("C,T", [(38, 480), (19, 480), (22, 1125)])
def test_full_size_batch_wolf_order_every_logit(C, T):
    """The order net_model_compute receives ([B][T][C_ALIGN]) at 65536 trials, every logit against the CPU integer path and against
    the module-order result of the same kernel family: super-rows of 2 samples (40-byte rows), of 4 samples (20-byte rows), and the
    24-byte rows of 22 x 1125 expanded in shared memory."""
    B = 65536
    net, arrays = _net(C, T, seed=4)
    x = synth.make_trials(B, C, T, seed=55)
    xw = _cuda(convert.to_wolf_input(x))
    n0 = net.launch_count
    y = net.forward(xw, layout="btc").cpu().numpy()
    assert net.last_path == "fused" and net.launch_count - n0 == 1
    ref = wolf_port.forward(arrays, C, T, x)
    bad = np.nonzero((y != ref).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} trials differ, first {bad[:5]}"
    del xw
    assert np.array_equal(net.forward(_cuda(x)).cpu().numpy(), y)


@pytest.mark.parametrize("C,T,layout", [(22, 1125, "bct"), (8, 480, "btc")])
def test_batch_beyond_one_launch(C, T, layout):
    """The upper end of BASELINE config 3's batch sweep and past it: 2^20 + 37 trials in ONE call, which the library splits into two
    launches (32-bit tensor-map coordinates / kMaxLaunchTrials).  The head, the trials on both sides of the split and the tail are
    compared with the CPU integer path; the whole result with the same trials run in smaller calls (determinism and per-trial
    independence at full size)."""
    free, _ = torch.cuda.mem_get_info()
    B = (1 << 20) + 37
    CA = C if layout == "bct" else (C + 3) // 4 * 4
    if free < 2.5 * B * CA * T:
        pytest.skip("not enough device memory for a 2^20-trial batch")
    net, arrays = _net(C, T, seed=9)
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device="cuda", generator=g)
    if layout == "btc":
        xw = torch.zeros((B, T, CA), dtype=torch.int8, device="cuda")
        xw[:, :, :C] = x.transpose(1, 2)
    else:
        xw = x
    n0 = net.launch_count
    y = net.forward(xw, layout=layout)
    assert net.last_path == "fused" and net.launch_count - n0 == 2
    for lo, hi in ((0, 2048), ((1 << 20) - 1024, (1 << 20) + 37)):
        ref = wolf_port.forward(arrays, C, T, x[lo:hi].cpu().numpy())
        assert np.array_equal(y[lo:hi].cpu().numpy(), ref), (lo, hi)
    step = 1 << 18                          # a multiple of 16 trials: every piece is 16-byte aligned again
    y2 = torch.cat([net.forward(xw[i:i + step], layout=layout) for i in range(0, B, step)])
    assert torch.equal(y, y2)


def test_full_size_batch_65536_every_logit():
    """BASELINE config 2: 64 x 480, B = 65536, every logit compared with the CPU integer path."""
    C, T, B = 64, 480, 65536
    net, arrays = _net(C, T, seed=0)
    x = synth.make_trials(B, C, T, seed=99)
    y = net.forward(_cuda(x)).cpu().numpy()
    assert net.last_path == "fused"
    ref = wolf_port.forward(arrays, C, T, x)      # all host cores
    assert np.array_equal(y, ref)
    # size-independent properties: per-trial independence (any permutation of the batch permutes
    # the logits), determinism, and a checksum of the logits against the oracle's
    perm = np.random.default_rng(1).permutation(B)
    y2 = net.forward(_cuda(x[perm])).cpu().numpy()
    assert np.array_equal(y2, y[perm])
    assert int(y.astype(np.int64).sum()) == int(ref.astype(np.int64).sum())


def test_host_entry_point_matches_device_path():
    C, T, B = 64, 480, 5000
    net, arrays = _net(C, T, seed=3)
    net.set_host_chunk(1024)
    x = synth.make_trials(B, C, T, seed=8)
    ref = wolf_port.forward(arrays, C, T, x)
    xp = torch.from_numpy(x).pin_memory()
    assert np.array_equal(net.forward_host(xp).numpy(), ref)          # pinned: direct copies
    assert np.array_equal(net.forward_host(torch.from_numpy(x)).numpy(), ref)   # pageable: staged


def test_float_pre_and_post_ops():
    C, T = 64, 480
    net, arrays = _net(C, T, seed=5)
    rng = np.random.default_rng(0)
    xf = rng.normal(0, 0.5, size=(33, C, T)).astype(np.float32)
    q = net.quantize_input(torch.from_numpy(xf).cuda())
    assert np.array_equal(q.cpu().numpy(), convert.quantize_input(xf, net.params.input_scale))   # functional.py:308-335
    y = net.forward(q)
    yf = net.dequantize_logits(y).cpu().numpy()
    ref = convert.dequantize(y.cpu().numpy(), net.params.output_scale)
    assert np.allclose(yf, ref, rtol=1e-6, atol=0)   # float32 vs the reference's float64: 1e-6 relative


@pytest.mark.parametrize("n,skip", [(1, 0), (15, 0), (16, 0), (4099, 0), (4099, 1), (65536 + 7, 3)])
def test_quantize_input_vector_body_and_tails(n, skip):
    """The 16-floats-per-thread body, its scalar tail and unaligned views give NumPy's bytes (values beyond +-s clip)."""
    net, _ = _net(64, 480, seed=5)
    rng = np.random.default_rng(n)
    xf = (rng.normal(0, 0.8, size=n + skip) * net.params.input_scale).astype(np.float32)
    xf[:: 7] *= 3.0
    xd = torch.from_numpy(xf).cuda()[skip:]                     # skip != 0: not 16-byte aligned -> scalar kernel
    q = net.quantize_input(xd)
    assert np.array_equal(q.cpu().numpy(), convert.quantize_input(xf[skip:], net.params.input_scale))


def test_errors_are_reported_not_printed():
    C, T = 64, 480
    net, arrays = _net(C, T, seed=1)
    with pytest.raises(ValueError):
        net.forward(torch.zeros((2, C, T + 1), dtype=torch.int8, device="cuda:0"))
    with pytest.raises(TypeError):
        net.forward(torch.zeros((2, C, T), dtype=torch.float32, device="cuda:0"))
    # a pointer that is not 16-byte aligned: the exact generic kernel (no alignment requirement) under AUTO, MIBMI_ERR_ALIGN
    # when a fused path is forced
    xs = synth.make_trials(2, C, T, seed=3)
    buf = torch.zeros((2 * C * T + 1,), dtype=torch.int8, device="cuda:0")
    buf[1:] = torch.from_numpy(xs).cuda().flatten()
    x = buf[1:].view(2, C, T)
    assert x.data_ptr() % 16 != 0
    assert np.array_equal(net.forward(x).cpu().numpy(), wolf_port.forward(arrays, C, T, xs))
    assert net.last_path == "generic" and net.last_kernel == "generic_forward_kernel"
    from mibminet_b200._lib import MibmiError, ERR_ALIGN
    net.set_path("fused")
    with pytest.raises(MibmiError) as ei:
        net.forward(x)
    assert ei.value.status == ERR_ALIGN


def test_sub_batch_views_of_unaligned_trials_run_exactly():
    """22 x 1125 / 19 x 480: a trial is not a multiple of 16 bytes, so x[1:] is a contiguous view at an odd address:
    it must work (generic kernel under AUTO), and 16-byte-aligned views stay on the fused kernel."""
    for C, T in ((22, 1125), (19, 480)):
        net, arrays = _net(C, T, seed=9)
        x = synth.make_trials(40, C, T, seed=10)
        ref = wolf_port.forward(arrays, C, T, x)
        xd = _cuda(x)
        for lo in (1, 3, 8, 16):
            v = xd[lo:]
            y = net.forward(v).cpu().numpy()
            assert np.array_equal(y, ref[lo:]), (C, T, lo)
            assert net.last_path == ("fused" if v.data_ptr() % 16 == 0 else "generic"), (C, T, lo)


@pytest.mark.parametrize("C,T", [(22, 1125), (38, 480), (64, 480)])
def test_host_entry_point_both_layouts_on_one_handle(C, T):
    """The staging buffers of mibmi_forward_int8_host are sized per call: module order (C * T bytes per trial) first, then Wolf
    order (T * align4(C) bytes per trial, larger for C % 4 != 0) on the same handle, and back."""
    net, arrays = _net(C, T, seed=12)
    net.set_host_chunk(256)
    B = 700
    x = synth.make_trials(B, C, T, seed=13)
    ref = wolf_port.forward(arrays, C, T, x)
    xw = convert.to_wolf_input(x)
    assert np.array_equal(net.forward_host(x, layout="bct").numpy(), ref)
    assert np.array_equal(net.forward_host(xw, layout="btc").numpy(), ref)
    assert np.array_equal(net.forward_host(x, layout="bct").numpy(), ref)
    net.set_host_chunk(512)                                   # larger chunk: reallocation by trial count
    assert np.array_equal(net.forward_host(torch.from_numpy(xw).pin_memory(), layout="btc").numpy(), ref)


@pytest.mark.parametrize("C,T", [(38, 480), (22, 1125)])
def test_wolf_order_calls_on_concurrent_streams(C, T, monkeypatch):
    """Wolf-order calls of ONE handle on different streams at the same time: scratch memory belongs to the call (allocated in
    stream order), so the results cannot mix."""
    monkeypatch.setenv("MIBMI_WOLF_CHUNK", "64")
    net, arrays = _net(C, T, seed=14)
    xs = [synth.make_trials(500, C, T, seed=20 + i) for i in range(3)]
    refs = [wolf_port.forward(arrays, C, T, x) for x in xs]
    xd = [_cuda(convert.to_wolf_input(x)) for x in xs]
    outs = [torch.empty((500, 4), dtype=torch.int8, device="cuda:0") for _ in xs]
    streams = [torch.cuda.Stream() for _ in xs]
    torch.cuda.synchronize()
    for rep in range(3):
        for st, x, o in zip(streams, xd, outs):
            with torch.cuda.stream(st):
                net.forward(x, layout="btc", out=o)
    torch.cuda.synchronize()
    for o, r in zip(outs, refs):
        assert np.array_equal(o.cpu().numpy(), r)


def _adversarial_arrays(C, T, seed):
    """Exported arrays (gen_net_header.py layout) built to stress the requantisation: layer-1 factors just above the smallest
    divisor the fused path takes and at 2^k +- 1, both signs; offsets that push |acc + off| to the bound the host proves
    (128 * sum|w| + |off|); weights with small integer steps so that the accumulator visits every residue modulo the factor."""
    rng = np.random.default_rng(seed)
    base = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=seed), C=C, T=T).to_wolf_arrays()
    a = {k: np.array(v, copy=True) for k, v in base.items()}
    CA = convert.align4(C)
    facs = np.array([4, -5, 7, -8, 9, 15, -16, 17, 31, -33, 63, 65, -127, 129, 255, -257], np.int32)
    w1 = np.zeros((16, CA), np.int8)
    for k in range(16):
        w1[k, :C] = rng.integers(-2, 3, C)                   # small steps: every residue is reached
        w1[k, rng.integers(0, C)] = 1
        if k % 3 == 0:
            w1[k, :C] = rng.integers(-127, 128, C)           # and full-range rows that saturate
    a["l1_weight_align"] = w1.ravel()
    a["l1_factor"] = facs
    off = np.array([0, 3, -2, 1000, -1000, 7, 129 * 16, -130 * 17, 12345, -54321, 1, -1, 20000, -20000, 99, -77], np.int32)
    a["l1_offset"] = off
    a["l2_factor"] = np.where(np.arange(16) % 2 == 0, 1, -1).astype(np.int32) * np.array(
        [8, 9, 15, 17, 31, 33, 63, 65, 127, 129, 255, 257, 511, 513, 4097, 65537], np.int32) * 8
    a["l3_factor"] = np.int32(-5)
    a["l4_factor"] = (np.array([4, 5, 7, 9, 15, 17, 31, 33, 63, 65, 127, 129, 255, 257, 1023, 1025], np.int32) *
                      np.where(np.arange(16) % 3 == 0, -1, 1)).astype(np.int32)
    a["l5_factor"] = np.int32(-337)
    return a


@pytest.mark.parametrize("C,T", [(64, 480), (22, 1125), (8, 480)])
def test_adversarial_requantisation_parameters(C, T):
    """Factors at 4, 5, 2^k +- 1 with both signs, large offsets, saturating and small-step weights: fused kernels == oracle."""
    for seed in (1, 2):
        arrays = _adversarial_arrays(C, T, seed)
        net = Int8Net(arrays, C=C, T=T, device="cuda:0")
        ok, why = net.fused_supported()
        assert ok, why
        x = np.concatenate([synth.make_trials(300, C, T, seed=seed), synth.make_trials(100, C, T, seed=seed + 1, dist="gauss"),
                            np.stack([np.full((C, T), v, np.int8) for v in (127, -127, -128, 0, 1, -1)])])
        ref, lay = wolf_port.forward(arrays, C, T, x, return_layers=True)
        y = net.forward(_cuda(x)).cpu().numpy()
        assert net.last_path == "fused"
        assert np.array_equal(y, ref)
        net.set_path("fused")
        _, taps = net.forward_taps(_cuda(x[:64]))
        for name in ("y1", "y2", "y3", "y4"):
            assert np.array_equal(taps[name].cpu().numpy(), lay[name][:64]), name
        assert (np.abs(lay["y1"].astype(int)) >= 127).any() and (np.abs(lay["y1"].astype(int)) < 100).any()


# ---- sliding-window streaming (BASELINE config 5) -------------------------------------------------------
def _stream_case(C, T, S, L, seed):
    net, arrays = _net(C, T, seed)
    x = synth.make_trials(S, C, L, seed=seed + 100, dist="gauss" if seed % 2 else "uniform")
    return net, arrays, x


@pytest.mark.parametrize("C,S,L,seed", [(64, 3, 480 + 16 * 70, 61), (64, 2, 480, 62), (64, 5, 480 + 16 * 31, 63),
                                         (64, 1, 480 + 16 * 32, 64), (38, 2, 480 + 16 * 40, 65), (8, 3, 480 + 16 * 45, 66),
                                         (16, 2, 480 + 16 * 33, 67), (24, 2, 480 + 16 * 64, 68), (19, 2, 480 + 16 * 37, 69)])
def test_stream_fused_vs_materialised_windows(C, S, L, seed):
    """Every window of the fused streaming kernel == the oracle on the materialised window, per layer."""
    T = 480
    net, arrays, x = _stream_case(C, T, S, L, seed)
    assert net.stream_supported()[0], net.stream_supported()[1]
    net.set_stream_mode("fused")
    ref, lay = wolf_port.forward_stream(arrays, C, T, x, return_layers=True)
    W = ref.shape[1]
    y, taps = net.forward_stream_taps(_cuda(x))
    assert net.last_path == "stream"
    y2 = lay["y2"]                                      # [S][W][16][60]
    e = taps["y2_edges"].cpu().numpy()
    assert np.array_equal(e[..., :4], y2[..., :4]), "left edge samples of layer 2"
    assert np.array_equal(e[..., 4:], y2[..., 56:]), "right edge samples of layer 2"
    s2 = taps["y2_stream"].cpu().numpy()               # [S][16][2 (W-1) + 60]
    for w in range(W):
        assert np.array_equal(s2[:, :, 2 * w + 4: 2 * w + 56], y2[:, w, :, 4:56]), f"interior of layer 2, window {w}"
    assert np.array_equal(taps["y3"].cpu().numpy(), lay["y3"]), "layer 3"
    assert np.array_equal(taps["y4"].cpu().numpy(), lay["y4"]), "layer 4"
    assert np.array_equal(y.cpu().numpy(), ref)
    assert np.array_equal(net.forward_stream(_cuda(x)).cpu().numpy(), ref)     # production instantiation


def test_stream_gather_fallback_any_hop():
    C, T = 64, 480
    net, arrays, x = _stream_case(C, T, 2, 480 + 700, 71)
    for hop in (16, 8, 5, 160):
        net.set_stream_mode("gather" if hop == 16 else "auto")
        y = net.forward_stream(_cuda(x), hop=hop).cpu().numpy()
        assert net.last_path == "fused"
        assert np.array_equal(y, wolf_port.forward_stream(arrays, C, T, x, hop=hop)), hop
    # 22 x 1125 has no fused streaming instantiation: auto falls back to materialised windows
    net2, arrays2, x2 = _stream_case(22, 1125, 2, 1125 + 16 * 9, 72)
    assert not net2.stream_supported()[0]
    assert np.array_equal(net2.forward_stream(_cuda(x2)).cpu().numpy(), wolf_port.forward_stream(arrays2, 22, 1125, x2))


def test_stream_config5_slice_every_logit():
    """BASELINE config 5 shape (64 ch, windows of 480, stride 16) on 64 subjects x 60 s at 160 Hz."""
    C, T, S, L = 64, 480, 64, 9600
    net, arrays, x = _stream_case(C, T, S, L, 73)
    y = net.forward_stream(_cuda(x)).cpu().numpy()
    assert net.last_path == "stream"
    assert y.shape == (S, (L - T) // 16 + 1, 4)
    assert np.array_equal(y, wolf_port.forward_stream(arrays, C, T, x))
    # shift property: dropping the first 16 samples of every stream drops the first window, nothing else
    y2 = net.forward_stream(_cuda(x[:, :, 16:])).cpu().numpy()
    assert np.array_equal(y2, y[:, 1:])


# ---- forward fused with the gather (multi-destination epilogue) -----------------------------------------
@pytest.mark.parametrize("C,T", [(64, 480), (22, 1125), (5, 64)])
def test_forward_gather_writes_every_destination(C, T):
    """mibmi_forward_int8_gather on one GPU: three destination buffers stand in for three ranks' symmetric
    buffers; rows outside [row0, row0 + B) stay untouched.  (5, 64) has no fused kernel: peer copies."""
    net, arrays = _net(C, T, seed=81)
    B, row0, rows = 333, 100, 600
    x = synth.make_trials(B, C, T, seed=82)
    ref = wolf_port.forward(arrays, C, T, x)
    bufs = [torch.full((rows, 4), 77, dtype=torch.int8, device="cuda:0") for _ in range(3)]
    net.forward_gather(_cuda(x), [b.data_ptr() for b in bufs], row0)
    for b in bufs:
        h = b.cpu().numpy()
        assert np.array_equal(h[row0:row0 + B], ref)
        assert (h[:row0] == 77).all() and (h[row0 + B:] == 77).all()


@pytest.mark.parametrize("C,T", [(22, 1125), (38, 480), (64, 480)])
def test_forward_gather_wolf_order(C, T, monkeypatch):
    """The fused gather with Wolf-order input: the chunked reorder path moves the peer pointers along with the chunks."""
    monkeypatch.setenv("MIBMI_WOLF_CHUNK", "150")
    net, arrays = _net(C, T, seed=83)
    B, row0, rows = 401, 52, 500                 # row0 * 4 is a multiple of 16: 16-byte peer stores; 3 chunks of 150 trials
    x = synth.make_trials(B, C, T, seed=84)
    ref = wolf_port.forward(arrays, C, T, x)
    bufs = [torch.full((rows, 4), 77, dtype=torch.int8, device="cuda:0") for _ in range(3)]
    net.forward_gather(_cuda(convert.to_wolf_input(x)), [b.data_ptr() for b in bufs], row0, layout="btc")
    assert net.last_path == "fused"
    for b in bufs:
        h = b.cpu().numpy()
        assert np.array_equal(h[row0:row0 + B], ref)
        assert (h[:row0] == 77).all() and (h[row0 + B:] == 77).all()


def test_torch_op_matches_the_reference_fixtures(golden):
    """torch.ops.mibminet.forward_int8 on the reference fixtures, both reorder_bn modes."""
    from mibminet_b200 import torch_op
    fx = golden
    dims = [fx["C"], fx["T"], 16, 1, 4]
    blob = torch_op.pack_blob(fx["arrays"], dims)
    y = torch.ops.mibminet.forward_int8(_cuda(fx["x"]), blob, dims, True)
    assert y.is_cuda and np.array_equal(y.cpu().numpy(), fx["golden_reorder"])
    y = torch.ops.mibminet.forward_int8(_cuda(fx["x"])[:, None], blob, dims, True)      # x[B, 1, C, T] as the module receives it
    assert np.array_equal(y.cpu().numpy(), fx["golden_reorder"])
