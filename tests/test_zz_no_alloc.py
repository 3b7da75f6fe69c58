"""mibmi_forward_stream_int8_noalloc (include/mibmi.h): the any-hop windowed scoring with its window scratch in a file-scope
`static __device__` array instead of a stream-ordered allocation -- an entry point that calls no cudaMalloc / cudaFree, for
callers under CUDA graph capture or in harnesses that forbid allocation.  No counterpart in the reference.
Called through the C-ABI (ctypes) and compared, bit for bit, with the existing entry point
mibmi_forward_stream_int8 on two calls in a row (the second smaller than the first: the shared array still holds the first
call's windows), for the tcgen05 kernel (64 x 480), the 22 x 1125 kernel and a shape that runs on the generic kernel.

This file is collected last on purpose (a failure here hides nothing)."""
import ctypes

import numpy as np
import pytest
import torch

from mibminet_b200 import _lib, convert, synth
from mibminet_b200.engine import Int8Net

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 32 * 1024 * 1024          # MIBMI_NOALLOC_SCRATCH_BYTES


def _net(C, T, seed):
    params = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=seed), C=C, T=T)
    return Int8Net(params, device="cuda:0")


def _noalloc(net, x, hop, out):
    """The new entry point, straight through the C-ABI; every tensor it touches exists before the call."""
    L = _lib.lib()
    S, Ls = x.shape[0], x.shape[2]
    stream = torch.cuda.current_stream(net.device).cuda_stream
    return L.mibmi_forward_stream_int8_noalloc(net._h, ctypes.c_void_p(x.data_ptr()), S, Ls, int(hop),
                                               ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(stream))


@pytest.mark.parametrize("C,T", [(64, 480), (22, 1125), (5, 64)])
def test_noalloc_stream_matches_existing_entry_point_twice(C, T):
    net = _net(C, T, seed=12)
    g = torch.Generator(device="cuda")
    g.manual_seed(C * 1000 + T)
    # call 1: hop 24 (the existing entry point takes its gather path); call 2: fewer, shorter streams, hop 16 (the existing entry
    # point takes the fused streaming kernel where the shape has one) -- the shared scratch still holds call 1's windows
    for S, extra, hop in ((7, 24 * 40, 24), (3, 16 * 9, 16), (7, 24 * 40, 24)):
        Ls = T + extra
        x = torch.randint(-127, 128, (S, C, Ls), dtype=torch.int8, device="cuda", generator=g)
        W = net.stream_windows(Ls, hop)
        assert S * W * C * T <= SCRATCH_BYTES
        ref = net.forward_stream(x, hop=hop)                       # existing entry point
        out = torch.full((S, W, net.N), -99, dtype=torch.int8, device="cuda")
        torch.cuda.synchronize()
        rc = _noalloc(net, x, hop, out)
        assert rc == 0, _lib.lib().mibmi_strerror(rc)
        torch.cuda.synchronize()
        assert torch.equal(out, ref), (C, T, S, hop)
        # and against the materialised windows on the per-trial path
        win = torch.stack([x[:, :, w * hop:w * hop + T] for w in range(W)], dim=1).reshape(S * W, C, T).contiguous()
        assert torch.equal(out.reshape(S * W, net.N), net.forward(win))


def test_noalloc_stream_rejects_what_does_not_fit():
    C, T = 64, 480
    net = _net(C, T, seed=12)
    max_windows = SCRATCH_BYTES // (C * T)                          # 1092
    hop = 8
    Ls = T + hop * (max_windows - 1)                                # exactly max_windows windows of one stream: fits
    x = torch.zeros((2, C, Ls), dtype=torch.int8, device="cuda")
    out = torch.empty((2, max_windows, net.N), dtype=torch.int8, device="cuda")
    assert _noalloc(net, x[:1], hop, out) == 0
    torch.cuda.synchronize()
    assert torch.equal(out[:1], net.forward_stream(x[:1], hop=hop))
    n0 = net.launch_count
    rc = _noalloc(net, x, hop, out)                                 # two streams: twice the scratch
    assert rc == -7 and net.launch_count == n0                      # MIBMI_ERR_TOO_LARGE, nothing launched
    msg = _lib.lib().mibmi_strerror(rc)
    assert b"scratch" in msg or "scratch" in str(msg)
    assert _noalloc(net, x[:1], 0, out) == -3                       # MIBMI_ERR_PARAM as in the existing entry point
