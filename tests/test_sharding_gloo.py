"""world_size-2 gloo test of the multi-GPU plumbing (CPU): shard arithmetic + logit gather.  The compute
leg is GPU-only, so here each rank's shard is evaluated by the oracle C port as a stand-in."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mibminet_b200.sharding import gather_logits, shard_bounds, shard_sizes


def test_shard_bounds_partition():
    for n in (0, 1, 7, 8, 65536, 65537, 4096 * 31):
        for world in (1, 2, 4, 8):
            b = [shard_bounds(n, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            s = shard_sizes(n, world)
            assert max(s) - min(s) <= 1 and sum(s) == n
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def _worker(rank, world, port, n_total, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "oracle")]
    import wolf_port
    from mibminet_b200 import convert, synth
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        C, T = 8, 480
        arrays = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=2), C=C, T=T).to_wolf_arrays()
        x = synth.make_trials(n_total, C, T, seed=9)          # same on every rank
        lo, hi = shard_bounds(n_total, world, rank)
        local = torch.from_numpy(wolf_port.forward(arrays, C, T, x[lo:hi], threads=1))
        full = gather_logits(local, n_total)
        ref = wolf_port.forward(arrays, C, T, x, threads=1)
        q.put((rank, bool(np.array_equal(full.numpy(), ref)), tuple(full.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [64, 37])   # equal and ragged shards
def test_gather_logits_world2_gloo(n_total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() + n_total) % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(ok and shape == (n_total, 4) for _, ok, shape in res)


def test_destination_ptrs_modes():
    """Peer-store destinations of the fused gather: own buffer first; 'root' = a gather (one remote destination, none on the
    root), 'all' = an all-gather (every peer), 'none' = local rows only."""
    from mibminet_b200.sharding import destination_ptrs
    ptrs = [100, 200, 300, 400]
    assert destination_ptrs(ptrs, 2, "all") == [300, 100, 200, 400]
    assert destination_ptrs(ptrs, 2, "root") == [300, 100]
    assert destination_ptrs(ptrs, 0, "root") == [100]
    assert destination_ptrs(ptrs, 3, "root", root=3) == [400]
    assert destination_ptrs(ptrs, 1, "none") == [200]
    with pytest.raises(ValueError):
        destination_ptrs(ptrs, 1, "ring")
