"""Two ranks on two GPUs (NCCL + CUDA symmetric memory): the rendezvous / barrier path of `sharding.PeerGather`, i.e. the fused
gather the multi-GPU bench lines use.  Skipped where fewer than two GPUs are visible (the round-end GPU tier runs on
one); run by tools/gpu_r2p.sh on a two-GPU box, log under profiles/.  Every mode is checked against the oracle on ALL trials:
  all   every rank ends up with every logit (all-gather through peer stores)
  root  rank 0 ends up with every logit, the other ranks with their own rows (gather)
  none  own rows only
for both the tcgen05 kernel with quads (64 x 480) and the 22 x 1125 kernel (groups of three), equal and ragged shards."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "oracle")]
    import wolf_port
    from mibminet_b200 import convert, synth
    from mibminet_b200.engine import Int8Net
    from mibminet_b200.sharding import PeerGather, shard_bounds
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ok, notes = True, []
    try:
        for (C, T), n_total in (((64, 480), 1024), ((64, 480), 1023), ((22, 1125), 599), ((16, 480), 130)):
            params = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=3), C=C, T=T)
            net = Int8Net(params, device=dev)
            x = synth.make_trials(n_total, C, T, seed=11)             # same on every rank
            ref = wolf_port.forward(params.to_wolf_arrays(), C, T, x)
            lo, hi = shard_bounds(n_total, world, rank)
            xl = torch.from_numpy(x[lo:hi]).to(dev)
            for mode in ("all", "root", "none"):
                pg = PeerGather(net, n_total, mode=mode)
                pg.buf.fill_(-77)
                torch.cuda.synchronize()
                dist.barrier()
                for _ in range(2):                                        # twice: a second forward overwrites the same rows
                    pg.forward(xl)
                full = pg.wait().cpu().numpy()
                own = np.array_equal(full[lo:hi], ref[lo:hi])
                if mode == "all" or (mode == "root" and rank == 0):
                    good = own and np.array_equal(full, ref)
                else:
                    other = np.ones(n_total, bool)
                    other[lo:hi] = False
                    good = own and bool((full[other] == -77).all())     # nobody wrote rows this rank should not receive
                ok = ok and good
                notes.append((C, T, n_total, mode, bool(good), net.last_kernel))
                dist.barrier()
                del pg
        q.put((rank, ok, notes))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_gather_two_gpus_all_root_none():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 90
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, ok, notes in sorted(res):
        print("rank", rank, notes)
        assert ok, notes
