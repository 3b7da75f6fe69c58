"""Trial-batch sharding across GPUs.

Trials are independent and the network is ~18 kB, so multi-GPU inference is: replicate the network,
give every rank a contiguous slice of the trials (streaming: a slice of the subjects), run the same
kernel, and gather the `int8 [B, N]` logits once at the end -- one NCCL all_gather over NVLink, or a
host-side concatenation.  There is no collective inside the network.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_items: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced partition: the first `n_items % world` ranks hold one item more."""
    if not (0 <= rank < world) or n_items < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(n_items: int, world: int) -> List[int]:
    return [shard_bounds(n_items, world, r)[1] - shard_bounds(n_items, world, r)[0] for r in range(world)]


def gather_logits(local: torch.Tensor, n_total: int, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All ranks get the logits of all `n_total` trials, in trial order.

    `local` is this rank's `[n_local, N]` int8 block as produced by `shard_bounds`.  Equal shards use a
    single all_gather_into_tensor; ragged shards are padded to the largest shard (4 B per trial, so the
    padding is noise)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = shard_sizes(n_total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank}: expected {sizes[rank]} local trials, got {local.shape[0]}")
    n_cls = local.shape[1]
    mx = max(sizes)
    if mx == 0:
        return local.new_empty((0, n_cls))
    send = local
    if local.shape[0] != mx:
        send = local.new_zeros((mx, n_cls))
        send[:local.shape[0]] = local
    out = local.new_empty((world * mx, n_cls))
    dist.all_gather_into_tensor(out, send.contiguous(), group=group)
    if all(s == mx for s in sizes):
        return out
    return torch.cat([out[r * mx:r * mx + sizes[r]] for r in range(world)], dim=0)


class ShardedInt8Net:
    """One `Int8Net` per rank; `forward(x_all_or_local)` returns the gathered logits on every rank."""

    def __init__(self, net, group: Optional[dist.ProcessGroup] = None):
        self.net = net
        self.group = group

    def forward_local(self, x_local: torch.Tensor, **kw) -> torch.Tensor:
        return self.net.forward(x_local, **kw)

    def forward(self, x_local: torch.Tensor, n_total: int, **kw) -> torch.Tensor:
        return gather_logits(self.forward_local(x_local, **kw), n_total, self.group)


def destination_ptrs(ptrs: List[int], rank: int, mode: str = "all", root: int = 0) -> List[int]:
    """Destination buffers of one rank's kernel, own buffer first (the kernel's primary destination)."""
    if mode == "all":
        return [ptrs[rank]] + [p for r, p in enumerate(ptrs) if r != rank]
    if mode == "root":
        return [ptrs[rank]] if rank == root else [ptrs[rank], ptrs[root]]
    if mode == "none":
        return [ptrs[rank]]
    raise ValueError("mode")


class PeerGather:
    """All-gather fused into the kernel (mibmi_forward_int8_gather): every rank owns a `[n_total, N]` int8 buffer
    in CUDA symmetric memory; a rank's kernel stores the logits of its own trials into the same rows of ALL
    buffers over NVLink (P2P stores from the epilogue), so no collective is launched per forward.

        pg = PeerGather(net, n_total)            # collective: allocation + rendezvous
        pg.forward(x_local)                      # asynchronous, no communication call
        ...
        full = pg.wait()                         # one barrier; `full` holds every rank's logits

    Raises RuntimeError when symmetric memory cannot be set up (the caller then uses `gather_logits`)."""

    MODES = ("all", "root", "none")

    def __init__(self, net, n_total: int, group: Optional[dist.ProcessGroup] = None, mode: str = "all", root: int = 0):
        """mode 'all': every rank ends up with every rank's logits (all-gather, N - 1 remote 16-byte stores per quad);
        'root': only rank `root`'s buffer is complete (a gather: one remote store per quad from the other ranks, none from the
        root); 'none': local rows only (no remote store at all; isolates the cost of the peer stores in measurements)."""
        import torch.distributed._symmetric_memory as symm_mem
        if mode not in self.MODES:
            raise ValueError(f"mode must be one of {self.MODES}")
        self.mode, self.root = mode, int(root)
        self.net = net
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.n_total = int(n_total)
        self.lo, self.hi = shard_bounds(self.n_total, self.world, self.rank)
        try:
            self.buf = symm_mem.empty((max(self.n_total, 1), net.N), dtype=torch.int8, device=net.device)
            self.handle = symm_mem.rendezvous(self.buf, self.group.group_name)
            ptrs = [int(p) for p in self.handle.buffer_ptrs]
        except Exception as e:   # noqa: BLE001 -- any failure means "not available here"
            raise RuntimeError(f"symmetric memory unavailable: {e}") from e
        if len(ptrs) != self.world:
            raise RuntimeError("symmetric memory rendezvous returned an unexpected number of peers")
        # own buffer first: the kernel's primary destination
        self.ptrs = destination_ptrs(ptrs, self.rank, mode, self.root)

    def forward(self, x_local: torch.Tensor, **kw) -> None:
        if x_local.shape[0] != self.hi - self.lo:
            raise ValueError(f"rank {self.rank}: expected {self.hi - self.lo} local trials, got {x_local.shape[0]}")
        self.net.forward_gather(x_local, self.ptrs, self.lo, **kw)

    def wait(self) -> torch.Tensor:
        """Barrier across the ranks (on the current stream): afterwards every row of the buffer is final."""
        self.handle.barrier()
        return self.buf[:self.n_total]
