"""Row sharding across the GPUs of one box (no reference precedent -- the reference
scales by replicas only, README.md:116-126).

Rank r owns the contiguous global id range ``shard_bounds(N, r, R)``; every rank answers the same
query batch over its own shard with any of the single-GPU engines, then the per-shard top-k lists
(k x {f32 value, i64 global id} per query, 120 KB per rank at Q=1024, k=10) are exchanged with ONE
all-gather over NCCL/NVLink and merged on the GPU by ``vdb_merge_topk`` (``gather_topk`` +
``merge_shards``), or -- the default on NVLink boxes -- pushed into every peer's exchange buffer and
merged by ONE kernel (``PeerExchange``, csrc/exchange.cu: peer-memory stores, system-scope arrival
counters, merge).  The payload is tiny, so what the fused kernel removes is launches and the
collective's latency.  There is no other data-path communication (codebooks and queries are
replicated).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous id range [lo, hi) of a rank: ceil(N/R) rows each, the last shard may be short/empty."""
    per = -(-int(n_total) // int(world))
    lo = min(int(n_total), rank * per)
    return lo, min(int(n_total), lo + per)


def pack_topk(d: torch.Tensor, i: torch.Tensor) -> torch.Tensor:
    """(f32 [Q,k], i64 [Q,k]) -> one int32 buffer [Q,k,3] so the exchange is a single collective."""
    buf = torch.empty(d.shape + (3,), dtype=torch.int32, device=d.device)
    buf[..., 0] = d.contiguous().view(torch.int32)
    buf[..., 1:] = i.contiguous().view(torch.int32).view(i.shape + (2,))
    return buf


def unpack_topk(buf: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    d = buf[..., 0].contiguous().view(torch.float32)
    i = buf[..., 1:].contiguous().view(torch.int64).squeeze(-1)
    return d, i


def gather_topk(d: torch.Tensor, i: torch.Tensor, group: Optional[dist.ProcessGroup] = None
                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather the per-shard lists: returns (d [R,Q,k], i [R,Q,k]) in rank order on every rank."""
    world = dist.get_world_size(group)
    mine = pack_topk(d, i)
    out = torch.empty((world * mine.shape[0],) + mine.shape[1:], dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(out, mine, group=group)           # rank-major concatenation along dim 0
    return unpack_topk(out.view((world,) + mine.shape))


def merge_shards(d: torch.Tensor, i: torch.Tensor, descending: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """k best of the R gathered lists per query, (value, id) order (shard r owns smaller ids than r+1)."""
    from . import ops
    return ops.merge_topk(d, i, descending=descending)


class PeerExchange:
    """Per-rank exchange buffer mapped into every peer + the call counter of vdb_exchange_merge_topk.

    ``capacity`` = largest Q*k of a call.  Every rank must make the same sequence of calls (same Q, k) on
    one stream -- true for this path: the query batch is replicated.  ``buffers`` (tests only) supplies the
    ranks' buffers directly, e.g. plain device tensors when several "ranks" share one GPU.
    """

    def __init__(self, capacity: int, device, group: Optional[dist.ProcessGroup] = None,
                 rank: Optional[int] = None, world: Optional[int] = None, buffers=None):
        from . import ops
        self.cap = int(capacity)
        self.device = torch.device(device)
        self.calls = 0
        self.sent = [0, 0]
        self.status = torch.zeros((1,), dtype=torch.int32, device=self.device)
        if buffers is not None:
            self.rank, self.world = int(rank), int(world)
            self._keep = buffers
            addrs = [int(b.data_ptr()) for b in buffers]
        else:
            import torch.distributed._symmetric_memory as symm
            group = group or dist.group.WORLD
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
            nbytes = ops.exchange_bytes(self.world, self.cap)
            buf = symm.empty((nbytes,), dtype=torch.uint8, device=self.device)
            buf.zero_()
            torch.cuda.synchronize(self.device)
            hdl = symm.rendezvous(buf, group.group_name)
            dist.barrier(group)                           # every rank's buffer is zeroed before anyone pushes
            self._keep = (buf, hdl)
            addrs = [int(p) for p in hdl.buffer_ptrs]
        if len(addrs) != self.world:
            raise ValueError(f"{len(addrs)} exchange buffers for {self.world} ranks")
        self.peers = torch.tensor(addrs, dtype=torch.int64, device=self.device)

    def exchange_merge(self, d: torch.Tensor, i: torch.Tensor, descending: bool = False
                       ) -> Tuple[torch.Tensor, torch.Tensor]:
        from . import ops
        parity = self.calls & 1
        target = self.sent[parity] + ops.exchange_grid(d.shape[0])
        # the call validates its arguments before it launches: the counters move only once the kernel is enqueued,
        # so a ValueError (Q*k over the capacity, ...) does not leave the arrival target ahead of what was pushed
        out = ops.exchange_merge_topk(d, i, self.peers, self.rank, self.world, self.cap, parity, target,
                                      self.status, descending)
        self.sent[parity] = target
        self.calls += 1
        return out

    def check(self) -> None:
        """Synchronising: raise if any call so far gave up waiting for a peer (such a call returned (inf, -1)
        lists for the affected queries, never a merge of stale slots).  Call it wherever the results are read."""
        if int(self.status.item()) != 0:
            raise RuntimeError("shard exchange: a peer never arrived (wait ran into its clock limit)")


def search_sharded(local_search, q: torch.Tensor, k: int, descending: bool = False,
                   group: Optional[dist.ProcessGroup] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``local_search(q, k) -> (d, global ids)`` on this rank's shard, then gather + merge."""
    d, i = local_search(q, k)
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return d, i
    gd, gi = gather_topk(d, i, group)
    return merge_shards(gd, gi, descending)
