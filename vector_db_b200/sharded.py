"""Row sharding across the GPUs of one box (no reference precedent -- the reference
scales by replicas only, README.md:116-126).

Rank r owns the contiguous global id range ``shard_bounds(N, r, R)``; every rank answers the same
query batch over its own shard with any of the single-GPU engines, then the per-shard top-k lists
(k x {f32 value, i64 global id} per query, 120 KB per rank at Q=1024, k=10) are exchanged with ONE
all-gather over NCCL/NVLink and merged on the GPU by ``vdb_merge_topk``.  The exchange is
latency-bound, so it is a single collective on a packed buffer, not a fused compute+collective
kernel; there is no other data-path communication (codebooks and queries are replicated).
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


def search_sharded(local_search, q: torch.Tensor, k: int, descending: bool = False,
                   group: Optional[dist.ProcessGroup] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``local_search(q, k) -> (d, global ids)`` on this rank's shard, then gather + merge."""
    d, i = local_search(q, k)
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return d, i
    gd, gi = gather_topk(d, i, group)
    return merge_shards(gd, gi, descending)
