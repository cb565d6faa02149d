#!/usr/bin/env python
"""Development probe: per-phase wall times of the multi-GPU bench step (torchrun, N ranks)."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import sharded  # noqa: E402
from vector_db_b200.indexing import ExactIndex  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
N, D, Q, K = 1_000_000, 384, 4096, 10
g = torch.Generator(device=dev).manual_seed(1234 + rank)
X = torch.randn((N, D), device=dev, generator=g)
Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(4321))
index = ExactIndex(D, "cosine", device=dev)
index.add_device(torch.arange(rank * N, (rank + 1) * N, dtype=torch.int64, device=dev), X)


def timed(name, fn, iters=10):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(iters):
        out = fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / iters * 1e3
    print(f"[rank {rank}] {name}: {dt:.3f} ms/iter", flush=True)
    return out


d, i = timed("search_device", lambda: index.search_device(Qd, K))
from vector_db_b200 import _lib  # noqa: E402
_lib.profile_enable(True)
for slot in ("exact_f32", "exact_tc", "rerank", "finalize"):
    _lib.profile_read(slot)
index.search_device(Qd, K)
torch.cuda.synchronize()
print(f"[rank {rank}] uncertified={index.last_uncertified} " + " ".join(f"{s}={_lib.profile_read(s)}" for s in ("exact_f32", "exact_tc", "rerank", "finalize")), flush=True)
_lib.profile_enable(False)
gd, gi = timed("gather", lambda: sharded.gather_topk(d, i))
timed("merge", lambda: sharded.merge_shards(gd, gi, descending=True))
timed("search+gather", lambda: sharded.gather_topk(*index.search_device(Qd, K)))
timed("full step", lambda: sharded.merge_shards(*sharded.gather_topk(*index.search_device(Qd, K)), descending=True))
ex = sharded.PeerExchange(Q * K, dev)
timed("full step, fused peer exchange", lambda: ex.exchange_merge(*index.search_device(Qd, K), descending=True))
timed("search_device again", lambda: index.search_device(Qd, K))


def ev_timed(name, fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    host_ms = (time.perf_counter() - t0) / iters * 1e3
    torch.cuda.synchronize()
    print(f"[rank {rank}] {name}: device {e0.elapsed_time(e1) / iters:.3f} ms/iter, host enqueue {host_ms:.3f} ms/iter", flush=True)


ev_timed("search_device (events)", lambda: index.search_device(Qd, K))
ev_timed("full step fused (events)", lambda: ex.exchange_merge(*index.search_device(Qd, K), descending=True))
ex.check()
dist.destroy_process_group()
