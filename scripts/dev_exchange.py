#!/usr/bin/env python
"""Multi-process check + timing of the fused shard exchange (torchrun, N ranks, one GPU each):
PeerExchange over torch symmetric memory against gather_topk + merge_shards on the same lists.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/dev_exchange.py
"""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import sharded  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
Q, K = 4096, 10
ex = sharded.PeerExchange(Q * K, dev)
print(f"[rank {rank}] exchange buffers mapped: {[hex(int(p)) for p in ex.peers.tolist()]}", flush=True)
ok = True
for call in range(6):
    g = torch.Generator(device=dev).manual_seed(100 * call + rank)
    d = torch.sort(torch.rand((Q, K), device=dev, generator=g), dim=1, descending=True).values
    i = torch.randint(0, 1_000_000, (Q, K), device=dev, generator=g) + rank * 1_000_000
    want = sharded.merge_shards(*sharded.gather_topk(d, i), descending=True)
    got = ex.exchange_merge(d, i, descending=True)
    torch.cuda.synchronize()
    ex.check()
    same = bool((got[0] == want[0]).all()) and bool((got[1] == want[1]).all())
    ok = ok and same
print(f"[rank {rank}] fused exchange == all-gather + merge: {ok}", flush=True)


def timed(name, fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / iters * 1e3
    print(f"[rank {rank}] {name}: {dt:.4f} ms/iter", flush=True)


timed("nccl all-gather + merge", lambda: sharded.merge_shards(*sharded.gather_topk(d, i), descending=True))
timed("fused peer exchange", lambda: ex.exchange_merge(d, i, descending=True))
ex.check()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
