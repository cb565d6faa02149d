#!/usr/bin/env python
"""Development probe: time the packed all-gather + merge of sharded.py on its own (torchrun, N ranks)."""
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
d = torch.rand((Q, K), device=dev).sort(dim=1).values
i = torch.randint(0, 1 << 40, (Q, K), device=dev)
for name, fn in (("gather", lambda: sharded.gather_topk(d, i)),
                 ("gather+merge", lambda: sharded.merge_shards(*sharded.gather_topk(d, i), descending=False))):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"{name}: {e0.elapsed_time(e1) / 20:.3f} ms/iter (wall {(time.perf_counter() - t0) * 50:.3f} ms/iter)", flush=True)
dist.destroy_process_group()
