#!/usr/bin/env python
"""How often the tensor-core scan's certificate rejects a query, by number of candidate bins (debug flags)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib  # noqa: E402
from vector_db_b200.indexing import ExactIndex  # noqa: E402

dev = torch.device("cuda", 0)
N, D, Q, K = 1_000_000, 384, 4096, 10
lib = _lib.load()
tot = {}
for seed in range(1234, 1234 + 8):
    X = torch.randn((N, D), device=dev, generator=torch.Generator(device=dev).manual_seed(seed), dtype=torch.float32)
    for metric in ("cosine", "l2"):
        index = ExactIndex(D, metric, device=dev)
        index.add_device(torch.arange(N, dtype=torch.int64, device=dev), X)
        for qs in (4321, 99):
            Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(qs))
            for nb in (32, 64, 128):
                lib.vdb_debug_set_tc_flags(nb << 12)
                index.search_device(Qd, K)
                u = index.last_uncertified
                tot[(metric, nb)] = tot.get((metric, nb), 0) + u
        del index
    print(seed, {f"{m}/{nb}": v for (m, nb), v in tot.items()}, flush=True)
lib.vdb_debug_set_tc_flags(0)
print("queries per cell:", 8 * 2 * Q)
