#!/usr/bin/env python
"""Dev probe for ncu: the launches of one list-major IVF batch search (1M x 384, K = 1024, Q = 4096, n_probe 4)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200.indexing import IvfIndex
N, D, K, Q = 1_000_000, 384, 1024, 4096
rng = np.random.default_rng(1)
X = rng.standard_normal((N, D), dtype=np.float32)
idx = IvfIndex(k=K)
idx.build_from_arrays(range(N), X, iters=4, init_rows=rng.choice(N, K, replace=False))
qd = torch.randn((Q, D), device="cuda")
idx.batch_engine = "tc"
for _ in range(2): idx.search_device(qd, 4, 10)
torch.cuda.synchronize()
print("MARK", flush=True)
torch.cuda.nvtx.range_push("one")
idx.search_device(qd, 4, 10)
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
