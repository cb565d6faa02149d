#!/usr/bin/env python
"""Training-path timings: ProductQuantizationService.compress (100 Lloyd rounds + encode, the reference's
pq.py:91-108) and IvfIndex.build_from_arrays (100 Lloyd rounds, ivf.py:75).

    python scripts/bench_train.py [N_pq=100000] [N_ivf=200000]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200.indexing import IvfIndex  # noqa: E402
from vector_db_b200.inference.pq import ProductQuantizationService  # noqa: E402

N_PQ = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
N_IVF = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
rng = np.random.default_rng(1234)
X = rng.standard_normal((max(N_PQ, N_IVF), 384), dtype=np.float32)
out = {}
svc = ProductQuantizationService(k=256, chunks=48, dim=384)
svc.compress(X[:4096])                      # warm-up (module load, workspaces)
torch.cuda.synchronize()
t0 = time.perf_counter()
codes = svc.compress(X[:N_PQ])
torch.cuda.synchronize()
out["pq_compress"] = {"rows": N_PQ, "m": 48, "K": 256, "lloyd_rounds": 100, "seconds": time.perf_counter() - t0,
                      "codes_shape": list(codes.shape), "reference_seconds_at_4000_rows": 38.0}
for K in (256, 1024):
    idx = IvfIndex(k=K)
    idx.build_from_arrays(range(20000), X[:20000], iters=2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    idx = IvfIndex(k=K)
    idx.build_from_arrays(range(N_IVF), X[:N_IVF], iters=100)
    torch.cuda.synchronize()
    out[f"ivf_build_K{K}"] = {"rows": N_IVF, "D": 384, "K": K, "lloyd_rounds": 100, "seconds": time.perf_counter() - t0}
print(json.dumps(out), flush=True)
