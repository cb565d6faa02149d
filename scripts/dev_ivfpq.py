#!/usr/bin/env python
"""IVF-PQ ADC scan alone: clustered rows, one trained index, kernel time per (n_probe, shortlist).

    python scripts/dev_ivfpq.py [N] [Q] [once]      # default 1_000_000 4096; `once` = one call per case (for ncu)
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from bench_paths import HBM, _time_slot, dev  # noqa: E402
from vector_db_b200.indexing import IvfPqIndex  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    Q = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    once = len(sys.argv) > 3
    D, K = 384, 1024
    g = torch.Generator(device=dev).manual_seed(1234)
    centres = torch.randn((1024, D), device=dev, generator=g)
    sigma = 0.15 * centres.norm(dim=1).mean() / (D ** 0.5)
    X = centres[torch.randint(0, 1024, (N,), device=dev, generator=g)] + sigma * torch.randn((N, D), device=dev, generator=g)
    Qd = centres[torch.randint(0, 1024, (Q,), device=dev, generator=g)] + sigma * torch.randn((Q, D), device=dev, generator=g)
    idx = IvfPqIndex(k=K, chunks=D // 8)
    idx.build_from_arrays(range(N), X.cpu().numpy(), iters=10, init_rows=np.random.default_rng(777).choice(N, K, replace=False))
    idx.train_pq(iters=10)
    idx._list_codes()
    torch.cuda.synchronize()
    rows = []
    for n_probe, R in ((4, 100), (4, 1000), (16, 1000)):
        fn = lambda: idx.search_device(Qd, n_probe, 10, shortlist=R)  # noqa: E731
        (d, i, scanned), k_ms, step_ms = _time_slot("adc_scan", fn, iters=1 if once else 5, warm=0 if once else 2)
        alg = float(scanned.sum().item()) * (D // 8)
        rows.append({"n_probe": n_probe, "shortlist": R, "adc_kernel_ms": round(k_ms, 3), "step_ms": round(step_ms, 3),
                     "rows_per_query": float(scanned.float().mean().item()), "max_rows": int(scanned.max().item()),
                     "frac": round(alg / (k_ms * 1e-3) / 1e9 / HBM, 4)})
    print(json.dumps({"path": "ivf_pq_adc", "N": N, "Q": Q, "rows": rows}), flush=True)


if __name__ == "__main__":
    main()
