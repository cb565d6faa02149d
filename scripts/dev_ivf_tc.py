#!/usr/bin/env python
"""Dev probe: list-major tensor-core IVF batch search -- uncertified rate, equality with the query-major scan, timing."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
from vector_db_b200.indexing import IvfIndex
n, dim, K, nq = [int(a) for a in (sys.argv[1:5] if len(sys.argv) > 4 else (120000, 64, 256, 700))]
kind = sys.argv[5] if len(sys.argv) > 5 else "clustered"
iters = int(sys.argv[6]) if len(sys.argv) > 6 else 4
rng = np.random.default_rng(1)
if kind == "clustered":
    centres = rng.standard_normal((K, dim)).astype(np.float32) * 3
    X = (centres[rng.integers(0, K, n)] + rng.standard_normal((n, dim))).astype(np.float32)
    Q = (centres[rng.integers(0, K, nq)] + rng.standard_normal((nq, dim))).astype(np.float32)
else:
    X = rng.standard_normal((n, dim), dtype=np.float32)
    Q = rng.standard_normal((nq, dim), dtype=np.float32)
idx = IvfIndex(k=K)
idx.build_from_arrays(range(n), X, iters=iters, init_rows=rng.choice(n, K, replace=False))
print(kind, "iters", iters)
sizes = [len(l) for l in idx.inverted_lists]
print("lists min/mean/max", min(sizes), sum(sizes) / len(sizes), max(sizes))
qd = torch.from_numpy(Q).cuda()
for n_probe in (1, 4, 16):
    for flags in (0, 16):
        _lib.load().vdb_debug_set_tc_flags(flags)
        idx.batch_engine = "scan"
        Ds, Is = idx.search_batch(Q, n_probe, 10)
        idx.batch_engine = "tc"
        Dt, It = idx.search_batch(Q, n_probe, 10)
        def t(engine):
            idx.batch_engine = engine
            idx.search_device(qd, n_probe, 10)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(5): idx.search_device(qd, n_probe, 10)
            torch.cuda.synchronize(); return (time.perf_counter() - t0) / 5 * 1e3
        print(f"n_probe {n_probe:2d} flags {flags:2d}: equal ids {np.array_equal(It, Is)} dists {np.array_equal(Dt, Ds)} "
              f"uncertified {idx.last_uncertified}/{nq}  tc {t('tc'):.3f} ms  scan {t('scan'):.3f} ms", flush=True)
_lib.load().vdb_debug_set_tc_flags(0)
