#!/usr/bin/env python
"""Dev probe: probe selection through the tensor cores (ops.ivf_probe_tc) against the numpy-order kernels."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
Q, D, K = 4096, 384, 1024
g = torch.Generator(device="cuda").manual_seed(1)
q = torch.randn((Q, D), device="cuda", generator=g)
cent = torch.randn((K, D), device="cuda", generator=g) * 0.3
ptr = torch.arange(K + 1, device="cuda", dtype=torch.int32) * 900
mirror, _, stats = ops.tc_prepare_rows(cent, "l2")
def T(f, n=20):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = f()
    e1.record(); torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) / n
for n_probe in (1, 4, 8, 12):
    (p0, x0), t0 = T(lambda: ops.ivf_probe(ops.ivf_centroid_dist(q, cent), n_probe, ptr))
    (p1, x1, fl, nu), t1 = T(lambda: ops.ivf_probe_tc(q, cent, mirror, stats, n_probe, ptr))
    _, t2 = T(lambda: ops.exact_topk_tc(q, cent, mirror, stats, n_probe + 4, "l2"))
    print(f"n_probe {n_probe}: numpy-order kernels {t0:.3f} ms; tc path {t1:.3f} ms (of which exact_topk_tc {t2:.3f}); probes equal {bool(torch.equal(p0, p1))} "
          f"prefix equal {bool(torch.equal(x0, x1))}; flagged {int(fl.sum())}/{Q}, rescanned {int(nu.item())}")
