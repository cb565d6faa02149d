#!/usr/bin/env python
"""Dev probe: exact cosine scan, single-CTA kernel (debug bit 5) against the CTA-pair kernel, same process; bit 0 = epilogue off."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
lib = _lib.load()
N, D, Q, K = 1_000_000, 384, int(sys.argv[1]) if len(sys.argv) > 1 else 4096, 10
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn((N, D), device="cuda", generator=g)
q = torch.randn((Q, D), device="cuda", generator=g)
mir, sq, stats = ops.tc_prepare_rows(X, "cosine")
def run(flags, reps=20):
    lib.vdb_debug_set_tc_flags(flags)
    out = ops.exact_topk_tc(q, X, mir, stats, K, "cosine")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): ops.exact_topk_tc(q, X, mir, stats, K, "cosine")
    e1.record(); torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) / reps
(d1, i1), t1 = run(32)
(d2, i2), t2 = run(0)
print(f"Q={Q}: single-CTA step {t1:.3f} ms, CTA-pair step {t2:.3f} ms; ids equal {bool((i1 == i2).all())} dists equal {bool((d1 == d2).all())}")
_, t1e = run(32 | 1); _, t2e = run(1)
print(f"  epilogue off (results invalid): single {t1e:.3f} ms, pair {t2e:.3f} ms")
for st in (4, 6, 8):
    _, t = run(st << 8)
    print(f"  pair, {st} stages: {t:.3f} ms")
lib.vdb_debug_set_tc_flags(0)
