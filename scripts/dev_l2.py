#!/usr/bin/env python
"""Dev probe: exact scan step time, cosine vs L2, 1M x 384, Q = 4096."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
N, D, Q, K = 1_000_000, 384, 4096, 10
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn((N, D), device="cuda", generator=g)
q = torch.randn((Q, D), device="cuda", generator=g)
for metric in ("cosine", "l2"):
    mir, sq, stats = ops.tc_prepare_rows(X, metric)
    out = ops.exact_topk_tc(q, X, mir, stats, K, metric)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.exact_topk_tc(q, X, mir, stats, K, metric)
    e1.record(); torch.cuda.synchronize()
    step = e0.elapsed_time(e1) / 20
    _lib.profile_enable(True)
    for _ in range(10): o = ops.exact_topk_tc(q, X, mir, stats, K, metric)
    torch.cuda.synchronize()
    parts = {s: _lib.profile_read(s) for s in ("exact_tc", "rerank", "exact_f32", "finalize")}
    _lib.profile_enable(False)
    unc = ops.last_uncertified() if hasattr(ops, "last_uncertified") else None
    print(f"{metric}: step {step:.3f} ms; per call: " + ", ".join(f"{k} {v[1] / 10:.3f} ms ({v[0] // 10} launches)" for k, v in parts.items()), "uncertified", unc)
