#!/usr/bin/env python
"""Dev probe: where the blocking ExactIndex.search call spends its time beyond the device step."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200.indexing import ExactIndex
N, D, Q, K = 1_000_000, 384, 4096, 10
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn((N, D), device="cuda", generator=g)
idx = ExactIndex(D, "cosine")
idx.add_device(torch.arange(N, device="cuda"), X) if hasattr(idx, "add_device") else idx.add(range(N), X.cpu().numpy())
qp = torch.randn((Q, D)).pin_memory()
qd = qp.cuda()
for _ in range(5): idx.search(qp, K)
def T(f, n=30):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("search(pinned)            %.3f ms" % T(lambda: idx.search(qp, K)))
print("search(pageable)          %.3f ms" % T(lambda: idx.search(qp.numpy().copy(), K)))
print("search_device + sync      %.3f ms" % T(lambda: (idx.search_device(qd, K), torch.cuda.synchronize())))
print("search_device back-to-back %.3f ms" % T(lambda: idx.search_device(qd, K)))
t0 = time.perf_counter(); 
for _ in range(30): idx.search_device(qd, K)
cpu = (time.perf_counter() - t0) / 30 * 1e3; torch.cuda.synchronize()
print("search_device CPU enqueue time %.3f ms" % cpu)
print("h2d 6.3MB + sync          %.3f ms" % T(lambda: (qd.copy_(qp, non_blocking=True), torch.cuda.synchronize())))
d, i = idx.search_device(qd, K)
pd, pi = torch.empty_like(d, device="cpu").pin_memory(), torch.empty_like(i, device="cpu").pin_memory()
print("d2h results + sync        %.3f ms" % T(lambda: (pd.copy_(d, non_blocking=True), pi.copy_(i, non_blocking=True), torch.cuda.synchronize())))
print("numpy copies of results   %.3f ms" % T(lambda: (pd.numpy().copy(), pi.numpy().copy())))
print("is_pinned()               %.4f ms" % T(lambda: qp.is_pinned(), 200))
print("torch.empty device        %.4f ms" % T(lambda: torch.empty((Q, D), dtype=torch.float32, device="cuda"), 200))
def manual():
    q = torch.empty((Q, D), dtype=torch.float32, device="cuda")
    q.copy_(qp, non_blocking=True)
    d, i = idx.search_device(q, K)
    pd.copy_(d, non_blocking=True); pi.copy_(i, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return pd.numpy().copy(), pi.numpy().copy()
print("manual sequence           %.3f ms" % T(manual))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20): idx.search(qp, K)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
