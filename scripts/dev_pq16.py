#!/usr/bin/env python
"""Dev probe: f16-accumulator PQ encode (mode 1) against the fp32-accumulator kernel (mode 0): codes equal, time, resolver share."""
import ctypes, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
lib = _lib.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 48
K = int(sys.argv[3]) if len(sys.argv) > 3 else 256
scale = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
d = 8 * m
g = torch.Generator(device="cuda").manual_seed(5)
x = torch.randn((n, d), device="cuda", generator=g) * scale
kind = sys.argv[5] if len(sys.argv) > 5 else "rand"
if kind == "rows":      # K rows (slightly moved) as centroids: many tiny minima; plus exact duplicates of centroids
    cb = x[torch.randperm(n, device="cuda", generator=g)[:K]].reshape(K, m, 8).permute(1, 0, 2).contiguous()
    cb = cb + 0.05 * scale * torch.randn(cb.shape, device="cuda", generator=g)
else:
    cb = torch.randn((m, K, 8), device="cuda", generator=g) * scale
if kind == "rows": x[:1000, :8] = cb[0, torch.arange(1000, device="cuda") % K]      # exact duplicates of centroids in sub-space 0
def run(mode):
    lib.vdb_debug_set_pq_tc_mode(mode)
    codes = ops.pq_encode(x, cb, engine="tc")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ops.pq_encode(x, cb, engine="tc")
    e1.record(); torch.cuda.synchronize()
    return codes, e0.elapsed_time(e1) / 5
c0, t0 = run(0)
c1, t1 = run(1)
if len(sys.argv) > 6:
    for dbg in (1, 3):
        _, td = run(1 | (dbg << 4))
        print(f'  debug bits {dbg & 255} hint {dbg >> 8} ns (1 = no epilogue work, 2 = no MMA, 4 = no split math, 8 = no row TMA): {td:.3f} ms', flush=True)
    lib.vdb_debug_set_pq_tc_mode(1)
st = (ctypes.c_ulonglong * 3)()
ws = ops._ws("pq_encode_tc", lib.vdb_pq_encode_tc_workspace(m), x.device)
_lib.check(lib.vdb_debug_pq_tc_stats(ctypes.c_void_p(ws.data_ptr()), m, st, None))
bad = (c0 != c1)
print(f"n={n} m={m} K={K} scale={scale}: fp32-acc {t0:.3f} ms, f16-acc {t1:.3f} ms; mismatching codes {int(bad.sum())} of {bad.numel()};"
      f" resolver items {st[0]} ({st[0] / (n * m) * 100:.2f} %), all-ones {st[1]}, in place {st[2]}", flush=True)
if bad.any():
    idx = bad.nonzero()[:8].cpu().numpy()
    for r, j in idx: print("  row", r, "sub", j, "fp32", int(c0[r, j]), "f16", int(c1[r, j]))
