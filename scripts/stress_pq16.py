#!/usr/bin/env python
"""Stress: the f16-accumulator PQ encoder over many shapes (codes must equal the fp32 kernel's; a hang trips the outer timeout)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
lib = _lib.load()
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
t_end = time.time() + float(sys.argv[2]) if len(sys.argv) > 2 else time.time() + 60
n_cases = 0
while time.time() < t_end:
    m = int(rng.choice([1, 2, 3, 4, 5, 7, 8, 12, 13, 16, 24, 31, 48, 64]))
    K = int(rng.choice([1, 2, 3, 16, 37, 100, 255, 256]))
    n = int(rng.choice([1, 2, 127, 128, 129, 255, 256, 257, 511, 513, 1000, 4096, 20000, 75777, 151000, 300001]))
    tiles = int(rng.choice([2, 4, 8]))
    scale = float(rng.choice([1e-3, 1.0, 50.0]))
    g = torch.Generator(device="cuda").manual_seed(int(rng.integers(1 << 30)))
    x = torch.randn((n, 8 * m), device="cuda", generator=g) * scale
    cb = torch.randn((m, K, 8), device="cuda", generator=g) * scale
    lib.vdb_debug_set_pq_tc_mode(1 | (tiles << 4))
    for rep in range(3):
        c1 = ops.pq_encode(x, cb, engine="tc")
    c0 = ops.pq_encode(x, cb, engine="f32")
    torch.cuda.synchronize()
    if not torch.equal(c0, c1):
        print("MISMATCH", n, m, K, tiles, scale, int((c0 != c1).sum())); sys.exit(1)
    n_cases += 1
lib.vdb_debug_set_pq_tc_mode(1 | (4 << 4))
print(f"stress ok: {n_cases} shapes x 3 runs, all codes equal")
