#!/usr/bin/env python
"""Dev probe: PQ encode (f16 kernel) time for tiles-per-round 2 / 4 / 8 and the two L2 hints of the row loads; run under
`ncu --metrics dram__bytes_read.sum` to see the DRAM traffic of each variant."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
lib = _lib.load()
n, m, K = 2_000_000, 48, 256
g = torch.Generator(device="cuda").manual_seed(5)
x = torch.randn((n, 8 * m), device="cuda", generator=g)
cb = torch.randn((m, K, 8), device="cuda", generator=g)
ref = None
for tiles in (2, 4, 8):
    for hint in (0, 1):
        lib.vdb_debug_set_pq_tc_mode(1 | (tiles << 4) | (hint << 8))
        c = ops.pq_encode(x, cb, engine="tc")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 1 if os.environ.get("ONE") else 5
        e0.record()
        for _ in range(reps): ops.pq_encode(x, cb, engine="tc")
        e1.record(); torch.cuda.synchronize()
        if ref is None: ref = c
        print(f"tiles {tiles} hint {'normal' if hint else 'evict_first'}: {e0.elapsed_time(e1) / reps:.3f} ms, codes equal {bool((c == ref).all())}", flush=True)
