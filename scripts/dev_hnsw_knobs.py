#!/usr/bin/env python
"""Dev probe: cfg-4 HNSW search kernel time against the resident-CTA cap (vdb_debug_set_hnsw_ctas_per_sm)."""
import os, random, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops
from vector_db_b200.indexing import HNSW
dev = torch.device("cuda", 0)
N, D, M, ef, Q, K = 1_000_000, 768, 32, 128, 10_000, 10
X = torch.randn((N, D), device=dev, generator=torch.Generator(device=dev).manual_seed(1234))
Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(4321))
idx = HNSW(M=M, ef_construction=200, rng=random.Random(42), device=dev)
idx.build_from_arrays(range(N), X)
vec = idx._vectors
entry_row, top = idx._read_state()
lib = _lib.load()
ref = None
for spec in [a if a[0] == "c" else "c" + a for a in (sys.argv[1:] or ["c8", "c9", "c10", "c12", "c16"])]:
    cap = int(spec[1:])                                               # cN: N resident CTAs per SM
    lib.vdb_debug_set_hnsw_ctas_per_sm(cap)
    def step():
        return ops.hnsw_search(Qd, vec.vecs, vec.n, idx._adj0, idx._adj_up, idx._up_ptr_dev, entry_row, top, ef, K,
                               want_stats=True, levels=idx._levels_dev)
    for _ in range(2):
        out = step()
    torch.cuda.synchronize()
    _lib.profile_enable(True); _lib.profile_read("hnsw_search")
    for _ in range(5):
        out = step()
    torch.cuda.synchronize()
    n, ms = _lib.profile_read("hnsw_search"); _lib.profile_enable(False)
    if ref is None:
        ref = out
    same = bool(torch.equal(out[1], ref[1]) and torch.equal(out[2], ref[2]))
    print(f"{spec:>4s}: kernel {ms / n:.3f} ms -> {Q / (ms / n) * 1e3:.0f} QPS  same_results={same}", flush=True)
