#!/usr/bin/env python
"""Dev probe: sequential-semantics HNSW build (BASELINE configs[0] shape) with 128 / 256 / 512 threads in the insert CTA."""
import os, sys, time, random
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib
from vector_db_b200.indexing import HNSW
from vector_db_b200.types import Node
lib = _lib.load()
N, D = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000, 384
X = np.random.default_rng(1234).standard_normal((N, D), dtype=np.float32)
nodes = [Node(id=i, embedding=X[i]) for i in range(N)]
ref = None
for nt in (128, 256, 512):
    lib.vdb_debug_set_hnsw_insert_threads(nt)
    idx = HNSW(M=16, ef_construction=200, rng=random.Random(42))
    t0 = time.perf_counter()
    idx.build_index(nodes)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    adj0, adj_up, up_ptr, levels, ids = idx._host_graph()
    same = True if ref is None else (np.array_equal(adj0, ref[0]) and np.array_equal(adj_up, ref[1]))
    if ref is None: ref = (adj0.copy(), adj_up.copy())
    print(f"threads {nt}: build {dt:.2f} s = {N / dt:.0f} inserts/s, evals {idx.distance_evaluations()}, graph equal to 128-thread build: {same}", flush=True)
lib.vdb_debug_set_hnsw_insert_threads(256)
