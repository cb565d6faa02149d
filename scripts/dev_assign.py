#!/usr/bin/env python
"""Where a k-means assignment round (IvfIndex._assign) spends its time, and how the number of reranked
bins for nearest-row calls changes it.  Development probe, not a bench line.

    python scripts/dev_assign.py [N=200000]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
D = 384
CHUNK = 65536
g = torch.Generator(device="cuda").manual_seed(7)
X = torch.randn((N, D), device="cuda", generator=g)
lib = _lib.load()
out = []
for K in (256, 1024, 4096):
    cent = X[torch.randperm(N, device="cuda", generator=g)[:K]].clone()
    # a few Lloyd rounds so that the centroids look like trained ones
    for _ in range(3):
        _, lab = ops.exact_topk_f32(X, cent, 1, "l2")
        lab = lab[:, 0]
        sums = torch.zeros_like(cent).index_add_(0, lab, X)
        cnt = torch.bincount(lab, minlength=K).clamp(min=1).unsqueeze(1)
        cent = sums / cnt
    _, ref = ops.exact_topk_f32(X, cent, 1, "l2")
    mirror, sq, max_norm = ops.tc_prepare_rows(cent, "l2")       # max_norm: the 2-float stats tensor
    for nbf, flags in (("trim", 0), ("all32", 16), ("trim16", 16 << 12), ("trim8", 8 << 12)):
        lib.vdb_debug_set_tc_flags(flags)
        for rep in range(3):
            _lib.profile_enable(rep == 2)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            unc = 0
            same = True
            for s in range(0, N, CHUNK):
                e = min(N, s + CHUNK)
                _, i, u = ops.exact_topk_tc(X[s:e], cent, mirror, max_norm, 1, "l2", return_uncertified=True)
                if rep == 2:
                    unc += int(u)
                    same = same and bool((i == ref[s:e]).all())
            torch.cuda.synchronize()
            if rep == 1:
                dt = time.perf_counter() - t0
        slots = {s: _lib.profile_read(s) for s in ("exact_tc", "rerank", "exact_f32", "finalize")}
        _lib.profile_enable(False)
        out.append({"K": K, "nbf": nbf, "ms": dt * 1e3, "uncertified": unc, "identical": same,
                    "slots_ms": {s: round(v[1], 3) for s, v in slots.items()}})
        print(json.dumps(out[-1]), flush=True)
    lib.vdb_debug_set_tc_flags(0)
    # the fp32 scan as the assignment engine, for comparison
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(0, N, CHUNK):
        ops.exact_topk_f32(X[s:min(N, s + CHUNK)], cent, 1, "l2")
    torch.cuda.synchronize()
    print(json.dumps({"K": K, "engine": "f32", "ms": (time.perf_counter() - t0) * 1e3}), flush=True)

# ---- one IvfIndex.build_from_arrays, phase by phase
from vector_db_b200.indexing import IvfIndex  # noqa: E402

Xh = X.cpu().numpy()
for K in (256, 1024):
    idx = IvfIndex(k=K)
    t0 = time.perf_counter()
    idx.build_from_arrays(range(N), Xh, iters=0)
    torch.cuda.synchronize()
    t_setup = time.perf_counter() - t0
    cent = X[:K].clone()
    ph = {"assign": 0.0, "csr": 0.0, "mean": 0.0}
    unc_rounds = []
    for r in range(20):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        lab = idx._assign(X, cent)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        ptr_, order = idx._labels_to_csr(lab)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        ops.ivf_centroid_mean(X, ptr_, order, cent)
        torch.cuda.synchronize(); t3 = time.perf_counter()
        ph["assign"] += t1 - t0; ph["csr"] += t2 - t1; ph["mean"] += t3 - t2
        unc_rounds.append(round((t1 - t0) * 1e3, 2))
    print(json.dumps({"K": K, "setup_iters0_s": t_setup, "per_round_ms": {k: v / 20 * 1e3 for k, v in ph.items()},
                      "assign_ms_by_round": unc_rounds}), flush=True)
