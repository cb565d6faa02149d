#!/usr/bin/env python
"""A/B of the candidate-bin count / trimming of the exact tensor-core scan on one box (debug flags)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib  # noqa: E402
from vector_db_b200.indexing import ExactIndex  # noqa: E402

dev = torch.device("cuda", 0)
N, D, Q, K = 1_000_000, 384, 4096, 10
lib = _lib.load()
for seed in (1234, 1235):
    X = torch.randn((N, D), device=dev, generator=torch.Generator(device=dev).manual_seed(seed))
    Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(4321))
    for metric in ("cosine", "l2"):
        index = ExactIndex(D, metric, device=dev)
        index.add_device(torch.arange(N, dtype=torch.int64, device=dev), X)
        for name, flags in (("nb64+trim", 0), ("nb32+trim", 32 << 12), ("nb32 no trim", (32 << 12) | 16), ("nb64+trim", 0)):
            lib.vdb_debug_set_tc_flags(flags)
            for _ in range(3):
                index.search_device(Qd, K)
            torch.cuda.synchronize()
            _lib.profile_enable(True)
            for s in ("exact_tc", "rerank", "exact_f32"):
                _lib.profile_read(s)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                index.search_device(Qd, K)
            e1.record()
            torch.cuda.synchronize()
            pr = {s: _lib.profile_read(s) for s in ("exact_tc", "rerank", "exact_f32")}
            _lib.profile_enable(False)
            print(f"seed {seed} {metric:6s} {name:13s}: step {e0.elapsed_time(e1) / 20:.3f} ms  uncertified {index.last_uncertified}  "
                  + "  ".join(f"{s} {v[1] / max(v[0], 1):.4f}x{v[0]}" for s, v in pr.items()), flush=True)
        lib.vdb_debug_set_tc_flags(0)
        del index
