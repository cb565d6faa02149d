#!/usr/bin/env python
"""Developer probe for the tensor-core exact scan: per-kernel CUDA-event times at BASELINE
configs[1] (1M x 384, Q = 4096, k = 10) for both metrics, the TMA+MMA-only ceiling (epilogue
arithmetic skipped), the number of uncertified queries, and a parity spot check against the fp32 scan.

    python scripts/dev_tc.py [N] [D] [Q]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vector_db_b200 import _lib, ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 384
Q = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
K = 10
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1234)
X = torch.randn((N, D), device=dev, generator=g)
Qd = torch.randn((Q, D), device=dev, generator=g)
lib = _lib.load()
SLOTS = ("exact_tc", "rerank", "finalize", "exact_f32")


def timed(metric, mir, mx, flags, iters=10):
    lib.vdb_debug_set_tc_flags(flags)
    for _ in range(3):
        out = ops.exact_topk_tc(Qd, X, mir, mx, K, metric, return_uncertified=True)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    for s in SLOTS:
        _lib.profile_read(s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = ops.exact_topk_tc(Qd, X, mir, mx, K, metric, return_uncertified=True)
    e1.record()
    torch.cuda.synchronize()
    prof = {s: _lib.profile_read(s) for s in SLOTS}
    _lib.profile_enable(False)
    lib.vdb_debug_set_tc_flags(0)
    per = {s: round(v[1] / max(v[0], 1), 4) for s, v in prof.items() if v[0]}
    total = e0.elapsed_time(e1) / iters
    tf = 2.0 * Q * N * D / (per["exact_tc"] * 1e-3) / 1e12
    print(f"{metric:6s} flags={flags} total {total:.3f} ms/step  kernels {per}  coarse {tf:.0f} TFLOP/s  "
          f"uncertified {out[2]}", flush=True)
    return out


for metric in ("cosine", "l2"):
    mir, sq, mx = ops.tc_prepare_rows(X, metric)                  # mx: the 2-float stats tensor
    timed(metric, mir, mx, 1)      # TMA + MMA only
    timed(metric, mir, mx, 1 | (3 << 8))      # same with a 3-deep TMA ring
    timed(metric, mir, mx, 2)      # + TMEM loads and bin minima, no threshold test / inserts
    timed(metric, mir, mx, 4)      # full epilogue, 16-entry bin lists
    timed(metric, mir, mx, 8)      # full epilogue without the threshold exchange between the two groups
    D_tc, I_tc, nbad = timed(metric, mir, mx, 0)
    ns = min(Q, 128)
    Dr, Ir = ops.exact_topk_f32(Qd[:ns], X, K, metric)
    same = bool(torch.equal(I_tc[:ns], Ir) and torch.equal(D_tc[:ns], Dr))
    print(f"{metric:6s} parity vs fp32 scan on {ns} queries: {'OK' if same else 'MISMATCH'}", flush=True)
    del mir
