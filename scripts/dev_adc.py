#!/usr/bin/env python
"""Development probe for the ADC threshold-collect path: per-phase times and the number of queries
that took the rescan path.  python scripts/dev_adc.py [N] [Q] [m] [R ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
m = int(sys.argv[3]) if len(sys.argv) > 3 else 48
Rs = [int(a) for a in sys.argv[4:]] or [10, 100, 400, 1000]
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(7)
codes = torch.randint(0, 256, (N, m), device=dev, dtype=torch.uint8, generator=g)
lut = torch.rand((Q, m, 256), device=dev, generator=g)


def al(v):
    return (v + 255) & ~255


for R in Rs:
    for it in range(3):
        _lib.profile_enable(True)
        _lib.profile_read("adc_scan"), _lib.profile_read("finalize")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.pq_adc_topk(codes, lut, R)
        e1.record()
        torch.cuda.synchronize()
        a, f = _lib.profile_read("adc_scan"), _lib.profile_read("finalize")
        _lib.profile_enable(False)
        ws = ops._ws("adc", 0, dev)
        head = al(Q * m * 256 * 4) + al(Q * 4)
        cnt = ws[head + al(Q * 4): head + al(Q * 4) + Q * 4].view(torch.int32)
        badn = ws[head + 3 * al(Q * 4): head + 3 * al(Q * 4) + 4].view(torch.int32)
        print(f"R={R} it={it} total={e0.elapsed_time(e1):.3f} scan={a} fin={f} flagged={int(badn[0])} "
              f"count min/mean/max={int(cnt.min())}/{float(cnt.float().mean()):.0f}/{int(cnt.max())}", flush=True)
