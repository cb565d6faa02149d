#!/usr/bin/env python
"""Dev probe: ADC threshold-collect on 16-bit lower-bound tables (8 queries per lookup) vs the 32-bit collect:
same results, times, collected-row counts.  python scripts/dev_adc_coarse.py [N] [Q] [m] [R ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
m = int(sys.argv[3]) if len(sys.argv) > 3 else 48
Rs = [int(a) for a in sys.argv[4:]] or [10, 100, 1000]
dsub = 8
D = m * dsub
dev = torch.device("cuda", 0)
lib = _lib.load()
g = torch.Generator(device=dev).manual_seed(7)
cb = torch.randn((m, 256, dsub), device=dev, generator=g)
codes = torch.empty((N, m), dtype=torch.uint8, device=dev)
step = 1_000_000
for s in range(0, N, step):
    x = torch.randn((min(step, N - s), D), device=dev, generator=g)
    codes[s:s + x.shape[0]] = ops.pq_encode(x, cb)
del x
q = torch.randn((Q, D), device=dev, generator=g)
lut = ops.pq_lut(q, cb)


def al(v):
    return (v + 255) & ~255


def run(R, mask=None, reps=3):
    best = 1e9
    for it in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        d, i = ops.pq_adc_topk(codes, lut, R, row_mask=mask)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    ws = ops._ws("adc", 0, dev)
    head = al(Q * m * 256 * 4) + al(Q * 4)
    cnt = ws[head + al(Q * 4): head + al(Q * 4) + Q * 4].view(torch.int32).clone()
    badn = int(ws[head + 3 * al(Q * 4): head + 3 * al(Q * 4) + 4].view(torch.int32)[0])
    return best, d, i, cnt, badn


for R in Rs:
    out = {}
    for mode in (0, 1):
        lib.vdb_debug_set_adc_coarse(mode)
        t, d, i, cnt, badn = run(R)
        out[mode] = (d, i)
        print(f"N={N} m={m} Q={Q} R={R} coarse={mode}: {t:.2f} ms = {Q / t * 1e3:.0f} QPS; collected min/mean/max "
              f"{int(cnt.min())}/{float(cnt.float().mean()):.0f}/{int(cnt.max())}, rescanned {badn}", flush=True)
    same = torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1])
    print(f"   results identical: {same}", flush=True)
# row mask + ragged last group
mask = torch.randint(-2**31, 2**31 - 1, ((N + 31) // 32,), device=dev, dtype=torch.int32, generator=g)
out = {}
for mode in (0, 1):
    lib.vdb_debug_set_adc_coarse(mode)
    t, d, i, cnt, badn = run(10, mask=mask, reps=2)
    out[mode] = (d, i)
    print(f"masked R=10 coarse={mode}: {t:.2f} ms, rescanned {badn}")
print("   results identical:", torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1]))
lut_full = lut
lut = lut_full[:13].contiguous()
Q = 13
out = {}
for mode in (0, 1):
    lib.vdb_debug_set_adc_coarse(mode)
    t, d, i, cnt, badn = run(10, reps=2)
    out[mode] = (d, i)
    print(f"Q=13 R=10 coarse={mode}: {t:.2f} ms, rescanned {badn}")
print("   results identical:", torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1]))
lib.vdb_debug_set_adc_coarse(1)
