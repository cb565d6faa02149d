#!/usr/bin/env python
"""north_star target check: recall@10 of the PQ path (ADC shortlist of R + exact fp32 rerank) on
1M x 384, m = 48, nbits = 8, against the exact scan, with the time of each stage.

    python scripts/bench_pq_recall.py [N=1000000] [Q=1024] [data=iid|clustered]

iid: N(0,1) rows (the reference's own tests use iid uniform rows) -- PQ recall on unstructured data is
poor by construction.  clustered: 1024 Gaussian blobs, sigma = 0.15 x centre norm
(SURVEY.md 8d), the shape real embedding sets have.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402
from vector_db_b200.indexing import ExactIndex  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
DATA = sys.argv[3] if len(sys.argv) > 3 else "iid"
D, M_SUB, K = 384, 48, 10
dev = torch.device("cuda", 0)
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
g = torch.Generator(device=dev).manual_seed(1234)
if DATA == "clustered":
    centres = torch.randn((1024, D), device=dev, generator=g)
    sigma = 0.15 * centres.norm(dim=1).mean() / (D ** 0.5)      # noise norm = 0.15 x centre norm
    X = centres[torch.randint(0, 1024, (N,), device=dev, generator=g)] + sigma * torch.randn((N, D), device=dev, generator=g)
    Qd = centres[torch.randint(0, 1024, (Q,), device=dev, generator=g)] + sigma * torch.randn((Q, D), device=dev, generator=g)
else:
    X = torch.randn((N, D), device=dev, generator=g)
    Qd = torch.randn((Q, D), device=dev, generator=g)

# train: 25 Lloyd rounds on a 100k-row sample, all sub-spaces per launch
sample = X[torch.randperm(N, device=dev, generator=g)[:100_000]].contiguous()
cb = sample[:256].view(256, M_SUB, 8).permute(1, 0, 2).contiguous()
for _ in range(25):
    ops.pq_kmeans_step(sample, cb)
codes = ops.pq_encode(X, cb)

ex = ExactIndex(D, "l2")
ex.add_device(torch.arange(N, device=dev), X)
_, Iex = ex.search_device(Qd, K)
truth = [set(r) for r in Iex.cpu().tolist()]


def timed(fn, iters=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) / iters


rows = []
lut, lut_ms = timed(lambda: ops.pq_lut(Qd, cb))
for R in (10, 100, 400, 1000, 2000, 4000, 8000):
    if R > N:
        break
    _lib.profile_enable(True)
    _lib.profile_read("adc_scan")
    (d, cand), adc_ms = timed(lambda: ops.pq_adc_topk(codes, lut, R))
    nk, kms = _lib.profile_read("adc_scan")
    _lib.profile_enable(False)
    (dr, ir), rr_ms = timed(lambda: ops.rerank_f32(Qd, X, cand, K, "l2"))
    got = ir.cpu().tolist()
    recall = sum(len(truth[j] & set(got[j])) for j in range(Q)) / (Q * K)
    adc_only = cand[:, :K].cpu().tolist()
    recall_adc = sum(len(truth[j] & set(adc_only[j])) for j in range(Q)) / (Q * K)
    alg = float(Q) * N * M_SUB
    rows.append({"R": R, "recall_at_10_after_rerank": round(recall, 4), "recall_at_10_adc_only": round(recall_adc, 4),
                 "adc_ms": round(adc_ms, 3), "adc_kernel_ms": round(kms / max(nk, 1), 3), "rerank_ms": round(rr_ms, 3),
                 "qps_adc_plus_rerank": round(Q / ((lut_ms + adc_ms + rr_ms) * 1e-3), 1),
                 "adc_frac_of_hbm_roofline": round(alg / (kms / max(nk, 1) * 1e-3) / 1e9 / HBM, 3)})
print(json.dumps({"path": "pq_recall", "data": DATA, "N": N, "D": D, "m": M_SUB, "Q": Q, "lut_ms": round(lut_ms, 3),
                  "rows": rows}), flush=True)
