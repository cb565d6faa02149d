#!/usr/bin/env python
"""Per-path measurements at the BASELINE.json config sizes that are NOT the headline bench line
(bench.py covers configs[1]): PQ encode + ADC scan (configs[2]), batched HNSW search (configs[3]).
Each prints one JSON line with the kernel's CUDA-event time and its roofline fraction against
MEASURED_PEAKS.json (algorithmic bytes as defined in DESIGN.md).

    python scripts/bench_paths.py adc  [N] [Q] [m] [R]     # default 10_000_000 1024 48 10
    python scripts/bench_paths.py encode [N] [m]           # default 10_000_000 48   (D = 8 m)
    python scripts/bench_paths.py hnsw [N] [D] [M] [ef] [Q] # default 200_000 768 32 128 10_000
    python scripts/bench_paths.py ivf [N] [D] [K] [Q] [iters] # default 1_000_000 384 1024 4096 10 (clustered rows)
"""
import json
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
HBM = PEAKS["hbm_gbs"]


def _time_slot(slot, fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    _lib.profile_read(slot)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    n, ms = _lib.profile_read(slot)
    _lib.profile_enable(False)
    return out, ms / max(n, 1), e0.elapsed_time(e1) / iters


def bench_adc(N=10_000_000, Q=1024, m=48, R=10):
    g = torch.Generator(device=dev).manual_seed(7)
    codes = torch.randint(0, 256, (N, m), device=dev, dtype=torch.uint8, generator=g)
    lut = torch.rand((Q, m, 256), device=dev, generator=g)
    (d, i), k_ms, step_ms = _time_slot("adc_scan", lambda: ops.pq_adc_topk(codes, lut, R), iters=3, warm=1)
    alg_bytes = float(Q) * N * m
    line = {"path": "adc_scan", "N": N, "Q": Q, "m": m, "R": R, "kernel_ms": k_ms, "step_ms": step_ms,
            "qps": Q / (step_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                         "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / HBM,
                         "algorithmic_bytes": alg_bytes, "model": "Q*N*m code bytes (per-query code stream)"},
            "lookups_per_s": alg_bytes / (k_ms * 1e-3)}
    # spot parity on 2 queries against a torch gather restatement (fp32, j ascending)
    acc = torch.zeros((2, N), device=dev)
    for j in range(m):
        acc += lut[:2, j, :].gather(1, codes[:, j].long().unsqueeze(0).expand(2, N))
    ref = torch.topk(acc, R, dim=1, largest=False)
    line["spot_check_ids_equal"] = bool(torch.equal(torch.sort(ref.indices, 1).values, torch.sort(i[:2], 1).values))
    print(json.dumps(line), flush=True)


def bench_encode(N=10_000_000, m=48):
    D = 8 * m
    g = torch.Generator(device=dev).manual_seed(11)
    cb = torch.randn((m, 256, 8), device=dev, generator=g)
    chunk = min(N, 2_000_000)                     # 2M x 384 fp32 = 3 GB per call; loop to cover N
    X = torch.randn((chunk, D), device=dev, generator=g)
    engine = os.environ.get("VDB_ENCODE_ENGINE", "auto")
    codes, k_ms, step_ms = _time_slot("pq_encode", lambda: ops.pq_encode(X, cb, engine=engine), iters=3, warm=1)
    alg_bytes = float(chunk) * D * 4 + float(chunk) * m
    line = {"path": "pq_encode", "engine": engine, "rows_per_call": chunk, "D": D, "m": m, "kernel_ms": k_ms,
            "vectors_per_s": chunk / (k_ms * 1e-3), "ms_for_N": k_ms * N / chunk, "N": N,
            "roofline": {"bound": "hbm", "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                         "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / HBM, "algorithmic_bytes": alg_bytes}}
    ns = 20000
    line["codes_equal_fp32_kernel_on_sample"] = bool(torch.equal(codes[:ns], ops.pq_encode(X[:ns], cb, engine="f32")))
    print(json.dumps(line), flush=True)


def bench_hnsw(N=200_000, D=768, M=32, ef=128, Q=10_000):
    from vector_db_b200.indexing import HNSW
    from vector_db_b200.types import Node
    rng = np.random.default_rng(1234)
    X = rng.standard_normal((N, D), dtype=np.float32)
    Qh = np.random.default_rng(4321).standard_normal((Q, D), dtype=np.float32)
    idx = HNSW(M=M, ef_construction=200, rng=random.Random(42))
    builder = os.environ.get("VDB_HNSW_BUILDER", "wave" if N > 20000 else "sequential")
    t0 = time.perf_counter()
    if builder == "wave":
        stats = idx.build_from_arrays(range(N), X)
    else:
        idx.build_index([Node(id=i, embedding=X[i]) for i in range(N)])
        stats = {}
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    variants = {}
    for lim in [int(v) for v in os.environ.get("VDB_HNSW_HASH_LIMITS", "").split(",") if v]:
        _lib.load().vdb_debug_set_hnsw_smem_hash_limit(lim)
        _, vk, _ = _time_slot("hnsw_search", lambda: idx.search_batch(Qh, 10, ef=ef, return_stats=True), iters=3, warm=1)
        variants[f"smem_hash_limit_{lim}"] = vk
    _lib.load().vdb_debug_set_hnsw_smem_hash_limit(48 * 1024)
    (d, i, st), k_ms, step_ms = _time_slot(
        "hnsw_search", lambda: idx.search_batch(Qh, 10, ef=ef, return_stats=True), iters=3, warm=1)
    evals, expans = float(st[:, 0].mean()), float(st[:, 1].mean())
    alg_bytes = Q * (evals * D * 4 + expans * 2 * M * 4)
    # recall against the exact scan on a sample
    from vector_db_b200.indexing import ExactIndex
    ex = ExactIndex(D, "l2")
    ex.add_device(torch.arange(N, device=dev), torch.from_numpy(X).to(dev))
    n_rec = min(512, Q)
    _, Iex = ex.search(Qh[:n_rec], 10)
    recall = float(np.mean([len(set(i[j]) & set(Iex[j])) / 10 for j in range(n_rec)]))
    # the reference's own search (oracle restatement of hnsw.py:330-374, pinned to the reference by the golden
    # fixtures) on the SAME graph, for a sample: recall parity gate (within 0.5 pt) and the CPU baseline
    from oracle.hnsw_oracle import search_arrays
    adj0, adj_up, up_ptr, _, _ = idx._host_graph()
    entry_row, top = idx._read_state()
    n_or = min(int(os.environ.get("VDB_HNSW_ORACLE_QUERIES", 64)), n_rec)
    t0 = time.perf_counter()
    same, rec_or, rec_gpu, ev_same = 0, 0.0, 0.0, 0
    for j in range(n_or):
        res, ev = search_arrays(X, adj0, up_ptr, adj_up, entry_row, top, Qh[j], 10, ef)
        ids_or = [r for _, r in res]
        same += int(ids_or == i[j].tolist())
        ev_same += int(ev == int(st[j, 0]))
        rec_or += len(set(ids_or) & set(Iex[j])) / 10
        rec_gpu += len(set(i[j]) & set(Iex[j])) / 10
    or_s = time.perf_counter() - t0
    oracle = {"queries": n_or, "identical_result_lists": same, "identical_eval_counts": ev_same,
              "recall_reference_search": rec_or / max(n_or, 1), "recall_gpu_search": rec_gpu / max(n_or, 1),
              "reference_search_qps_1core": n_or / or_s if n_or else None}
    line = {"path": "hnsw_search", "N": N, "D": D, "M": M, "ef": ef, "Q": Q, "builder": builder, "build_stats": stats,
            "vs_reference_search_same_graph": oracle,
            "build_s": build_s, "kernel_ms_variants": variants,
            "inserts_per_s": N / build_s, "kernel_ms": k_ms, "step_ms": step_ms, "qps": Q / (step_ms * 1e-3),
            "evals_per_query": evals, "expansions_per_query": expans, "recall_at_10_vs_exact": recall,
            "roofline": {"bound": "hbm", "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                         "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / HBM, "algorithmic_bytes": alg_bytes,
                         "model": "evals*D*4 + expansions*2M*4 per query, counted by the kernel"}}
    print(json.dumps(line), flush=True)


def bench_ivf(N=1_000_000, D=384, K=1024, Q=4096, iters=10):
    """IVF on clustered rows: train, then batched search at several n_probe with recall@10
    against the exact scan; roofline = bytes of the probed lists' rows / list-scan kernel time."""
    from vector_db_b200.indexing import ExactIndex, IvfIndex, IvfPqIndex
    g = torch.Generator(device=dev).manual_seed(1234)
    centres = torch.randn((1024, D), device=dev, generator=g)
    sigma = 0.15 * centres.norm(dim=1).mean() / (D ** 0.5)
    X = centres[torch.randint(0, 1024, (N,), device=dev, generator=g)] + sigma * torch.randn((N, D), device=dev, generator=g)
    Qd = centres[torch.randint(0, 1024, (Q,), device=dev, generator=g)] + sigma * torch.randn((Q, D), device=dev, generator=g)
    idx = IvfPqIndex(k=K, chunks=D // 8)
    Xh = X.cpu().numpy()
    t0 = time.perf_counter()
    idx.build_from_arrays(range(N), Xh, iters=iters, init_rows=np.random.default_rng(777).choice(N, K, replace=False))
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    ex = ExactIndex(D, "l2")
    ex.add_device(torch.arange(N, device=dev), X)
    _, Iex = ex.search_device(Qd, 10)
    truth = [set(r) for r in Iex.cpu().tolist()]
    sizes = [len(l) for l in idx.inverted_lists]
    rows = []

    def ev_time(fn, iters=5):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return out, e0.elapsed_time(e1) / iters

    cd, cdist_ms = ev_time(lambda: ops.ivf_centroid_dist(Qd, idx._centroids_dev))
    _, probe_ms = ev_time(lambda: ops.ivf_probe(cd, 16, idx._csr()[0]))
    for n_probe in (1, 4, 16, 64):
        if n_probe > K:
            break
        (d, i, scanned), k_ms, step_ms = _time_slot("ivf_scan", lambda: IvfIndex.search_device(idx, Qd, n_probe, 10), iters=5, warm=2)
        got = i.cpu().tolist()
        recall = sum(len(truth[j] & set(got[j])) for j in range(Q)) / (Q * 10)
        alg = float(scanned.sum().item()) * (D * 4 + 4)
        rows.append({"n_probe": n_probe, "recall_at_10": round(recall, 4), "scan_kernel_ms": round(k_ms, 3),
                     "step_ms": round(step_ms, 3), "qps": round(Q / (step_ms * 1e-3), 1),
                     "rows_scanned_per_query": float(scanned.float().mean().item()),
                     "roofline": {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                                  "frac": alg / (k_ms * 1e-3) / 1e9 / HBM, "algorithmic_bytes": alg}})
    # IVF-PQ on the same lists: m = D/8 sub-spaces, ADC shortlist + exact rerank
    t0 = time.perf_counter()
    idx.train_pq(iters=25)
    idx._list_codes()
    torch.cuda.synchronize()
    pq_train_s = time.perf_counter() - t0
    pq_rows = []
    for n_probe, R in ((4, 100), (4, 1000), (16, 1000), (64, 1000)):
        if n_probe > K:
            continue
        (d, i, scanned), k_ms, step_ms = _time_slot("adc_scan", lambda: idx.search_device(Qd, n_probe, 10, shortlist=R), iters=5, warm=2)
        got = i.cpu().tolist()
        recall = sum(len(truth[j] & set(got[j])) for j in range(Q)) / (Q * 10)
        alg = float(scanned.sum().item()) * (D // 8)
        pq_rows.append({"n_probe": n_probe, "shortlist": R, "recall_at_10": round(recall, 4), "adc_kernel_ms": round(k_ms, 3),
                        "step_ms": round(step_ms, 3), "qps": round(Q / (step_ms * 1e-3), 1),
                        "rows_scanned_per_query": float(scanned.float().mean().item()),
                        "roofline": {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                                     "frac": alg / (k_ms * 1e-3) / 1e9 / HBM, "algorithmic_bytes": alg}})
    print(json.dumps({"path": "ivf_search", "N": N, "D": D, "K": K, "Q": Q, "train_iters": iters, "build_s": build_s,
                      "pq_train_encode_s": pq_train_s, "ivf_pq_rows": pq_rows,
                      "centroid_dist_ms": cdist_ms, "probe16_ms": probe_ms,
                      "list_size_min_mean_max": [min(sizes), sum(sizes) / len(sizes), max(sizes)], "rows": rows}), flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "adc"
    args = [int(a) for a in sys.argv[2:]]
    {"adc": bench_adc, "encode": bench_encode, "hnsw": bench_hnsw, "ivf": bench_ivf}[what](*args)
