#!/usr/bin/env python
"""Batched HNSW search over a row-sharded database: every rank owns ONE complete HNSW graph over
its rows (BASELINE.json configs[3] shape per GPU: N x 768, M = 32, ef = 128), every query searches all R graphs,
the per-shard top-k lists are exchanged and merged (fused peer-memory kernel, or NCCL all-gather + merge).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node R --master-addr 127.0.0.1 --master-port P \
        scripts/bench_sharded_hnsw.py [rows_per_gpu=1000000] [Q=10000] [steps=5]

Weak scaling: rows per GPU fixed, the index grows with R; value = queries/s over the whole (R x rows) index.
"""
import json
import os
import random
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops, sharded  # noqa: E402
from vector_db_b200.indexing import HNSW  # noqa: E402

N_LOCAL = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 5
D, M, EF, K = 768, 32, 128, 10
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
HBM = peaks.get("hbm_gbs", 6650.0)

lo = rank * N_LOCAL                                              # global ids: contiguous range per shard
rng = np.random.default_rng(1234 + rank)
X = rng.standard_normal((N_LOCAL, D), dtype=np.float32)
g = HNSW(M=M, ef_construction=200, rng=random.Random(42 + rank), device=dev)
t0 = time.perf_counter()
g.build_from_arrays(np.arange(lo, lo + N_LOCAL), X)
torch.cuda.synchronize(dev)
build_s = time.perf_counter() - t0
del X
Qd = torch.from_numpy(np.random.default_rng(4321).standard_normal((Q, D), dtype=np.float32)).to(dev)
vec = g._vectors
entry_row, top = g._read_state()


def local_search(q, k):
    d, r, _ = ops.hnsw_search(q, vec.vecs, vec.n, g._adj0, g._adj_up, g._up_ptr_dev, entry_row, top, EF, k,
                              alive=None, flt=None, want_stats=False)
    return d, torch.where(r >= 0, vec.ids[r.clamp(min=0).long()], torch.full_like(r, -1, dtype=torch.int64))


peer = None
if world > 1:
    try:
        peer = sharded.PeerExchange(Q * K, dev)
    except Exception as e:                                       # noqa: BLE001
        print(f"peer exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
    flag = torch.tensor([1 if peer is not None else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 0:
        peer = None


def step():
    d, i = local_search(Qd, K)
    if world == 1:
        return d, i
    if peer is not None:
        return peer.exchange_merge(d, i)
    gd, gi = sharded.gather_topk(d, i)
    return sharded.merge_shards(gd, gi)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


for _ in range(3):
    out = step()
barrier()
_lib.profile_enable(True)
_lib.profile_read("hnsw_search")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
barrier()
e0.record()
for _ in range(STEPS):
    out = step()
e1.record()
barrier()
ms = torch.tensor([e0.elapsed_time(e1) / STEPS], device=dev)
nk, kms = _lib.profile_read("hnsw_search")
_lib.profile_enable(False)
kern = torch.tensor([kms / max(nk, 1)], device=dev)
bld = torch.tensor([build_s], device=dev)
# traversal statistics of this shard (evaluations / expansions per query) for the byte model
_, _, st = ops.hnsw_search(Qd[:2000], vec.vecs, vec.n, g._adj0, g._adj_up, g._up_ptr_dev, entry_row, top, EF, K,
                           alive=None, flt=None, want_stats=True)
stats = st.float().mean(0)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dist.all_reduce(kern, op=dist.ReduceOp.MAX)
    dist.all_reduce(bld, op=dist.ReduceOp.MAX)
    dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    stats /= world
if peer is not None:
    peer.check()
if rank == 0:
    ev, ex = float(stats[0]), float(stats[1])
    alg = Q * (ev * D * 4 + ex * 2 * M * 4)                       # bytes per GPU and step
    kms_ = float(kern.item())
    print(json.dumps({
        "path": "sharded_hnsw", "n_gpus": world, "rows_per_gpu": N_LOCAL, "rows_total": N_LOCAL * world, "D": D, "M": M,
        "ef": EF, "Q": Q, "k": K, "scaling": "weak", "ms_per_step": float(ms.item()),
        "qps": Q / (float(ms.item()) * 1e-3), "hnsw_kernel_ms_max_over_ranks": kms_,
        "build_s_max_over_ranks": float(bld.item()), "evals_per_query_per_shard": ev, "expansions_per_query_per_shard": ex,
        "exchange": "single GPU" if world == 1 else ("fused peer-memory kernel" if peer is not None else "nccl all-gather + merge"),
        "roofline": {"bound": "hbm", "achieved": alg / (kms_ * 1e-3) / 1e9, "peak": HBM, "unit": "GB/s",
                     "frac": alg / (kms_ * 1e-3) / 1e9 / HBM, "algorithmic_bytes_per_gpu": alg}}), flush=True)
if world > 1:
    dist.destroy_process_group()
