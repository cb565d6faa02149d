#!/usr/bin/env python
"""BASELINE.json configs[4]: PQ (m=16, nbits=8) index over N x 128 vectors sharded by row across the
GPUs of one box; every rank encodes and scans its own shard, per-shard top-k lists are merged after
ONE packed NCCL all-gather (vector_db_b200/sharded.py).  Strong scaling: N is the TOTAL row count.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node R --master-addr 127.0.0.1 --master-port P \
        scripts/bench_sharded_adc.py [N_total=100000000] [Q=1024] [steps=5]

The shard is generated on the device from seed+rank (never materialised on the host), encoded in
chunks with the tensor-core encoder and dropped; only the uint8 codes stay resident.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops, sharded  # noqa: E402

N_TOTAL = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 5
D, M_SUB, K = 128, 16, 10
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
HBM = peaks.get("hbm_gbs", 6650.0)

lo, hi = sharded.shard_bounds(N_TOTAL, rank, world)
n_local = hi - lo

# codebooks: 10 Lloyd rounds on a 200k-row sample, trained on rank 0 and broadcast (replicated, 131 KB)
g0 = torch.Generator(device=dev).manual_seed(777)
sample = torch.randn((200_000, D), device=dev, generator=g0)
cb = sample[:256].view(256, M_SUB, 8).permute(1, 0, 2).contiguous()
if rank == 0:
    for _ in range(10):
        ops.pq_kmeans_step(sample, cb)
if world > 1:
    dist.broadcast(cb, 0)
del sample

# shard generation + encode, chunked
codes = torch.empty((n_local, M_SUB), dtype=torch.uint8, device=dev)
g = torch.Generator(device=dev).manual_seed(1234 + rank)
CH = 2_000_000
enc0, enc1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
enc_ms = 0.0
for s in range(0, n_local, CH):
    e = min(n_local, s + CH)
    X = torch.randn((e - s, D), device=dev, generator=g)
    enc0.record()
    codes[s:e] = ops.pq_encode(X, cb)
    enc1.record()
    torch.cuda.synchronize()
    enc_ms += enc0.elapsed_time(enc1)
    del X
gq = torch.Generator(device=dev).manual_seed(4321)
Qd = torch.randn((Q, D), device=dev, generator=gq)


def local_search(q, k):
    lut = ops.pq_lut(q, cb)
    d, rows = ops.pq_adc_topk(codes, lut, k)
    return d, torch.where(rows >= 0, rows + lo, rows)          # global ids: contiguous range per shard


def step():
    return sharded.search_sharded(local_search, Qd, K)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


for _ in range(3):
    out = step()
barrier()
_lib.profile_enable(True)
_lib.profile_read("adc_scan")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
barrier()
e0.record()
for _ in range(STEPS):
    out = step()
e1.record()
barrier()
ms = torch.tensor([e0.elapsed_time(e1) / STEPS], device=dev)
nk, kms = _lib.profile_read("adc_scan")
_lib.profile_enable(False)
kern = torch.tensor([kms / max(nk, 1)], device=dev)
enc = torch.tensor([enc_ms], device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dist.all_reduce(kern, op=dist.ReduceOp.MAX)
    dist.all_reduce(enc, op=dist.ReduceOp.MAX)
if rank == 0:
    alg = float(Q) * n_local * M_SUB                            # per GPU
    line = {"path": "sharded_adc", "n_gpus": world, "rows_total": N_TOTAL, "rows_per_gpu": n_local, "Q": Q, "m": M_SUB,
            "k": K, "ms_per_step": float(ms.item()), "qps": Q / (float(ms.item()) * 1e-3), "scaling": "strong",
            "adc_kernel_ms_max_over_ranks": float(kern.item()), "encode_ms_max_over_ranks": float(enc.item()),
            "encode_vectors_per_s_per_gpu": n_local / (float(enc.item()) * 1e-3),
            "allgather_payload_bytes_per_rank": Q * K * 12,
            "roofline_per_gpu": {"bound": "hbm", "achieved": alg / (float(kern.item()) * 1e-3) / 1e9, "peak": HBM,
                                 "unit": "GB/s", "frac": alg / (float(kern.item()) * 1e-3) / 1e9 / HBM,
                                 "algorithmic_bytes": alg},
            "first_result": [float(out[0][0, 0]), int(out[1][0, 0])]}
    print(json.dumps(line), flush=True)
if world > 1:
    dist.destroy_process_group()
