"""world_size-2 gloo test (CPU) of the multi-GPU host logic: shard bounds, the packed all-gather of
per-shard top-k lists, and that merging them by (value, shard, position) gives the global
(dist, id) top-k.  The merge itself is a CUDA kernel on the product path (vdb_merge_topk, covered by
the -m gpu tests); here the oracle's lexsort plays its role."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O
from vector_db_b200 import sharded


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, d, nq, k, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[n // 2 + 3] = X[5]                                   # duplicate across shards -> (dist, id) tie
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[5]
    lo, hi = sharded.shard_bounds(n, rank, world)
    Dl, Il = O.exact_l2_topk(X[lo:hi], np.arange(lo, hi, dtype=np.int64), Q, k)     # this rank's shard
    gd, gi = sharded.gather_topk(torch.from_numpy(Dl), torch.from_numpy(Il))
    assert gd.shape == (world, nq, k) and gi.dtype == torch.int64
    if rank == 0:
        out["gd"], out["gi"] = gd.numpy().copy(), gi.numpy().copy()
    dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for n, w in [(10, 2), (7, 4), (100_000_000, 8), (3, 8), (0, 2)]:
        spans = [sharded.shard_bounds(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:])) and all(lo <= hi for lo, hi in spans)


def test_pack_unpack_roundtrip():
    d = torch.tensor([[0.5, float("inf")], [1e-30, 3.0]])
    i = torch.tensor([[7, -1], [2**40 + 3, 0]], dtype=torch.int64)
    d2, i2 = sharded.unpack_topk(sharded.pack_topk(d, i))
    assert torch.equal(d, d2) and torch.equal(i, i2)


def test_two_rank_gather_and_merge_equals_global_topk():
    n, d, nq, k, world = 4001, 24, 9, 10, 2
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, n, d, nq, k, out), nprocs=world, join=True)
        gd, gi = out["gd"], out["gi"]
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[n // 2 + 3] = X[5]
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[5]
    Dg, Ig = O.exact_l2_topk(X, np.arange(n, dtype=np.int64), Q, k)
    # merge rule of vdb_merge_topk: (value, shard, position) ascending
    for qi in range(nq):
        vals = gd[:, qi, :].reshape(-1)
        order = np.lexsort((np.arange(vals.size), vals))[:k]
        np.testing.assert_array_equal(gi[:, qi, :].reshape(-1)[order], Ig[qi])
        np.testing.assert_array_equal(vals[order], Dg[qi])
    assert Ig[0, 0] == 5 and Ig[0, 1] == n // 2 + 3
