#!/usr/bin/env python
"""Small-shape drivers for compute-sanitizer: one kernel family per invocation so each tool's log
names one protocol.   python scripts/sanitize_small.py exact_tc|pq_encode_tc|exchange|hnsw|adc|static
Each family checks its result against the fp32 / reference-order path, so a hazard that changes the answer is seen too."""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_db_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)


def exact_tc():
    for metric, (n, d, q) in (("l2", (20_000, 128, 200)), ("cosine", (12_000, 192, 130))):
        X = torch.randn((n, d), device=dev, generator=g)
        Q = torch.randn((q, d), device=dev, generator=g)
        mir, _, stats = ops.tc_prepare_rows(X, metric)
        D, I = ops.exact_topk_tc(Q, X, mir, stats, 10, metric)
        Dr, Ir = ops.exact_topk_f32(Q, X, 10, metric)
        assert torch.equal(I, Ir) and torch.equal(D, Dr)


def pq_encode_tc():
    X = torch.randn((9_000, 128), device=dev, generator=g)
    cb = torch.randn((16, 256, 8), device=dev, generator=g)
    a = ops.pq_encode(X, cb, engine="tc")
    b = ops.pq_encode(X, cb, engine="f32")
    assert torch.equal(a, b)


def adc():
    codes = torch.randint(0, 256, (50_000, 16), device=dev, dtype=torch.uint8, generator=g)
    lut = torch.rand((9, 16, 256), device=dev, generator=g)
    d, i = ops.pq_adc_topk(codes, lut, 10)
    d2, i2 = ops.pq_adc_topk(codes, lut, 600)
    assert torch.equal(i, i2[:, :10])


def exchange():
    from vector_db_b200.sharded import PeerExchange
    world, Q, k = 4, 64, 10
    cap = Q * k
    bufs = [torch.zeros((ops.exchange_bytes(world, cap),), dtype=torch.uint8, device=dev) for _ in range(world)]
    ex = [PeerExchange(cap, dev, rank=r, world=world, buffers=bufs) for r in range(world)]
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    for _ in range(3):
        ds = [torch.rand((Q, k), device=dev, generator=g).sort(dim=1).values for _ in range(world)]
        ids = [torch.arange(Q * k, device=dev).view(Q, k) + 10_000 * r for r in range(world)]
        want = ops.merge_topk(torch.stack(ds), torch.stack(ids))
        torch.cuda.synchronize()
        outs = []
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                outs.append(ex[r].exchange_merge(ds[r], ids[r]))
        torch.cuda.synchronize()
        for r in range(world):
            ex[r].check()
            assert torch.equal(outs[r][1], want[1])


def hnsw():
    from vector_db_b200.indexing import HNSW
    from vector_db_b200.types import Node
    rng = np.random.default_rng(2)
    X = rng.standard_normal((400, 32)).astype(np.float32)
    idx = HNSW(M=8, ef_construction=40, rng=random.Random(42))
    idx.build_index([Node(id=i, embedding=X[i]) for i in range(400)])
    D, I = idx.search_batch(X[:64], 5, ef=40)
    assert (I[:, 0] == np.arange(64)).all()
    idx2 = HNSW(M=8, ef_construction=40, rng=random.Random(42))
    idx2.build_from_arrays(range(400), X, wave_max=128, seed=64)
    D2, I2 = idx2.search_batch(X[:64], 5, ef=40, filter_ids=set(range(0, 400, 2)))
    assert (I2[::2, 0] == np.arange(0, 64, 2)).all()


def static():
    lib = _lib.load()
    X = torch.randn((5_000, 64), device=dev, generator=g)
    Q = torch.randn((33, 64), device=dev, generator=g)
    od = torch.empty((33, 10), device=dev)
    oi = torch.empty((33, 10), dtype=torch.int64, device=dev)
    for _ in range(2):
        _lib.check(lib.vdb_exact_topk_f32_static(Q.data_ptr(), 33, X.data_ptr(), None, None, 5_000, 64, 10, 0, od.data_ptr(),
                                                 oi.data_ptr(), _lib.stream_ptr(dev)))
    Dr, Ir = ops.exact_topk_f32(Q, X, 10, "l2")
    assert torch.equal(oi, Ir) and torch.equal(od, Dr)


if __name__ == "__main__":
    for name in sys.argv[1:]:
        {"exact_tc": exact_tc, "pq_encode_tc": pq_encode_tc, "exchange": exchange, "hnsw": hnsw, "adc": adc, "static": static}[name]()
        torch.cuda.synchronize()
        print(f"{name}: results ok", flush=True)
