"""GPU: the fused shard exchange + merge kernel (csrc/exchange.cu) against vdb_merge_topk.

One GPU is enough for the protocol: R "ranks" are R streams of the same device, each with its own exchange
buffer (plain device memory -- on one device every buffer is trivially peer-visible) and its own
PeerExchange; the R kernels run concurrently and wait on one another through the arrival counters exactly
as R processes would.  The multi-process form (torch symmetric memory over NVLink) is exercised by
scripts/dev_exchange.py under ``gpurun --gpus 2`` and by ``bench.py --gpus N``.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lists(rng, world, Q, k, dev, short_from=None):
    """Per-shard sorted lists with global ids (shard r owns ids [r*1000, (r+1)*1000)) and some ties."""
    ds, ids = [], []
    for r in range(world):
        d = np.sort(np.round(rng.random((Q, k)).astype(np.float32), 2), axis=1)       # rounding -> cross-shard ties
        i = np.sort(rng.integers(0, 1000, size=(Q, k)), axis=1).astype(np.int64) + 1000 * r
        if short_from is not None and r >= short_from:                                 # a shard with fewer than k rows
            d[:, k // 2:] = np.inf
            i[:, k // 2:] = -1
        ds.append(torch.from_numpy(d).to(dev))
        ids.append(torch.from_numpy(i).to(dev))
    return ds, ids


@pytest.mark.parametrize("world,Q,k,descending", [(2, 64, 10, False), (4, 100, 10, True), (8, 37, 16, False),
                                                  (3, 1, 1, False)])
def test_exchange_merge_equals_gather_then_merge(world, Q, k, descending):
    from vector_db_b200 import ops
    from vector_db_b200.sharded import PeerExchange
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(world * 1000 + Q)
    cap = Q * k + 5
    bufs = [torch.zeros((ops.exchange_bytes(world, cap),), dtype=torch.uint8, device=dev) for _ in range(world)]
    ex = [PeerExchange(cap, dev, rank=r, world=world, buffers=bufs) for r in range(world)]
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    torch.cuda.synchronize()
    for call in range(5):                                   # several calls: both parities, counters keep running
        ds, ids = _lists(rng, world, Q, k, dev, short_from=(world - 1 if call == 2 else None))
        if descending:                                      # similarity lists: largest first
            ds = [torch.flip(torch.where(torch.isinf(d), -d, d), dims=[1]).contiguous() for d in ds]
            ids = [torch.flip(i, dims=[1]).contiguous() for i in ids]
        want_d, want_i = ops.merge_topk(torch.stack(ds), torch.stack(ids), descending=descending)
        torch.cuda.synchronize()
        outs = []
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                outs.append(ex[r].exchange_merge(ds[r], ids[r], descending=descending))
        torch.cuda.synchronize()
        for r in range(world):
            ex[r].check()
            np.testing.assert_array_equal(outs[r][1].cpu().numpy(), want_i.cpu().numpy())
            np.testing.assert_array_equal(outs[r][0].cpu().numpy(), want_d.cpu().numpy())


def test_exchange_reports_a_missing_peer_instead_of_hanging():
    from vector_db_b200 import ops
    from vector_db_b200.sharded import PeerExchange
    dev = torch.device("cuda", 0)
    world, Q, k = 2, 8, 4
    bufs = [torch.zeros((ops.exchange_bytes(world, Q * k),), dtype=torch.uint8, device=dev) for _ in range(world)]
    ex0 = PeerExchange(Q * k, dev, rank=0, world=world, buffers=bufs)
    d = torch.zeros((Q, k), device=dev)
    i = torch.zeros((Q, k), dtype=torch.int64, device=dev)
    out_d, out_i = ex0.exchange_merge(d, i)                 # rank 1 never calls
    with pytest.raises(RuntimeError, match="never arrived"):
        ex0.check()
    # a timed-out exchange must not return a merge of stale slots: (inf, -1) instead
    assert torch.isinf(out_d).all() and (out_i == -1).all()


def test_exchange_rejects_bad_arguments():
    from vector_db_b200 import ops
    from vector_db_b200.sharded import PeerExchange
    dev = torch.device("cuda", 0)
    buf = [torch.zeros((ops.exchange_bytes(1, 16),), dtype=torch.uint8, device=dev)]
    ex = PeerExchange(16, dev, rank=0, world=1, buffers=buf)
    with pytest.raises(ValueError, match="exceeds the exchange capacity"):
        ex.exchange_merge(torch.zeros((8, 4), device=dev), torch.zeros((8, 4), dtype=torch.int64, device=dev))  # 32 > cap
    assert ops.exchange_bytes(17, 16) == 0 and ops.exchange_bytes(0, 16) == 0
    # the rejected call must not have advanced the arrival counters: the next valid call completes at once
    assert ex.calls == 0 and ex.sent == [0, 0]
    d = torch.rand((4, 4), device=dev).sort(dim=1).values
    i = torch.arange(16, dtype=torch.int64, device=dev).view(4, 4)
    od, oi = ex.exchange_merge(d, i)
    ex.check()
    assert torch.equal(od, d) and torch.equal(oi, i)
