"""Metadata pre-filter: the reference's StorageService.filter_by_metadata
(src/vector_db/services/storage_service.py:106-128) against the GPU column filter.  The golden id sets in
tests/golden/metadata_filter.json come from the reference itself (oracle/gen_golden.py metadata)."""
import json
import os
import random

import numpy as np
import pytest

from oracle import oracle as O


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "metadata_filter.json")) as fh:
        return json.load(fh)


def test_oracle_restatement_equals_the_reference(golden):
    live = [(i, md) for i, md in enumerate(golden["metadatas"]) if i not in set(golden["deleted"])]
    for flt, want in zip(golden["filters"], golden["expect"]):
        assert sorted(O.filter_by_metadata(live, flt)) == want, flt


@pytest.mark.gpu
def test_gpu_filter_equals_the_reference_id_sets(golden):
    from vector_db_b200.indexing import ExactIndex
    mds = golden["metadatas"]
    rng = np.random.default_rng(0)
    X = rng.standard_normal((len(mds), 16)).astype(np.float32)
    idx = ExactIndex(16, "l2")
    ids = [1000 + 3 * i for i in range(len(mds))]                  # external ids != rows
    idx.add(ids, X, metadata=mds)
    for i in golden["deleted"]:
        idx.store.mark_dead(ids[i])
    for flt, want in zip(golden["filters"], golden["expect"]):
        got = idx.store.bitmap_to_ids(idx.metadata_bitmap(flt))
        assert sorted(got) == [ids[i] for i in want], flt
    # a filter dict, the bitmap and the host-built id set give the same search results
    Q = rng.standard_normal((7, 16)).astype(np.float32)
    flt = {"parity": 1, "flag": False}
    live = [(ids[i], md) for i, md in enumerate(mds) if i not in set(golden["deleted"])]
    id_set = O.filter_by_metadata(live, flt)
    d0, i0 = idx.search(Q, 5, filter_ids=id_set)
    d1, i1 = idx.search(Q, 5, filter_ids=flt)
    d2, i2 = idx.search(Q, 5, filter_ids=idx.metadata_bitmap(flt))
    np.testing.assert_array_equal(i0, i1)
    np.testing.assert_array_equal(i0, i2)
    np.testing.assert_array_equal(d0, d1)
    assert set(i0.ravel().tolist()) <= id_set | {-1}
    # NaN matches nothing; a re-attached metadata dict replaces the old one
    assert idx.store.bitmap_to_ids(idx.metadata_bitmap({"score": float("nan")})) == set()
    idx.store.set_metadata([ids[1]], [{"parity": 7}])
    assert idx.store.bitmap_to_ids(idx.metadata_bitmap({"parity": 7})) == {ids[1]}
    assert ids[1] not in idx.store.bitmap_to_ids(idx.metadata_bitmap({"color": "green"}))


@pytest.mark.gpu
def test_hnsw_metadata_filter_kwarg_equals_filter_ids():
    from vector_db_b200.indexing import HNSW
    from vector_db_b200.types import Node
    rng = np.random.default_rng(4)
    X = rng.standard_normal((500, 24)).astype(np.float32)
    g = HNSW(M=8, ef_construction=40, rng=random.Random(42))
    mds = [{"shard": i % 5, "lang": "en" if i % 3 else "de"} for i in range(500)]
    g.build_index([Node(id=i, embedding=X[i], metadata=mds[i]) for i in range(500)])
    g.delete_node(10)
    flt = {"shard": 0, "lang": "de"}
    want_ids = O.filter_by_metadata([(i, mds[i]) for i in range(500) if i != 10], flt)
    for qi in (3, 77, 240):
        a = [(n.id, float(d)) for n, d in g.search(X[qi], k=8, ef=60, filter_ids=want_ids)]
        b = [(n.id, float(d)) for n, d in g.search(X[qi], k=8, ef=60, metadata_filter=flt)]
        assert a == b and all(i in want_ids for i, _ in b) and len(b) > 0
    # a node re-inserted with new metadata is seen with the new values
    g.insert_node(Node(id=15, embedding=X[15], metadata={"shard": 99}))
    assert [n.id for n, _ in g.search(X[15], k=3, ef=60, metadata_filter={"shard": 99})] == [15]


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 31, 32, 33, 1023, 1024, 1025, 1_000_003])
def test_filter_kernel_bitmap_equals_numpy_at_ragged_sizes(n):
    """vdb_filter_eq through the C ABI on random code columns: every word of the bitmap, including the ragged
    last one, the AND of several keys and the alive mask."""
    import torch
    from vector_db_b200 import _lib
    lib = _lib.load()
    dev = _lib.require_cuda()
    rng = np.random.default_rng(n)
    c1 = rng.integers(0, 4, n).astype(np.int32)
    c2 = rng.integers(0, 3, n).astype(np.int32)
    alive = rng.random(n) < 0.9
    words = (n + 31) // 32

    def pack(bits):
        padded = np.zeros(words * 32, dtype=np.uint8)
        padded[:n] = bits
        return np.packbits(padded, bitorder="little").view(np.uint32)

    t1, t2 = torch.from_numpy(c1).to(dev), torch.from_numpy(c2).to(dev)
    ta = torch.from_numpy(pack(alive).view(np.int32)).to(dev)
    out = torch.full((words,), 0x5A5A5A5A, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.vdb_filter_eq(t1.data_ptr(), n, 2, ta.data_ptr(), out.data_ptr(), 1, st))
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), pack((c1 == 2) & alive))
    _lib.check(lib.vdb_filter_eq(t2.data_ptr(), n, 1, None, out.data_ptr(), 0, st))
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), pack((c1 == 2) & (c2 == 1) & alive))
