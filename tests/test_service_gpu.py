"""GPU: IndexingService + MMapNodeStorage drop-in behaviour (tests/services/test_indexing_service.py,
tests/inference/test_mmap_vector_store.py restated) and the memmap -> pinned -> HBM staging path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(tmp_path, extra=""):
    p = tmp_path / "config.yaml"
    p.write_text("index:\n  M: 8\n  ef_construction: 50\n  flush_threshold: 20\n" + extra +
                 "vector_db:\n  dimension: 16\n  capacity: 256\n")
    return str(p)


def test_indexing_service_dropin(tmp_path):
    from vector_db_b200.services import IndexingService
    from vector_db_b200.storage import MMapNodeStorage
    from vector_db_b200.types import Node
    st = MMapNodeStorage(tmp_path / "db.embeddings.npy", tmp_path / "db.metadata.npy", dim=16, capacity=256)
    svc = IndexingService(st, _cfg(tmp_path))
    assert not svc.is_index_loaded() and svc.get_index_size() == 0
    assert svc.index.M == 8 and svc.index.ef_construction == 50 and svc.flush_threshold == 20
    rng = np.random.default_rng(0)
    X = rng.random((40, 16), dtype=np.float32)
    for i in range(40):
        svc.insert_node(Node(id=i, embedding=X[i], content=f"doc {i}", metadata={"parity": i % 2}))
    assert svc.get_index_size() == 40
    assert svc.index_file.exists() and not svc._index_modified          # flush_threshold reached -> saved
    res = svc.search(X[5], k=3, ef=50, pq_chunks=4, custom_arg=1)        # extra kwargs ignored (hnsw.py:338)
    assert len(res) == 3 and res[0][0].id == 5 and res[0][1] < 1e-6
    assert res[0][0].content == "doc 5" and res[0][0].metadata == {"parity": 1}
    even = {i for i in range(40) if i % 2 == 0}
    assert all(n.id in even for n, _ in svc.search(X[5], k=5, ef=50, filter_ids=even))
    assert svc.search(X[5], k=5, filter_ids=set()) == []
    svc.delete_node(5)
    assert svc._index_modified and all(n.id != 5 for n, _ in svc.search(X[5], k=5, ef=50))
    svc.force_save_index()
    a = [(n.id, float(d)) for n, d in svc.search(X[7], k=5, ef=64)]
    st.close()
    st2 = MMapNodeStorage(tmp_path / "db.embeddings.npy", tmp_path / "db.metadata.npy", dim=16, capacity=256)
    svc2 = IndexingService(st2, _cfg(tmp_path, "  ef_search: 64\n"))
    assert svc2.is_index_loaded() and svc2.get_index_size() == 39 and svc2.ef_search == 64
    b = [(n.id, float(d)) for n, d in svc2.search(X[7], k=5)]
    assert a == b


def test_stage_memmap_to_hbm(tmp_path):
    from vector_db_b200.storage import MMapNodeStorage
    from vector_db_b200.types import Node
    dim = 24
    st = MMapNodeStorage(tmp_path / "s.embeddings.npy", tmp_path / "s.metadata.npy", dim=dim, capacity=300)
    rng = np.random.default_rng(1)
    X = rng.standard_normal((200, dim)).astype(np.float32)
    # write layer 1 directly (vectorised) the way a bulk loader would, then reopen to index it
    st.embeddings["id"][:200] = np.arange(200) * 7
    st.embeddings["embedding"][:200] = X
    st.embeddings.flush()
    st.close()
    st = MMapNodeStorage(tmp_path / "s.embeddings.npy", tmp_path / "s.metadata.npy", dim=dim, capacity=300)
    assert st.size() == 200
    st.delete(21)                                   # slot 3 becomes a hole (id 0, zero vector)
    ids, vecs = st.stage_to_device(chunk_records=64)
    keep = np.array([i for i in range(200) if i != 3])
    np.testing.assert_array_equal(ids.cpu().numpy(), keep * 7)
    np.testing.assert_array_equal(vecs.cpu().numpy(), X[keep])
    from vector_db_b200.indexing import ExactIndex
    idx = ExactIndex(dim, engine="f32")
    idx.add_device(ids, vecs)
    D, I = idx.search(X[10:12], 1)
    assert I[:, 0].tolist() == [70, 77] and np.all(D[:, 0] == 0)


def test_concurrent_search_and_insert_from_threads():
    """The reference's callers are Starlette threadpool threads with no locks (api/app.py:120,166).  Here:
    one thread inserts 300 nodes in order while four threads search; nothing raises, every hit is a node that
    exists, and the graph ends up equal to the one a single-threaded build of the same insertion order gives."""
    import random
    import threading
    from vector_db_b200.indexing import HNSW
    from vector_db_b200.types import Node
    rng = np.random.default_rng(3)
    X = rng.standard_normal((300, 32)).astype(np.float32)
    g = HNSW(M=8, ef_construction=40, rng=random.Random(42))
    for i in range(20):
        g.insert_node(Node(id=i, embedding=X[i]))
    errors, hits = [], []
    stop = threading.Event()

    def searcher(seed):
        r = np.random.default_rng(seed)
        try:
            while not stop.is_set():
                q = X[int(r.integers(0, 20))]
                res = g.search(q, k=5, ef=30)
                assert 1 <= len(res) <= 5 and res[0][1] < 1e-6          # a stored row finds itself first
                assert all(0 <= n.id < 300 for n, _ in res)
                hits.append(len(res))
        except Exception as e:                                            # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=searcher, args=(s,)) for s in range(4)]
    for t in threads:
        t.start()
    try:
        for i in range(20, 300):
            g.insert_node(Node(id=i, embedding=X[i]))
    finally:
        stop.set()
        for t in threads:
            t.join(timeout=60)
    assert not errors, errors[:1]
    assert len(hits) > 0
    ref = HNSW(M=8, ef_construction=40, rng=random.Random(42))
    for i in range(300):
        ref.insert_node(Node(id=i, embedding=X[i]))
    assert g.entry_node_id == ref.entry_node_id and g.max_level == ref.max_level
    for i in range(300):
        assert g.node_store[i].neighbors == ref.node_store[i].neighbors


def test_memory_mapping_service_dropin(tmp_path):
    """tests/inference/test_mmap_vector_store.py restated against the GPU-backed facade: constructor validation,
    write / read / get_embedding / delete / search contracts and messages, persistence across a reopen."""
    from vector_db_b200.inference import MemoryMappingService
    cfg = tmp_path / "config.yaml"
    cfg.write_text("index:\n  M: 4\n  ef_construction: 10\n")
    base = str(tmp_path / "vectors")
    with pytest.raises(ValueError, match="Dimension must be greater than 0"):
        MemoryMappingService(base, dim=0, capacity=10, config_path=str(cfg))
    with pytest.raises(ValueError, match="Capacity must be greater than 0"):
        MemoryMappingService(base, dim=4, capacity=0, config_path=str(cfg))
    with pytest.raises(ValueError, match="config_path is required"):
        MemoryMappingService(base, dim=4, capacity=10)
    svc = MemoryMappingService(base, dim=4, capacity=10, config_path=str(cfg), index_file=str(tmp_path / "index.pkl"))
    assert svc.size == 0 and svc.dim == 4 and svc.capacity == 10 and svc.index.M == 4 and svc.index.ef_construction == 10
    with pytest.raises(TypeError, match="Embedding must be a numpy array"):
        svc.write([1.0, 2.0, 3.0, 4.0])
    with pytest.raises(ValueError, match="Embedding must be a 1D array"):
        svc.write(np.zeros((2, 2), dtype=np.float32))
    with pytest.raises(ValueError, match="Embedding must be of dimension 4"):
        svc.write(np.zeros(3, dtype=np.float32))
    embs = [np.array([1.0, 0.0, 0.0, 0.0]), np.array([0.0, 1.0, 0.0, 0.0], dtype=np.float32),
            np.array([0.0, 0.0, 1.0, 0.0], dtype=np.float32)]
    ids = [svc.write(e, content=f"node_{i}", metadata={"n": i}) for i, e in enumerate(embs)]     # float64 is cast
    assert ids == [0, 1, 2] and svc.size == 3
    node = svc.read(1)
    assert node.id == 1 and node.content == "node_1" and node.metadata == {"n": 1}
    np.testing.assert_array_equal(node.embedding, embs[1])
    np.testing.assert_array_equal(svc.get_embedding(0), embs[0].astype(np.float32))
    with pytest.raises(TypeError, match="Node ID must be an integer"):
        svc.read("0")
    with pytest.raises(IndexError, match="Node 999 not found"):
        svc.read(999)
    res = svc.search(np.array([1.0, 0.0, 0.0, 0.0], dtype=np.float32), k=2, ef=10)
    assert 0 < len(res) <= 2 and res[0][0].id == 0 and res[0][1] < 1e-6
    more = svc.write_many(np.eye(4, dtype=np.float32)[3:4] * 2.0, contents=["batched"])
    assert more == [3] and svc.read(3).content == "batched" and svc.size == 4
    svc.delete(1)
    assert svc.size == 3
    with pytest.raises(IndexError):
        svc.read(1)
    assert all(n.id != 1 for n, _ in svc.search(embs[1].astype(np.float32), k=3, ef=10))
    before = [(n.id, float(d)) for n, d in svc.search(embs[2], k=3, ef=10)]
    svc.storage.close()
    del svc
    svc2 = MemoryMappingService(base, dim=4, capacity=10, config_path=str(cfg), index_file=str(tmp_path / "index.pkl"))
    assert svc2.size == 3 and svc2.read(0).content == "node_0"
    assert [(n.id, float(d)) for n, d in svc2.search(embs[2], k=3, ef=10)] == before
