"""CPU-side checks: the C-ABI library loads and exports every symbol include/vdb_b200.h declares,
the ctypes table mirrors the header, and the host-side storage logic keeps the reference's formats."""
import ctypes
import os
import re

import numpy as np
import pytest

import vector_db_b200
from vector_db_b200 import _lib
from vector_db_b200.storage import InMemoryNodeStorage, MMapNodeStorage
from vector_db_b200.types import Node

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "vdb_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vdb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from vector_db_b200.build import build
    path = build()
    lib = ctypes.CDLL(path)
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in vdb_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes table out of step with the header"
    lib.vdb_abi_version.restype = ctypes.c_int
    assert lib.vdb_abi_version() == 4


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.VdbError):
        _lib.require_cuda()
    from vector_db_b200.indexing import ExactIndex
    with pytest.raises(_lib.VdbError):
        ExactIndex(8)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "vector_db_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_mmap_storage_layout_and_roundtrip(tmp_path):
    dim = 6
    st = MMapNodeStorage(tmp_path / "db.embeddings.npy", tmp_path / "db.metadata.npy", dim=dim, capacity=32)
    assert st.embeddings.dtype.itemsize == 8 + 4 * dim            # storage.py:135-144 record stride
    rng = np.random.default_rng(0)
    vecs = rng.standard_normal((5, dim)).astype(np.float32)
    for i in range(5):
        st.save(Node(id=10 + i, embedding=vecs[i], content=f"c{i}", metadata={"k": i}))
    assert st.size() == 5 and st.get_next_id() == 15
    np.testing.assert_array_equal(st.get_embedding(12), vecs[2])
    st.delete(11)
    assert st.get(11) is None and st.size() == 4
    with pytest.raises(KeyError):
        st.get_embedding(11)
    st.save(Node(id=99, embedding=vecs[0]))                        # reuses the lowest free slot (1)
    assert st._id_to_index[99] == 1
    st.close()
    # raw bytes: id at offset 0, embedding at offset 8 of each record
    raw = np.fromfile(tmp_path / "db.embeddings.npy", dtype=np.uint8).reshape(32, 8 + 4 * dim)
    assert raw[0, :8].view(np.int64)[0] == 10
    np.testing.assert_array_equal(raw[0, 8:].view(np.float32), vecs[0])
    st2 = MMapNodeStorage(tmp_path / "db.embeddings.npy", tmp_path / "db.metadata.npy", dim=dim, capacity=32)
    assert sorted(st2.get_all_ids()) == [10, 12, 13, 14, 99] and st2.size() == 5
    n = st2.get(13)
    assert n.content == "c3" and n.metadata == {"k": 3}
    with pytest.raises(RuntimeError):
        small = MMapNodeStorage(tmp_path / "s.e", tmp_path / "s.m", dim=2, capacity=1)
        small.save(Node(id=0, embedding=np.ones(2, np.float32)))
        small.save(Node(id=1, embedding=np.ones(2, np.float32)))


def test_in_memory_storage_contract():
    st = InMemoryNodeStorage()
    st.add(Node(id=4, embedding=np.ones(3, np.float32)))
    assert st.size() == 1 and st.get_next_id() == 5 and st.get(7) is None
    with pytest.raises(KeyError):
        st.get_embedding(7)


def test_pq_service_validation_messages():
    from vector_db_b200.inference.pq import ProductQuantizationService as PQ
    for kw, msg in (((0, 2, 4), "k must be greater than 0"), ((2, 0, 4), "chunks must be greater than 0"),
                    ((2, 2, 0), "dim must be greater than 0"), ((2, 3, 4), "dim must be divisible by chunks")):
        with pytest.raises(ValueError, match=msg):
            PQ(*kw)
    svc = PQ(k=4, chunks=2, dim=4)
    with pytest.raises(TypeError):
        svc._validate_embeddings([[1, 2, 3, 4]])
    with pytest.raises(ValueError, match="2D"):
        svc._validate_embeddings(np.zeros(4, np.float32))
    with pytest.raises(ValueError, match="dimension must be 4"):
        svc._validate_embeddings(np.zeros((2, 5), np.float32))
    ch = svc._chunk_embeddings(np.arange(8, dtype=np.float32).reshape(2, 4))
    assert len(ch) == 2 and len(ch[0]) == 2
    np.testing.assert_array_equal(ch[1][0], [2, 3])
    np.testing.assert_array_equal(ch[0][1], [4, 5])
