"""Import shim that runs the REFERENCE's own hot-path test files against this repo's drop-in.

TEST INFRASTRUCTURE.  ``oracle/fetch_ref_suite.py`` (run in the build container, where /root/reference exists) puts a copy
of the reference's ``src/vector_db`` package and of its hot-path test files under ``oracle/_ref/ref_suite/`` (git-ignored,
travels to the GPU box with the snapshot -- nothing of the reference enters the history) next to a two-line conftest
that calls :func:`install`.  ``install`` builds the package a reference maintainer would end up with after following
INTEGRATION.md: the reference's ``vector_db`` with the hot-path modules REPLACED by ``vector_db_b200``'s, so
``from vector_db.indexing.hnsw import HNSW`` in the reference's tests resolves to the CUDA-backed class.

Modes (env VDB_REF_SHIM_MODE):
  full     every module this repo mirrors is swapped in (types, storage, util, indexing, inference.pq,
           inference.mmap_vector_store, services.indexing_service); the rest (services.storage_service, ...) stays
           the reference's own code and is injected duck-typed -- the default;
  hotpath  only the compute modules are swapped (indexing.hnsw, indexing.ivf, inference.pq, util.distance, util.math,
           services.indexing_service, inference.mmap_vector_store); ``vector_db.types`` / ``vector_db.storage`` stay the
           reference's, so our classes run over the reference's Node and NodeStorage objects.
"""
import importlib
import os
import sys

SWAP_COMPUTE = ["util.distance", "util.math", "indexing.hnsw", "indexing.ivf", "inference.pq",
                "inference.mmap_vector_store", "services.indexing_service"]
SWAP_DATA = ["types", "storage"]


def install(ref_src: str, repo_root: str) -> dict:
    mode = os.environ.get("VDB_REF_SHIM_MODE", "full")
    for p in (repo_root, ref_src):
        if p not in sys.path:
            sys.path.insert(0, p)
    swapped = {}
    names = (SWAP_DATA if mode == "full" else []) + SWAP_COMPUTE
    # data modules first: the reference's remaining modules (services.storage_service) import them by name
    for name in names:
        ours = importlib.import_module(f"vector_db_b200.{name}")
        sys.modules[f"vector_db.{name}"] = ours
        swapped[name] = ours.__name__
    import vector_db                                        # the reference's package object (its __init__ is tolerant)
    for name in names:
        parent, _, leaf = f"vector_db.{name}".rpartition(".")
        try:
            setattr(importlib.import_module(parent), leaf, sys.modules[f"vector_db.{name}"])
        except Exception:                                   # noqa: BLE001 -- a parent __init__ that needs absent extras
            pass
    return {"mode": mode, "swapped": swapped}
