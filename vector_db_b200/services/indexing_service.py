"""IndexingService with the reference's interface (src/vector_db/services/indexing_service.py:14-144):
reads ``index.{M, ef_construction, flush_threshold}`` (and, as a superset, an optional
``index.ef_search`` default) from the YAML config, owns an ``HNSW(rng=random.Random(42), ...)`` and
forwards ``insert_node / delete_node / search / save_index``.  The index it owns is the GPU one."""
from __future__ import annotations

import random
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np
import yaml

from ..indexing.hnsw import HNSW
from ..types import Node


class IndexingService:
    def __init__(self, storage, config_path: str, index_file: Optional[str] = None):
        self.storage = storage
        self.config_path = Path(config_path)
        with open(self.config_path, "r") as f:
            config = yaml.safe_load(f) or {}
        index_config = config.get("index", {}) or {}
        M = index_config.get("M", 16)
        ef_construction = index_config.get("ef_construction", 200)
        self.ef_search = index_config.get("ef_search")          # optional default for search(ef=...)
        if index_file:
            self.index_file = Path(index_file)
        else:
            base = Path(storage.embedding_file).parent / Path(storage.embedding_file).stem.replace(".embeddings", "")
            self.index_file = base.with_suffix(".index.pkl")
        self._index_loaded = self.index_file.exists()
        self.index = HNSW(M=M, ef_construction=ef_construction, rng=random.Random(42), storage=storage,
                          index_file=None)
        # Persistence policy: the reference pickles the whole graph on EVERY insert because its HNSW
        # owns index_file (hnsw.py:224-225), which makes flush_threshold moot; here the service owns
        # the file and flush_threshold decides (indexing_service.py:137-144).
        self.index.index_file = self.index_file
        if self._index_loaded:
            self.index.load_index()
        self.index.index_file = None
        self._index_modified = False
        self.flush_threshold = index_config.get("flush_threshold", 1000)

    def is_index_loaded(self) -> bool:
        return self._index_loaded

    def insert_node(self, node: Node) -> None:
        self.index.insert_node(node)
        self._index_modified = True
        if self._should_flush():
            self.save_index()

    def insert_nodes(self, nodes: List[Node]) -> None:
        """Batched insert (one kernel launch per 2048 nodes)."""
        self.index.build_index(nodes)
        self._index_modified = True
        if self._should_flush():
            self.save_index()

    def delete_node(self, node_id: int) -> None:
        self.index.delete_node(node_id)
        self._index_modified = True

    def search(self, query: np.ndarray, k: int, **kwargs) -> List[Tuple[Node, float]]:
        if "ef" not in kwargs and self.ef_search is not None:
            kwargs["ef"] = self.ef_search
        return self.index.search(query, k=k, **kwargs)

    def _write(self) -> None:
        self.index.index_file = self.index_file
        try:
            self.index.save_index()
        finally:
            self.index.index_file = None

    def save_index(self) -> None:
        if self._index_modified:
            self._write()
            self._index_modified = False

    def force_save_index(self) -> None:
        self._write()
        self._index_modified = False

    def get_index_size(self) -> int:
        return len(self.index.node_store)

    def _should_flush(self) -> bool:
        return self.get_index_size() >= self.flush_threshold
