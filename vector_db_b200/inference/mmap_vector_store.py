"""MemoryMappingService: the older storage + index facade of the reference
(src/vector_db/inference/mmap_vector_store.py:12-177) over the two-layer memmap storage and the GPU HNSW.

It is the second caller of the search boundary (mmap_vector_store.py:122 insert, :157 delete, :177 search;
SURVEY.md 8b).  Same constructor, methods, attributes, validation messages and file names
(``<file_path>.embeddings.npy``, ``.metadata.npy``, ``.index.pkl``), so a directory written by the reference's
service opens here and vice versa; the index work runs on the device.
"""
from __future__ import annotations

import random
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import yaml

from ..indexing.hnsw import HNSW
from ..storage import MMapNodeStorage
from ..types import Node


class MemoryMappingService:
    def __init__(self, file_path: str, dim: int, capacity: int, config_path: Optional[str] = None,
                 index_file: Optional[str] = None):
        if dim <= 0:
            raise ValueError("Dimension must be greater than 0")          # mmap_vector_store.py:39-40
        if capacity <= 0:
            raise ValueError("Capacity must be greater than 0")           # :41-42
        self.file_path = Path(file_path)
        self.dim = dim
        self.capacity = capacity
        self.size = 0
        if config_path is None:
            raise ValueError("config_path is required")                   # :50-51
        with open(config_path, "r") as f:
            config = yaml.safe_load(f) or {}
        index_config = config.get("index", {})
        M = index_config.get("M", 16)                                      # :54-55
        ef_construction = index_config.get("ef_construction", 200)
        self.storage = MMapNodeStorage(
            embedding_file=self.file_path.with_suffix(".embeddings.npy"),
            metadata_file=self.file_path.with_suffix(".metadata.npy"),
            dim=dim, capacity=capacity)
        index_path = Path(index_file) if index_file else self.file_path.with_suffix(".index.pkl")
        self.index = HNSW(M=M, ef_construction=ef_construction, rng=random.Random(42), storage=self.storage,
                          index_file=index_path)                           # :69-77
        self.size = self.storage.size()

    def write(self, embedding: np.ndarray, content: Optional[str] = None,
              metadata: Optional[Dict[str, Any]] = None) -> int:
        """Store one embedding and insert it into the index; returns the node id (mmap_vector_store.py:83-125)."""
        if not isinstance(embedding, np.ndarray):
            raise TypeError("Embedding must be a numpy array")
        if embedding.dtype != np.float32:
            embedding = embedding.astype(np.float32)
        if embedding.ndim != 1:
            raise ValueError("Embedding must be a 1D array")
        if embedding.size != self.dim:
            raise ValueError(f"Embedding must be of dimension {self.dim}")
        node_id = self.storage.get_next_id()
        node = Node(id=node_id, embedding=embedding, content=content, metadata=metadata or {})
        self.storage.save(node)
        self.index.insert_node(node)
        self.size = self.storage.size()
        return node_id

    def write_many(self, embeddings: np.ndarray, contents: Optional[List[Optional[str]]] = None,
                   metadatas: Optional[List[Optional[Dict[str, Any]]]] = None) -> List[int]:
        """Batched ``write`` (additive): the same ids, storage records and graph as writing the rows one by one,
        with the index inserts of the whole batch in few kernel launches."""
        if not isinstance(embeddings, np.ndarray):
            raise TypeError("Embedding must be a numpy array")
        if embeddings.ndim != 2 or embeddings.shape[1] != self.dim:
            raise ValueError(f"Embedding must be of dimension {self.dim}")
        embeddings = embeddings.astype(np.float32, copy=False)
        nodes = []
        for r in range(embeddings.shape[0]):
            node = Node(id=self.storage.get_next_id(), embedding=embeddings[r],
                        content=contents[r] if contents else None,
                        metadata=(metadatas[r] if metadatas else None) or {})
            self.storage.save(node)
            nodes.append(node)
        self.index.build_index(nodes)
        self.size = self.storage.size()
        return [n.id for n in nodes]

    def read(self, node_id: int) -> Node:
        if not isinstance(node_id, int):
            raise TypeError("Node ID must be an integer")                  # :137-138
        node = self.storage.get(node_id)
        if node is None:
            raise IndexError(f"Node {node_id} not found")                  # :141-142
        return node

    def get_embedding(self, node_id: int) -> np.ndarray:
        return self.storage.get_embedding(node_id)

    def delete(self, node_id: int) -> None:
        self.index.delete_node(node_id)                                    # :157
        if hasattr(self.storage, "delete"):
            self.storage.delete(node_id)
        self.size = self.storage.size()

    def search(self, query: np.ndarray, k: int, ef: int = 50) -> List[Tuple[Node, float]]:
        return self.index.search(query, k=k, ef=ef)                        # :177
