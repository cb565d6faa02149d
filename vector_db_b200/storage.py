"""Node storage with the reference's interface and on-disk formats, plus the HBM staging path.

Mirrors src/vector_db/storage.py: ``NodeStorage`` ABC (:15-51), ``InMemoryNodeStorage`` (:54-97) and
the two-layer ``MMapNodeStorage`` (:100-306).  File formats are byte-compatible with the reference
(layer 1: records ``{int64 id; float32 embedding[dim]}``, stride 8+4*dim, storage.py:135-144;
layer 2: ``{int64 id; U10240 content; U5120 metadata_json}``, storage.py:163-181), so a database
written by the reference opens here and vice versa.  What changes:

  * the O(capacity) Python scans on open / on every new save (storage.py:148-159,183-196) are
    vectorised numpy passes plus a free-slot heap;
  * ``stage_to_device`` streams layer 1 through a pinned host ring into SoA tensors in HBM: the
    strided DMA of ``vdb_stage_aos_h2d`` de-interleaves ids and vectors in flight.
"""
from __future__ import annotations

import heapq
import json
from abc import ABC, abstractmethod
from pathlib import Path
from typing import Dict, List, Optional, Tuple, Union

import numpy as np

from .types import Node


class NodeStorage(ABC):
    """Abstract base class for node storage backends (same methods as the reference ABC)."""

    @abstractmethod
    def save(self, node: Node) -> None: ...

    @abstractmethod
    def get(self, node_id: int) -> Optional[Node]: ...

    @abstractmethod
    def get_embedding(self, node_id: int) -> np.ndarray: ...

    @abstractmethod
    def get_all_ids(self) -> List[int]: ...

    @abstractmethod
    def size(self) -> int: ...

    @abstractmethod
    def close(self) -> None: ...

    @abstractmethod
    def add(self, node: Node) -> None: ...


class InMemoryNodeStorage(NodeStorage):
    def __init__(self):
        self._nodes: Dict[int, Node] = {}
        self._next_id = 0

    def save(self, node: Node) -> None:
        self._nodes[node.id] = node
        if node.id >= self._next_id:
            self._next_id = node.id + 1

    def add(self, node: Node) -> None:
        self.save(node)

    def get(self, node_id: int) -> Optional[Node]:
        return self._nodes.get(node_id)

    def get_embedding(self, node_id: int) -> np.ndarray:
        node = self._nodes.get(node_id)
        if node is None:
            raise KeyError(f"Node {node_id} not found")
        return node.embedding

    def get_all_ids(self) -> List[int]:
        return list(self._nodes.keys())

    def size(self) -> int:
        return len(self._nodes)

    def close(self) -> None:
        pass

    def delete(self, node_id: int) -> None:
        self._nodes.pop(node_id, None)

    def get_next_id(self) -> int:
        return self._next_id


_CONTENT_CHARS = 10240
_META_CHARS = 5120


class MMapNodeStorage(NodeStorage):
    """Two-layer memmap storage, byte-compatible with the reference's files."""

    def __init__(self, embedding_file: Union[str, Path], metadata_file: Optional[Union[str, Path]], dim: int,
                 capacity: int = 1_000_000):
        """``metadata_file=None`` opens layer 1 only (embeddings: what ``stage_to_device`` / ``get_embedding`` need);
        ``save`` / ``get`` / ``delete`` then leave layer 2 alone."""
        self.embedding_file = Path(embedding_file)
        self.metadata_file = Path(metadata_file) if metadata_file is not None else None
        self.dim = dim
        self.capacity = capacity
        self._init_embedding_memmap()
        self._init_metadata_memmap()

    # -- layer 1 ------------------------------------------------------------------------------
    def _init_embedding_memmap(self) -> None:
        mode = "r+" if self.embedding_file.exists() else "w+"
        dtype = np.dtype([("id", np.int64), ("embedding", np.float32, (self.dim,))])
        self.embeddings = np.memmap(self.embedding_file, dtype=dtype, mode=mode, shape=(self.capacity,))
        self._id_to_index: Dict[int, int] = {}
        self._free: List[int] = []       # heap of freed slots below the high-water mark
        self._hw = 0                     # every slot >= _hw is free
        if mode == "r+":
            # slot is valid iff id >= 0 and the embedding is not all zeros (storage.py:156)
            step = max(1, (64 << 20) // self.embeddings.dtype.itemsize)
            for s in range(0, self.capacity, step):
                blk = self.embeddings[s:s + step]
                valid = (blk["id"] >= 0) & np.any(blk["embedding"] != 0.0, axis=1)
                idx = np.nonzero(valid)[0]
                if idx.size:
                    ids = blk["id"][idx]
                    self._id_to_index.update(zip(ids.tolist(), (idx + s).tolist()))
                    self._hw = int(idx[-1]) + s + 1
            used = np.zeros(self._hw, dtype=bool)
            if self._id_to_index:
                used[np.fromiter(self._id_to_index.values(), dtype=np.int64)] = True
            self._free = np.nonzero(~used)[0].tolist()
            heapq.heapify(self._free)
        self._size = len(self._id_to_index)

    # -- layer 2 ------------------------------------------------------------------------------
    def _init_metadata_memmap(self) -> None:
        if self.metadata_file is None:
            self.metadata = None
            return
        mode = "r+" if self.metadata_file.exists() else "w+"
        dtype = np.dtype([("id", np.int64), ("content", f"U{_CONTENT_CHARS}"), ("metadata_json", f"U{_META_CHARS}")])
        self.metadata = np.memmap(self.metadata_file, dtype=dtype, mode=mode, shape=(self.capacity,))

    def _find_free_slot(self) -> int:
        """Lowest free slot -- the slot the reference's linear scan (storage.py:183-196) finds."""
        if self._size >= self.capacity:
            raise RuntimeError(f"Storage capacity ({self.capacity}) exceeded")
        if self._free:
            return heapq.heappop(self._free)
        idx = self._hw
        self._hw += 1
        return idx

    def save(self, node: Node) -> None:
        if node.id in self._id_to_index:
            idx = self._id_to_index[node.id]
        else:
            idx = self._find_free_slot()
            self._id_to_index[node.id] = idx
            self._size += 1
        self.embeddings[idx]["id"] = node.id
        self.embeddings[idx]["embedding"] = np.asarray(node.embedding).astype(np.float32)
        if self.metadata is None:
            self.embeddings.flush()
            return
        metadata_json = json.dumps(node.metadata) if node.metadata else ""
        content = node.content or ""
        if len(content) > _CONTENT_CHARS:
            content = content[:_CONTENT_CHARS - 1]
        if len(metadata_json) > _META_CHARS:
            metadata_json = metadata_json[:_META_CHARS - 1]
        self.metadata[idx]["id"] = node.id
        self.metadata[idx]["content"] = content
        self.metadata[idx]["metadata_json"] = metadata_json
        self.embeddings.flush()
        self.metadata.flush()

    def get(self, node_id: int) -> Optional[Node]:
        idx = self._id_to_index.get(node_id)
        if idx is None:
            return None
        embedding = self.embeddings[idx]["embedding"].copy()
        if self.metadata is None:
            return Node(id=node_id, embedding=embedding, content=None, metadata={})
        content = str(self.metadata[idx]["content"]).strip("\x00") or None
        metadata_json = str(self.metadata[idx]["metadata_json"]).strip("\x00") or None
        metadata = json.loads(metadata_json) if metadata_json else {}
        return Node(id=node_id, embedding=embedding, content=content, metadata=metadata)

    def get_embedding(self, node_id: int) -> np.ndarray:
        idx = self._id_to_index.get(node_id)
        if idx is None:
            raise KeyError(f"Node {node_id} not found")
        return self.embeddings[idx]["embedding"]

    def get_all_ids(self) -> List[int]:
        return list(self._id_to_index.keys())

    def size(self) -> int:
        return self._size

    def close(self) -> None:
        if hasattr(self, "embeddings"):
            del self.embeddings
        if hasattr(self, "metadata"):
            del self.metadata

    def get_next_id(self) -> int:
        if not self._id_to_index:
            return 0
        return max(self._id_to_index.keys()) + 1

    def add(self, node: Node) -> None:
        self.save(node)

    def delete(self, node_id: int) -> None:
        idx = self._id_to_index.pop(node_id, None)
        if idx is None:
            return
        # the reference clears the slot and sets id 0, not -1 (storage.py:295-299)
        self.embeddings[idx]["id"] = 0
        self.embeddings[idx]["embedding"] = 0.0
        if self.metadata is not None:
            self.metadata[idx]["id"] = 0
            self.metadata[idx]["content"] = ""
            self.metadata[idx]["metadata_json"] = ""
        self._size -= 1
        heapq.heappush(self._free, idx)
        self.embeddings.flush()
        if self.metadata is not None:
            self.metadata.flush()

    # -- HBM staging ----------------------------------------------------------------------------
    def stage_to_device(self, device=None, chunk_records: int = 8192, threads: Optional[int] = None, depth: Optional[int] = None
                        ) -> Tuple["torch.Tensor", "torch.Tensor"]:
        """Upload every valid record to HBM as SoA tensors ``(ids i64 [n], vecs f32 [n,dim])`` in slot
        order.  ``threads`` host workers each take every ``threads``-th chunk of ``chunk_records`` records: memcpy out
        of the memmap into a pinned buffer, one contiguous DMA, a kernel that splits the id and embedding columns
        (vdb_stage_aos_pageable_h2d); ``depth`` buffers in all (default two per worker); invalid slots are dropped on
        the device.  ``self.last_stage`` records bytes / seconds of the last upload."""
        import os
        import time

        import torch

        from . import _lib

        dev = _lib.require_cuda() if device is None else torch.device(device)
        lib = _lib.load()
        hw = self._hw
        ids = torch.empty(hw, dtype=torch.int64, device=dev)
        vecs = torch.empty((hw, self.dim), dtype=torch.float32, device=dev)
        if hw == 0:
            return ids, vecs
        rec = self.embeddings.dtype.itemsize
        chunk = max(1, min(chunk_records, hw))
        if threads is None:
            try:
                cpus = len(os.sched_getaffinity(0))
            except Exception:                                   # noqa: BLE001
                cpus = os.cpu_count() or 2
            threads = max(1, min(16, cpus))
        threads = max(1, min(int(threads), 64))
        depth = 2 * threads if depth is None else max(2, min(int(depth), 128))
        ring = getattr(self, "_stage_ring", None)
        if ring is None or ring.numel() < depth * chunk * rec:
            ring = self._stage_ring = torch.empty(depth * chunk * rec, dtype=torch.uint8).pin_memory()
        dring = torch.empty(depth * chunk * rec, dtype=torch.uint8, device=dev)
        stream = torch.cuda.current_stream(dev)
        raw = self.embeddings.view(np.uint8).reshape(self.capacity, rec)
        t0 = time.perf_counter()
        _lib.check(lib.vdb_stage_aos_pageable_h2d(raw.ctypes.data, hw, self.dim, ids.data_ptr(), vecs.data_ptr(),
                                                  ring.data_ptr(), dring.data_ptr(), chunk * rec, depth, int(threads),
                                                  stream.cuda_stream))
        stream.synchronize()
        dt = time.perf_counter() - t0
        self.last_stage = {"bytes": hw * rec, "seconds": dt, "gb_per_s": hw * rec / dt / 1e9, "threads": int(threads),
                           "depth": depth, "chunk_bytes": chunk * rec}
        valid = (ids >= 0) & (vecs != 0).any(dim=1)
        if bool(valid.all()):
            return ids, vecs
        keep = valid.nonzero().squeeze(1)
        return ids.index_select(0, keep), vecs.index_select(0, keep)
