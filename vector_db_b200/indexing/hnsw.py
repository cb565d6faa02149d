"""HNSW index with the reference's API (src/vector_db/indexing/hnsw.py), graph and vectors in HBM.

Same constructor, methods and public attributes as the reference class (hnsw.py:30-64,143,227,279,
330,376,400): ``insert_node / build_index / delete_node / search / save_index / load_index`` and
``M, M_max, M_max0, ef_construction, rng, storage, index_file, node_store, entry_node_id, max_level,
level_mult``.  What changes is where the work happens:

  * the graph is a fixed-degree adjacency table in HBM (``adj0`` int32 [rows, 2M], ``adj_up`` int32
    [upper rows, M], -1 padded, ROW indices) next to the fp32 vectors (device_store.DeviceVectors);
  * ``search`` runs the batched traversal kernel (``search_batch`` exposes the many-query form);
  * ``insert_node`` / ``build_index`` run the sequential-semantics insert kernel, so the graph equals
    the reference's node-by-node build for the same ``random.Random`` level draws;
  * ``node_store`` is a read-only mapping view materialised from the device arrays on demand.
"""
from __future__ import annotations

import math
import pickle
import random
import threading
from collections.abc import Mapping
from pathlib import Path
from typing import Dict, Iterable, List, Optional, Sequence, Set, Tuple, Union

import numpy as np
import torch

from .. import _lib, ops
from ..device_store import DeviceVectors
from ..storage import InMemoryNodeStorage, NodeStorage
from ..types import Node

_INSERT_CHUNK = 2048   # rows per insert-kernel launch (one CTA walks them in order)


class _NodeStoreView(Mapping):
    """Read-only ``Dict[id, InternalNode]`` facade over the device adjacency (hnsw.py:57)."""

    def __init__(self, index: "HNSW"):
        self._index = index

    def _snap(self):
        return self._index._host_graph()

    def __len__(self):
        return len(self._index._graph_ids)

    def __iter__(self):
        return iter(list(self._index._graph_ids.keys()))

    def __contains__(self, node_id):
        return node_id in self._index._graph_ids

    def __getitem__(self, node_id):
        row = self._index._graph_ids[node_id]
        adj0, adj_up, up_ptr, levels, ids = self._snap()
        node = HNSW.InternalNode(node_id)
        for lvl in range(int(levels[row]) + 1):
            lst = adj0[row] if lvl == 0 else adj_up[up_ptr[row] + lvl - 1]
            linked = [int(ids[r]) for r in lst if r >= 0]
            # a level key exists if it was created at insert time or a later insert linked to it
            if lvl <= self._index._link_top[row] or linked:
                node.neighbors[lvl] = linked
        return node


class HNSW:
    class InternalNode:
        """Graph-only record: id + ``neighbors[level] -> list of ids`` (hnsw.py:23-28)."""

        def __init__(self, node_id: int):
            self.id = node_id
            self.neighbors: Dict[int, List[int]] = {}

    def __init__(self, M: int, ef_construction: int, rng: random.Random, storage: Optional[NodeStorage] = None,
                 index_file: Optional[Union[str, Path]] = None, device=None):
        if M < 2:
            raise ValueError("M must be >= 2")
        self.M = M
        self.M_max = M
        self.M_max0 = M * 2
        self.ef_construction = ef_construction
        self.rng = rng
        self.storage = storage or InMemoryNodeStorage()
        self.index_file = Path(index_file) if index_file else None
        self.level_mult = 1 / math.log(M)
        self._device = _lib.require_cuda() if device is None else torch.device(device)
        self._lock = threading.RLock()
        self._reset_graph(dim=None)
        if self.index_file and self.index_file.exists():
            self.load_index()

    # ------------------------------------------------------------------ device state
    def _reset_graph(self, dim: Optional[int]) -> None:
        self._vectors: Optional[DeviceVectors] = DeviceVectors(dim, self._device) if dim else None
        self._graph_ids: Dict[int, int] = {}          # id -> row, live graph nodes in insertion order
        self._levels: List[int] = []                  # per row
        self._up_ptr: List[int] = [0]                 # per row + 1
        self._link_top: List[int] = []                # highest level that has a neighbours entry
        self._adj0 = None
        self._adj_up = None
        self._levels_dev = None
        self._up_ptr_dev = None
        self._state = torch.tensor([-1, -1, 0, 0], dtype=torch.int32, device=self._device)
        self._state_host: Optional[Tuple[int, int]] = (-1, -1)
        self._snapshot = None
        self._meta_stale = set()                      # rows whose metadata changed since the columns were built

    def _ensure_capacity(self, rows: int, up_rows: int) -> None:
        dev = self._device
        if self._adj0 is None or self._adj0.shape[0] < rows:
            cap = max(rows, int((self._adj0.shape[0] if self._adj0 is not None else 0) * 1.5) + 1024)
            adj0 = torch.full((cap, self.M_max0), -1, dtype=torch.int32, device=dev)
            lv = torch.zeros((cap,), dtype=torch.int32, device=dev)
            up = torch.zeros((cap + 1,), dtype=torch.int32, device=dev)
            if self._adj0 is not None:
                n = self._adj0.shape[0]
                adj0[:n].copy_(self._adj0)
                lv[:n].copy_(self._levels_dev)
                up[:n + 1].copy_(self._up_ptr_dev)
            self._adj0, self._levels_dev, self._up_ptr_dev = adj0, lv, up
        if self._adj_up is None or self._adj_up.shape[0] < max(up_rows, 1):
            cap = max(up_rows, int((self._adj_up.shape[0] if self._adj_up is not None else 0) * 1.5) + 256)
            adj_up = torch.full((cap, self.M_max), -1, dtype=torch.int32, device=dev)
            if self._adj_up is not None:
                adj_up[:self._adj_up.shape[0]].copy_(self._adj_up)
            self._adj_up = adj_up

    def _read_state(self) -> Tuple[int, int]:
        if self._state_host is None:
            s = self._state.cpu().tolist()
            self._state_host = (int(s[0]), int(s[1]))
        return self._state_host

    def _host_graph(self):
        if self._snapshot is None:
            n = len(self._levels)
            self._snapshot = (self._adj0[:n].cpu().numpy(), self._adj_up[:max(self._up_ptr[-1], 1)].cpu().numpy(),
                              np.asarray(self._up_ptr, dtype=np.int64), np.asarray(self._levels, dtype=np.int64),
                              self._vectors.ids[:n].cpu().numpy())
        return self._snapshot

    def distance_evaluations(self) -> int:
        """Running count of distance evaluations made by inserts (parity check against the reference)."""
        s = self._state.cpu().numpy().astype(np.int64)
        return int((int(s[3]) & 0xFFFFFFFF) << 32 | (int(s[2]) & 0xFFFFFFFF))

    # ------------------------------------------------------------------ reference attributes
    @property
    def node_store(self) -> Mapping:
        return _NodeStoreView(self)

    @property
    def entry_node_id(self) -> Optional[int]:
        row, _ = self._read_state()
        if row < 0:
            return None
        return int(self._vectors.ids[row].item())

    @property
    def max_level(self) -> int:
        return self._read_state()[1]

    def sample_level(self) -> int:
        """Level draw of hnsw.py:66-68 (one ``rng.random()`` per new node)."""
        return int(-math.log(self.rng.random()) * self.level_mult)

    # ------------------------------------------------------------------ insert
    def insert_node(self, node: Node) -> None:
        self.build_index([node])

    def build_index(self, nodes: Sequence[Node]) -> None:
        """Insert nodes in order (hnsw.py:279-287 -> :143-225), batched into few kernel launches."""
        with self._lock:
            fresh_ids, fresh_vecs, fresh_set = [], [], set()
            for node in nodes:
                self.storage.save(node)                         # hnsw.py:150
                emb = np.asarray(node.embedding, dtype=np.float32).reshape(-1)
                if self._vectors is None:
                    self._vectors = DeviceVectors(emb.shape[0], self._device)
                if emb.shape[0] != self._vectors.dim:
                    raise ValueError(f"embedding dimension {emb.shape[0]} != index dimension {self._vectors.dim}")
                if node.id in self._graph_ids or node.id in fresh_set:
                    if node.id in self._graph_ids:              # idempotent insert (hnsw.py:153-155);
                        self._vectors.append([node.id], emb[None, :])   # storage was overwritten, mirror it
                        self._meta_stale.add(self._graph_ids[node.id])
                    continue
                fresh_ids.append(node.id)
                fresh_set.add(node.id)
                fresh_vecs.append(emb)
            if not fresh_ids:
                return
            first = self._vectors.n
            if first != len(self._levels):
                raise RuntimeError("device rows out of step with the graph")
            self._vectors.append(fresh_ids, np.stack(fresh_vecs))
            levels = [self.sample_level() for _ in fresh_ids]
            entry_row, top = self._read_state()
            for j, (nid, lvl) in enumerate(zip(fresh_ids, levels)):
                row = first + j
                self._graph_ids[nid] = row
                self._levels.append(lvl)
                self._up_ptr.append(self._up_ptr[-1] + lvl)
                if entry_row < 0:
                    self._link_top.append(lvl)      # first node: neighbors[l] = [] for l <= level (hnsw.py:164-171)
                    entry_row, top = row, lvl
                else:
                    self._link_top.append(min(lvl, top))   # levels above the old top get no entry (hnsw.py:179-191)
                    if lvl > top:
                        top = lvl
            n = len(self._levels)
            self._ensure_capacity(n, self._up_ptr[-1])
            self._levels_dev[first:n].copy_(torch.tensor(levels, dtype=torch.int32))
            self._up_ptr_dev[first:n + 1].copy_(torch.tensor(self._up_ptr[first:n + 1], dtype=torch.int32))
            lib = _lib.load()
            nbytes = lib.vdb_hnsw_insert_workspace(self.ef_construction, self.M_max0)
            ws = ops._ws("hnsw_insert", nbytes, self._device)
            vec = self._vectors
            ws[:4].zero_()                                      # sticky overflow flag of vdb_hnsw_insert: zeroed once
            for s in range(first, n, _INSERT_CHUNK):
                cnt = min(_INSERT_CHUNK, n - s)
                _lib.check(lib.vdb_hnsw_insert(
                    vec.vecs.data_ptr(), vec.n, vec.dim, self._adj0.data_ptr(), self.M_max0, self._adj_up.data_ptr(),
                    self.M_max, self._up_ptr_dev.data_ptr(), self._levels_dev.data_ptr(), s, cnt, self.M,
                    self.ef_construction, self._state.data_ptr(), vec.alive.data_ptr() if vec.n_dead else None,
                    ws.data_ptr(), ws.numel(), _lib.stream_ptr(self._device)))
            if int(ws[:4].view(torch.int32).item()) != 0:
                raise _lib.VdbError("hnsw insert: visited table overflow (ef_construction * degree too large)")
            self._state_host = None
            self._snapshot = None
            if self.index_file:
                self.save_index()                               # hnsw.py:224-225

    # ------------------------------------------------------------------ wave-parallel build
    def build_from_arrays(self, ids: Sequence[int], vectors: np.ndarray, wave_max: int = 16384, seed: int = 1024,
                          save_to_storage: bool = False) -> dict:
        """Build the whole index from scratch with the wave-parallel GPU builder (``vdb_hnsw_build_search`` +
        ``vdb_hnsw_reverse_merge``): levels are drawn with the reference's ``sample_level`` (hnsw.py:66-68) and
        built top-down; on each level a brute-force kNN seed of ``seed`` nodes is grown in waves whose members
        all search the graph built so far with ``ef_construction`` and link to their M closest.  This is what
        makes the 1M-row configurations buildable (the sequential-semantics ``build_index`` stays the parity
        builder); the graph is NOT the reference's graph.  Returns build statistics."""
        with self._lock:
            if self._graph_ids:
                raise ValueError("build_from_arrays builds from scratch: the index already holds nodes")
            on_device = isinstance(vectors, torch.Tensor) and vectors.is_cuda     # rows already in HBM (staged / generated)
            X = vectors.to(torch.float32).contiguous() if on_device else np.ascontiguousarray(vectors, dtype=np.float32)
            ids = [int(i) for i in ids]
            n, dim = X.shape
            if len(set(ids)) != n:
                raise ValueError("ids must be unique and match the rows of `vectors`")
            if save_to_storage:
                Xh = X.cpu().numpy() if on_device else X
                for i, nid in enumerate(ids):
                    self.storage.save(Node(id=nid, embedding=Xh[i]))
            dev = self._device
            self._reset_graph(dim=None)
            self._vectors = DeviceVectors(dim, dev, capacity=n)
            if on_device:
                self._vectors.append_device(torch.as_tensor(ids, dtype=torch.int64, device=dev), X)
            else:
                self._vectors.append(ids, X)
            levels = np.fromiter((self.sample_level() for _ in range(n)), dtype=np.int64, count=n)
            top = int(levels.max())
            up_ptr = np.zeros(n + 1, dtype=np.int64)
            up_ptr[1:] = np.cumsum(levels)
            self._levels = levels.tolist()
            self._up_ptr = up_ptr.tolist()
            self._link_top = self._levels[:]
            self._graph_ids = {nid: r for r, nid in enumerate(ids)}
            self._ensure_capacity(n, int(up_ptr[-1]))
            self._levels_dev[:n].copy_(torch.from_numpy(levels.astype(np.int32)))
            self._up_ptr_dev[:n + 1].copy_(torch.from_numpy(up_ptr.astype(np.int32)))
            adjd0 = torch.full(self._adj0.shape, float("inf"), dtype=torch.float32, device=dev)
            adjd_up = torch.full(self._adj_up.shape, float("inf"), dtype=torch.float32, device=dev)
            entry_row = int(np.flatnonzero(levels == top)[0])        # first node that reached the top level
            lib = _lib.load()
            vec = self._vectors
            M = self.M
            up_ptr_dev = self._up_ptr_dev
            waves = 0
            ws = ops._ws("hnsw_build", lib.vdb_hnsw_build_search_workspace(wave_max, self.ef_construction, self.M_max0), dev)
            ws[:4].zero_()                                           # sticky overflow flag of vdb_hnsw_build_search
            for level in range(top, -1, -1):
                rows = np.flatnonzero(levels >= level)
                rows = rows[np.lexsort((rows, -levels[rows]))]       # upper-level nodes first, then by row
                rows_t = torch.from_numpy(rows.astype(np.int32)).to(dev)
                adj, adjd = (self._adj0, adjd0) if level == 0 else (self._adj_up, adjd_up)
                deg = adj.shape[1]

                def lists_of(r: torch.Tensor) -> torch.Tensor:
                    r = r.long()
                    return r if level == 0 else (up_ptr_dev[r].long() + (level - 1))

                n0 = min(len(rows), max(seed, 2))
                if n0 >= 2:                                          # brute-force kNN seed
                    s_rows = rows_t[:n0].long()
                    Xs = vec.vecs[s_rows].contiguous()
                    kk = min(M + 1, n0)
                    d, i = ops.exact_topk_f32(Xs, Xs, kk, "l2")
                    keep = i != torch.arange(n0, device=dev).unsqueeze(1)
                    order = torch.argsort((~keep).to(torch.int8), dim=1, stable=True)[:, :kk - 1]
                    nb = s_rows[torch.gather(i, 1, order)].to(torch.int32)
                    nd = torch.gather(d, 1, order)
                    li = lists_of(s_rows)
                    adj[li, :kk - 1] = nb
                    adjd[li, :kk - 1] = nd
                pos = n0
                out_rows = torch.empty((wave_max, M), dtype=torch.int32, device=dev)
                out_d = torch.empty((wave_max, M), dtype=torch.float32, device=dev)
                while pos < len(rows):
                    b = int(min(max(pos // 4, 256), wave_max, len(rows) - pos))
                    wave = rows_t[pos:pos + b].contiguous()
                    _lib.check(lib.vdb_hnsw_build_search(
                        vec.vecs.data_ptr(), vec.n, vec.dim, self._adj0.data_ptr(), self.M_max0, self._adj_up.data_ptr(),
                        self.M_max, up_ptr_dev.data_ptr(), wave.data_ptr(), b, level, entry_row, top, int(rows[0]),
                        self.ef_construction, M, out_rows.data_ptr(), out_d.data_ptr(), ws.data_ptr(), ws.numel(),
                        _lib.stream_ptr(dev)))
                    fr, fd = out_rows[:b], out_d[:b]
                    li = lists_of(wave)
                    adj[li, :M] = fr
                    adjd[li, :M] = fd
                    valid = fr >= 0
                    tgt = fr[valid].long()
                    src = wave.unsqueeze(1).expand(b, M)[valid]
                    dd = fd[valid]
                    key = (tgt << 32) | dd.view(torch.int32).long()   # distances are >= 0: their bits sort like the values
                    key, perm = torch.sort(key)
                    tgt_sorted = key >> 32
                    uniq, counts = torch.unique_consecutive(tgt_sorted, return_counts=True)
                    seg_start = torch.zeros(uniq.numel() + 1, dtype=torch.int32, device=dev)
                    seg_start[1:] = torch.cumsum(counts, 0).to(torch.int32)
                    seg_lists = lists_of(uniq).to(torch.int32)
                    in_rows = src[perm].contiguous()
                    in_d = dd[perm].contiguous()
                    _lib.check(lib.vdb_hnsw_reverse_merge(adj.data_ptr(), adjd.data_ptr(), deg, seg_lists.data_ptr(),
                                                          seg_start.data_ptr(), int(uniq.numel()), in_rows.data_ptr(),
                                                          in_d.data_ptr(), _lib.stream_ptr(dev)))
                    pos += b
                    waves += 1
            if int(ws[:4].view(torch.int32).item()) != 0:
                raise _lib.VdbError("hnsw build: visited table overflow (ef_construction * degree too large)")
            st = (entry_row, top)
            self._state[:2].copy_(torch.tensor(st, dtype=torch.int32))
            self._state_host = st
            self._snapshot = None
            if self.index_file:
                self.save_index()
            return {"nodes": n, "top_level": top, "waves": waves, "entry_row": entry_row}

    # ------------------------------------------------------------------ delete
    def delete_node(self, node_id: int) -> None:
        """hnsw.py:227-277: unlink the node from the lists of ITS OWN neighbours, drop it, fix the
        entry point, delete from storage.  One-way links pointing at it stay and become tombstones."""
        with self._lock:
            row = self._graph_ids.get(node_id)
            if row is None:
                return
            dev = self._device
            for lvl in range(self._levels[row] + 1):
                lst_t = self._adj0[row] if lvl == 0 else self._adj_up[self._up_ptr[row] + lvl - 1]
                for nb in [int(r) for r in lst_t.cpu().tolist() if r >= 0]:
                    if self._levels[nb] < lvl:
                        continue
                    tgt = self._adj0[nb] if lvl == 0 else self._adj_up[self._up_ptr[nb] + lvl - 1]
                    cur = [r for r in tgt.cpu().tolist() if r >= 0 and r != row]
                    cur = cur + [-1] * (tgt.shape[0] - len(cur))
                    tgt.copy_(torch.tensor(cur, dtype=torch.int32, device=dev))
                lst_t.fill_(-1)
            del self._graph_ids[node_id]
            self._vectors.mark_dead(node_id)
            entry_row, top = self._read_state()
            if entry_row == row:
                if self._graph_ids:
                    # highest-level surviving node, first in insertion order (hnsw.py:253-266)
                    best_row, best_lvl = -1, -2
                    self._snapshot = None
                    view = self.node_store
                    for nid, r in self._graph_ids.items():
                        keys = view[nid].neighbors.keys()
                        lvl = max(keys) if keys else -1
                        if lvl > best_lvl:
                            best_row, best_lvl = r, lvl
                    new_state = (best_row, best_lvl)
                else:
                    new_state = (-1, -1)
                self._state[:2].copy_(torch.tensor(new_state, dtype=torch.int32))
                self._state_host = new_state
            if hasattr(self.storage, "delete"):
                self.storage.delete(node_id)
            self._snapshot = None
            if self.index_file:
                self.save_index()

    # ------------------------------------------------------------------ search
    def _metadata_bitmap(self, metadata_filter: dict) -> torch.Tensor:
        """Row bitmap of a metadata equality filter, evaluated on the GPU (storage_service.py:106-128 semantics).
        Rows that came in without going through insert_node / build_index (load_index, build_from_arrays) get
        their metadata from the storage on first use."""
        vec = self._vectors
        meta = vec.meta
        todo = sorted(self._meta_stale) + list(range(meta.rows, vec.n))   # re-inserted ids, then rows not seen yet
        if todo:
            host_ids = vec.ids[:vec.n].cpu().numpy()
            for r in todo:
                try:
                    node = self.storage.get(int(host_ids[r]))
                except Exception:                               # noqa: BLE001 -- missing nodes are skipped (hnsw.py:120-121)
                    node = None
                meta.set_row(r, node.metadata if node else None)
            meta.rows = max(meta.rows, vec.n)
            self._meta_stale.clear()
        return vec.metadata_bitmap(metadata_filter)

    def search_batch(self, queries: np.ndarray, k: int, ef: int = 50, filter_ids=None,
                     return_stats: bool = False, metadata_filter: Optional[dict] = None):
        """Many queries at once: (dist f32 [Q,k], ids i64 [Q,k]) ordered by (dist,id), padded (+inf,-1).
        ``filter_ids``: id set or row bitmap; ``metadata_filter``: {key: value} evaluated on the GPU (additive
        kwarg: the reference builds the id set on the host, api/app.py:184-198)."""
        with self._lock:
            queries = np.ascontiguousarray(queries, dtype=np.float32)
            if queries.ndim == 1:
                queries = queries[None, :]
            Q = queries.shape[0]
            entry_row, top = self._read_state()
            if entry_row < 0 or self._vectors is None:
                d = np.full((Q, k), np.inf, np.float32)
                i = np.full((Q, k), -1, np.int64)
                return (d, i, np.zeros((Q, 2), np.int64)) if return_stats else (d, i)
            vec = self._vectors
            if queries.shape[1] != vec.dim:
                raise ValueError(f"query dimension {queries.shape[1]} != index dimension {vec.dim}")
            q = torch.from_numpy(queries).to(self._device)
            flt = vec.bitmap_from_ids(filter_ids) if filter_ids is not None else None
            if metadata_filter is not None:
                mb = self._metadata_bitmap(metadata_filter)
                flt = mb if flt is None else torch.bitwise_and(flt[:mb.numel()], mb)
            alive = vec.alive if vec.n_dead else None
            kk = min(k, ef)
            d, r, st = ops.hnsw_search(q, vec.vecs, vec.n, self._adj0, self._adj_up, self._up_ptr_dev, entry_row, top,
                                       ef, kk, alive=alive, flt=flt, want_stats=return_stats, levels=self._levels_dev)
            ids = torch.where(r >= 0, vec.ids[r.clamp(min=0).long()], torch.full_like(r, -1, dtype=torch.int64))
            d_h, i_h = d.cpu().numpy(), ids.cpu().numpy()
            if kk < k:
                d_h = np.concatenate([d_h, np.full((Q, k - kk), np.inf, np.float32)], axis=1)
                i_h = np.concatenate([i_h, np.full((Q, k - kk), -1, np.int64)], axis=1)
            if return_stats:
                return d_h, i_h, st.cpu().numpy().astype(np.int64)
            return d_h, i_h

    def search(self, query: np.ndarray, k: int, **kwargs) -> List[Tuple[Node, float]]:
        """Same contract as hnsw.py:330-374: kwargs ``ef`` (default 50) and ``filter_ids``; all other
        kwargs are ignored; returns ``[(Node, np.float32 distance)]`` sorted by (distance, id)."""
        ef = kwargs.get("ef", 50)
        filter_ids = kwargs.get("filter_ids")
        if not self._graph_ids:
            return []
        query = np.asarray(query)
        if query.ndim != 1:
            raise ValueError("query must be 1D array")
        d, i = self.search_batch(query[None, :], k, ef=ef, filter_ids=filter_ids, metadata_filter=kwargs.get("metadata_filter"))
        results = []
        for dist, node_id in zip(d[0], i[0]):
            if node_id < 0:
                break
            node = self.storage.get(int(node_id))
            if node:
                results.append((node, np.float32(dist)))
        return results

    # ------------------------------------------------------------------ persistence (hnsw.py:376-423)
    def save_index(self) -> None:
        if self.index_file is None:
            return
        with self._lock:
            view = self.node_store
            data = {
                "M": self.M, "M_max": self.M_max, "M_max0": self.M_max0, "ef_construction": self.ef_construction,
                "max_level": self.max_level, "entry_node_id": self.entry_node_id,
                "node_store": {nid: {"id": nid, "neighbors": view[nid].neighbors} for nid in view},
            }
            with open(self.index_file, "wb") as f:
                pickle.dump(data, f)

    def load_index(self) -> None:
        if self.index_file is None or not self.index_file.exists():
            return
        with open(self.index_file, "rb") as f:
            data = pickle.load(f)
        self.load_graph(data)

    def load_graph(self, data: dict, vectors: Optional[Dict[int, np.ndarray]] = None) -> None:
        """Import a graph in the reference's pickle schema; embeddings come from ``vectors`` or storage."""
        with self._lock:
            self.M = data["M"]
            self.M_max = data["M_max"]
            self.M_max0 = data["M_max0"]
            self.ef_construction = data["ef_construction"]
            self.level_mult = 1 / math.log(self.M)
            store = data["node_store"]
            ids = list(store.keys())
            self._reset_graph(dim=None)
            if not ids:
                return
            get = (lambda i: vectors[i]) if vectors is not None else self.storage.get_embedding
            X = np.stack([np.asarray(get(i), dtype=np.float32) for i in ids])
            self._vectors = DeviceVectors(X.shape[1], self._device, capacity=len(ids))
            self._vectors.append(ids, X)
            row = {nid: r for r, nid in enumerate(ids)}
            n = len(ids)
            levels = []
            for nid in ids:
                nb = store[nid]["neighbors"]
                levels.append(max(nb.keys()) if nb else 0)
            # The reference gives a node inserted above the old top NO lists for the levels above it
            # (hnsw.py:179-191) but makes it the entry point at its full level (hnsw.py:215-217): the pickle then
            # carries max_level > max(entry's keys).  Reserve the entry's rows up to max_level (empty lists) so the
            # greedy descent reads its own rows, and later inserts at those levels link into them.
            entry = data["entry_node_id"]
            if entry is not None and entry in row:
                levels[row[entry]] = max(levels[row[entry]], int(data["max_level"]))
            up_ptr = np.zeros(n + 1, dtype=np.int64)
            up_ptr[1:] = np.cumsum(levels)
            adj0 = np.full((n, self.M_max0), -1, dtype=np.int32)
            adj_up = np.full((max(int(up_ptr[-1]), 1), self.M_max), -1, dtype=np.int32)
            for r, nid in enumerate(ids):
                for lvl, lst in store[nid]["neighbors"].items():
                    rows = [row[x] for x in lst if x in row]
                    tgt = adj0[r] if lvl == 0 else adj_up[up_ptr[r] + lvl - 1]
                    tgt[:len(rows)] = rows[:tgt.shape[0]]
            self._graph_ids = row
            self._levels = [int(l) for l in levels]
            self._up_ptr = [int(x) for x in up_ptr]
            self._link_top = [max(store[nid]["neighbors"].keys()) if store[nid]["neighbors"] else -1 for nid in ids]
            self._ensure_capacity(n, int(up_ptr[-1]))
            self._adj0[:n].copy_(torch.from_numpy(adj0))
            self._adj_up[:adj_up.shape[0]].copy_(torch.from_numpy(adj_up))
            self._levels_dev[:n].copy_(torch.tensor(self._levels, dtype=torch.int32))
            self._up_ptr_dev[:n + 1].copy_(torch.from_numpy(up_ptr.astype(np.int32)))
            st = (row[entry] if entry is not None else -1, int(data["max_level"]))
            self._state[:2].copy_(torch.tensor(st, dtype=torch.int32))
            self._state_host = st
            self._snapshot = None
