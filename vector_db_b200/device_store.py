"""HBM-resident mirror of the embeddings an index searches.

Layout (all torch CUDA tensors, struct-of-arrays, sized for 180 GB of HBM3e):
    vecs   f32 [cap, D]   row-major, row = insertion order
    ids    i64 [cap]      external node id of each row
    alive  u32 bitmap     bit r cleared = tombstone (deleted node still linked in a graph)
    sqnorm f32 [n], bf16 mirror [n, D]   built lazily for the tensor-core scan
Host side keeps the id -> row dict (ids are arbitrary Python ints in the reference).
Growth is geometric; appends are batched H2D copies from pinned memory.
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional

import numpy as np
import torch

from . import _lib, ops


class DeviceVectors:
    def __init__(self, dim: int, device=None, capacity: int = 1024):
        self.device = _lib.require_cuda() if device is None else torch.device(device)
        self.dim = int(dim)
        self.n = 0
        self._cap = 0
        self.vecs = torch.empty((0, self.dim), dtype=torch.float32, device=self.device)
        self.ids = torch.empty((0,), dtype=torch.int64, device=self.device)
        self.alive = torch.empty((0,), dtype=torch.int32, device=self.device)
        self._alive_host = np.zeros(0, dtype=np.uint32)
        self.id_to_row: Dict[int, int] = {}
        self.n_dead = 0
        self._prepared_n = -1
        self._prepared_metric = -1
        self._sqnorm = None
        self._bf16 = None
        self._tc_stats = None
        self._reserve(capacity)

    # -- capacity ---------------------------------------------------------------------------
    def _reserve(self, need: int) -> None:
        if need <= self._cap:
            return
        cap = max(need, int(self._cap * 1.5) + 1024)
        vecs = torch.empty((cap, self.dim), dtype=torch.float32, device=self.device)
        ids = torch.full((cap,), -1, dtype=torch.int64, device=self.device)
        alive = torch.zeros(((cap + 31) // 32,), dtype=torch.int32, device=self.device)
        alive_host = np.zeros((cap + 31) // 32, dtype=np.uint32)
        if self.n:
            vecs[:self.n].copy_(self.vecs[:self.n])
            ids[:self.n].copy_(self.ids[:self.n])
            alive[:self.alive.numel()].copy_(self.alive)
            alive_host[:self._alive_host.shape[0]] = self._alive_host
        self._alive_host = alive_host
        self.vecs, self.ids, self.alive, self._cap = vecs, ids, alive, cap

    # -- mutation ---------------------------------------------------------------------------
    def append(self, ids: Iterable[int], vectors: np.ndarray) -> np.ndarray:
        """Append rows (host arrays); returns the row indices.  Existing ids are overwritten in place."""
        vectors = np.ascontiguousarray(vectors, dtype=np.float32).reshape(-1, self.dim)
        ids = [int(i) for i in ids]
        rows = np.empty(len(ids), dtype=np.int64)
        new_pos = []
        for j, i in enumerate(ids):
            r = self.id_to_row.get(i)
            if r is None:
                r = self.n + len(new_pos)
                self.id_to_row[i] = r
                new_pos.append(j)
            rows[j] = r
        self._reserve(self.n + len(new_pos))
        src = torch.from_numpy(vectors)
        if len(new_pos) == len(ids):
            dst = slice(self.n, self.n + len(ids))
            self.vecs[dst].copy_(src.pin_memory() if src.numel() > (1 << 16) else src, non_blocking=True)
            self.ids[dst].copy_(torch.tensor(ids, dtype=torch.int64))
        else:
            r_t = torch.from_numpy(rows).to(self.device)
            self.vecs.index_copy_(0, r_t, src.to(self.device))
            self.ids.index_copy_(0, r_t, torch.tensor(ids, dtype=torch.int64, device=self.device))
        old_n = self.n
        self.n += len(new_pos)
        self._set_alive(rows, True)
        if len(new_pos) != len(ids) or old_n != self.n:
            self._prepared_n = -1
        return rows

    def append_device(self, ids: torch.Tensor, vecs: torch.Tensor) -> None:
        """Bulk append of device tensors (from MMapNodeStorage.stage_to_device); ids must be new."""
        cnt = int(ids.numel())
        host_ids = ids.cpu().tolist()
        self._reserve(self.n + cnt)
        self.vecs[self.n:self.n + cnt].copy_(vecs)
        self.ids[self.n:self.n + cnt].copy_(ids)
        for j, i in enumerate(host_ids):
            self.id_to_row[int(i)] = self.n + j
        rows = np.arange(self.n, self.n + cnt)
        self.n += cnt
        self._set_alive(rows, True)
        self._prepared_n = -1

    def _set_alive(self, rows: np.ndarray, value: bool) -> None:
        """Update the host mirror of the bitmap and upload the touched word range."""
        if len(rows) == 0:
            return
        rows = np.asarray(rows, dtype=np.int64)
        words, bits = rows >> 5, (np.uint32(1) << (rows & 31).astype(np.uint32))
        if value:
            np.bitwise_or.at(self._alive_host, words, bits)
        else:
            np.bitwise_and.at(self._alive_host, words, ~bits)
        lo, hi = int(words.min()), int(words.max()) + 1
        self.alive[lo:hi].copy_(torch.from_numpy(self._alive_host[lo:hi].view(np.int32)))

    def mark_dead(self, node_id: int) -> Optional[int]:
        r = self.id_to_row.pop(int(node_id), None)
        if r is None:
            return None
        self._set_alive(np.array([r]), False)
        self.n_dead += 1
        return r

    def bitmap_from_ids(self, ids) -> torch.Tensor:
        """Row bitmap (uint32 words as int32 tensor) with the bits of the given external ids set.  A device
        tensor is taken to BE such a bitmap already (metadata_bitmap's result) and is passed through."""
        if isinstance(ids, torch.Tensor):
            if ids.dtype != torch.int32 or ids.numel() < (max(self.n, 1) + 31) // 32:
                raise ValueError("a row bitmap must be int32 words covering every row")
            return ids.to(self.device)
        words = np.zeros((max(self.n, 1) + 31) // 32, dtype=np.uint32)
        for i in ids:
            r = self.id_to_row.get(int(i))
            if r is not None:
                words[r // 32] |= np.uint32(1 << (r % 32))
        return torch.from_numpy(words.view(np.int32)).to(self.device)

    # -- metadata columns ----------------------------------------------------
    @property
    def meta(self):
        if getattr(self, "_meta", None) is None:
            from .metadata import DeviceMetadata
            self._meta = DeviceMetadata(self.device)
        return self._meta

    def set_metadata(self, ids: Iterable[int], metadatas) -> None:
        """Attach metadata dicts to stored ids (ids that are not stored are skipped)."""
        for i, md in zip(ids, metadatas):
            r = self.id_to_row.get(int(i))
            if r is not None:
                self.meta.set_row(r, md)

    def metadata_bitmap(self, filter_dict) -> torch.Tensor:
        """Row bitmap of the live rows whose metadata matches every key of the filter, evaluated on the GPU
        (storage_service.py:106-128 semantics); usable wherever ``filter_ids`` is accepted."""
        words = (max(self.n, 1) + 31) // 32
        return self.meta.filter_bitmap(filter_dict, self.n, alive=self.alive[:words])

    def bitmap_to_ids(self, bitmap: torch.Tensor) -> set:
        w = bitmap[:(max(self.n, 1) + 31) // 32].cpu().numpy().view(np.uint32)
        rows = np.nonzero(np.unpackbits(w.view(np.uint8), bitorder="little")[:self.n])[0]
        return set(self.ids[:self.n].cpu().numpy()[rows].tolist())

    # -- derived arrays for the tensor-core scan ------------------------------------------------
    def prepared(self, metric):
        """(operand mirror bf16, stats f32 [2] on the device) for the tensor-core scan of `metric`
        -- rebuilt after every mutation; no host synchronisation."""
        m = ops._metric(metric)
        if self._prepared_n != self.n or self._prepared_metric != m:
            mir, sq, stats = ops.tc_prepare_rows(self.vecs[:self.n], m)
            self._sqnorm, self._bf16, self._tc_stats = sq, mir, stats
            self._prepared_n, self._prepared_metric = self.n, m
        return self._bf16, self._tc_stats

    def rows_to_ids(self, rows: np.ndarray) -> np.ndarray:
        host_ids = self.ids[:self.n].cpu().numpy()
        rows = np.asarray(rows)
        out = np.full(rows.shape, -1, dtype=np.int64)
        ok = rows >= 0
        out[ok] = host_ids[rows[ok]]
        return out
