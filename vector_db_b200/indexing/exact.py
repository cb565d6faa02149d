"""Exact brute-force index: batched L2 / cosine top-k over rows resident in HBM.

The reference has no brute-force search API; it has the semantics (the per-pair loop + tuple
sort of tests/indexing/test_hnsw.py:128-134 and IvfIndex.search with n_probe == k,
src/vector_db/indexing/ivf.py:170-199).  This class is that scan as a batched GPU operator:
results are ordered by (distance, id) ascending for L2 (non-squared distances) and by
(-similarity, id) for cosine, padded with (+/-inf, -1).
"""
from __future__ import annotations

from typing import Callable, Iterable, Iterator, Optional, Sequence, Set, Tuple

import numpy as np
import torch

from .. import _lib, ops
from ..device_store import DeviceVectors


class ExactIndex:
    def __init__(self, dim: int, metric: str = "l2", device=None, engine: str = "auto"):
        self.dim = int(dim)
        self.metric = ops._metric(metric)
        self.engine = engine          # "auto" | "f32" | "tc"
        self.store = DeviceVectors(dim, device)
        self._pin = None
        self._last_uncertified = 0    # int or 0-dim device tensor (read lazily: reading synchronises)

    @property
    def last_uncertified(self) -> int:
        """Queries of the last tensor-core search that the certificate rejected (re-scanned in fp32)."""
        return int(self._last_uncertified)

    # -- population -----------------------------------------------------------------------------
    def add(self, ids: Sequence[int], vectors: np.ndarray, metadata: Optional[Sequence[dict]] = None) -> None:
        ids = list(ids)
        self.store.append(ids, vectors)
        if metadata is not None:
            self.store.set_metadata(ids, metadata)

    def metadata_bitmap(self, metadata_filter: dict) -> torch.Tensor:
        """GPU-evaluated equality filter (storage_service.py:106-128 semantics) as a row bitmap; pass it as
        ``filter_ids`` to the searches."""
        return self.store.metadata_bitmap(metadata_filter)

    def add_device(self, ids: torch.Tensor, vecs: torch.Tensor) -> None:
        self.store.append_device(ids, vecs)

    def __len__(self) -> int:
        return self.store.n

    # -- search -----------------------------------------------------------------------------------
    def search_device(self, q: torch.Tensor, k: int, filter_ids=None) -> Tuple[torch.Tensor, torch.Tensor]:
        """``filter_ids``: a set of node ids, a metadata filter dict (evaluated on the GPU), or a row bitmap."""
        st = self.store
        if isinstance(filter_ids, dict):
            filter_ids = st.metadata_bitmap(filter_ids)
        mask = st.bitmap_from_ids(filter_ids) if filter_ids is not None else None
        if mask is not None and st.n_dead and not isinstance(filter_ids, torch.Tensor):
            mask = torch.bitwise_and(mask, st.alive[:mask.numel()])
        if mask is None and st.n_dead:
            mask = st.alive
        use_tc = (self.engine in ("auto", "tc") and mask is None and ops.tc_supported(q.shape[0], st.n, self.dim, k, self.metric))
        if self.engine == "tc" and not use_tc:
            raise ValueError("tensor-core engine does not support this shape / filter")
        if use_tc:
            mir, stats = st.prepared(self.metric)
            d, i, self._last_uncertified = ops.exact_topk_tc(q, st.vecs, mir, stats, k, self.metric, ids=st.ids,
                                                            n_rows=st.n, return_uncertified=True)
            return d, i
        return ops.exact_topk_f32(q, st.vecs, k, self.metric, ids=st.ids, row_mask=mask, n_rows=st.n)

    def search(self, queries: np.ndarray, k: int, filter_ids=None) -> Tuple[np.ndarray, np.ndarray]:
        """Host in / host out, one blocking call.  Queries already in PINNED host memory (a pinned torch tensor, or a numpy
        view of one) are copied to the device straight from there; pageable queries go through a pinned staging buffer in
        chunks (the DMA of one chunk runs under the host copy of the next).  The results come back through pinned buffers
        with one synchronisation."""
        if isinstance(queries, torch.Tensor):
            src = queries.detach().to(torch.float32).reshape(-1, self.dim).contiguous()
        else:
            src = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dim))
        nq = src.shape[0]
        dev = self.store.device
        q = torch.empty((nq, self.dim), dtype=torch.float32, device=dev)
        if src.is_pinned():
            q.copy_(src, non_blocking=True)
        else:
            if self._pin is None or self._pin.shape[0] < nq:
                self._pin = torch.empty((max(nq, 64), self.dim), dtype=torch.float32).pin_memory()
            step = max(256, -(-nq // 4))
            for s in range(0, nq, step):
                e = min(nq, s + step)
                self._pin[s:e].copy_(src[s:e])
                q[s:e].copy_(self._pin[s:e], non_blocking=True)
        d, i = self.search_device(q, k, filter_ids)
        po = getattr(self, "_pin_out", None)
        if po is None or po[0].shape[0] < nq or po[0].shape[1] != k:
            po = self._pin_out = (torch.empty((max(nq, 64), k), dtype=torch.float32).pin_memory(),
                                  torch.empty((max(nq, 64), k), dtype=torch.int64).pin_memory())
        po[0][:nq].copy_(d, non_blocking=True)
        po[1][:nq].copy_(i, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return po[0][:nq].numpy().copy(), po[1][:nq].numpy().copy()

    def search_batches(self, batches: Iterable, k: int, post: Optional[Callable] = None, depth: int = 2
                       ) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
        """Host in / host out over a STREAM of query batches, software-pipelined: the H2D copy of batch n+1 and the
        D2H copy of batch n-1 run on their own CUDA streams under the kernels of batch n, so a serving loop pays
        max(copy, compute) per batch instead of their sum.  `batches` yields (nq, dim) float32 arrays (numpy, or
        pinned torch tensors, which are copied from directly); results come back in order, `depth - 1` batches
        behind the submissions.  `post(d, i) -> (d, i)`, if given, runs on the compute stream after each search
        (the multi-GPU shard merge).  Every batch's input and output still cross PCIe -- nothing is cached."""
        dev = self.store.device
        cur = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_streams", None) is None:
            self._copy_streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        s_in, s_out = self._copy_streams
        slots = [dict(q=None, pin=None, d=None, i=None, dh=None, ih=None, nq=0, busy=False,
                      h2d=torch.cuda.Event(), done=torch.cuda.Event(), d2h=torch.cuda.Event()) for _ in range(max(1, depth))]

        def collect(slot):
            slot["d2h"].synchronize()
            slot["busy"] = False
            nq = slot["nq"]
            out = (slot["dh"][:nq].numpy().copy(), slot["ih"][:nq].numpy().copy())
            slot["d"] = slot["i"] = None
            return out

        count = 0
        for b in batches:
            slot = slots[count % len(slots)]
            if slot["busy"]:
                yield collect(slot)
            if isinstance(b, torch.Tensor) and b.is_pinned() and b.dtype == torch.float32 and b.is_contiguous():
                src = b.reshape(-1, self.dim)
            else:
                host = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, self.dim)
                if slot["pin"] is None or slot["pin"].shape[0] < host.shape[0]:
                    slot["pin"] = torch.empty((max(host.shape[0], 64), self.dim), dtype=torch.float32).pin_memory()
                src = slot["pin"][:host.shape[0]]
                src.copy_(torch.from_numpy(host))
            nq = src.shape[0]
            if slot["q"] is None or slot["q"].shape[0] < nq:
                slot["q"] = torch.empty((nq, self.dim), dtype=torch.float32, device=dev)
            with torch.cuda.stream(s_in):
                s_in.wait_event(slot["done"])              # the slot's previous search has consumed its query buffer
                slot["q"][:nq].copy_(src, non_blocking=True)
                slot["h2d"].record(s_in)
            cur.wait_event(slot["h2d"])
            d, i = self.search_device(slot["q"][:nq], k)
            if post is not None:
                d, i = post(d, i)
            slot["done"].record(cur)
            if slot["dh"] is None or slot["dh"].shape[0] < nq or slot["dh"].shape[1] != d.shape[1]:
                slot["dh"] = torch.empty((nq, d.shape[1]), dtype=d.dtype).pin_memory()
                slot["ih"] = torch.empty((nq, i.shape[1]), dtype=i.dtype).pin_memory()
            with torch.cuda.stream(s_out):
                s_out.wait_event(slot["done"])
                d.record_stream(s_out)
                i.record_stream(s_out)
                slot["dh"][:nq].copy_(d, non_blocking=True)
                slot["ih"][:nq].copy_(i, non_blocking=True)
                slot["d2h"].record(s_out)
            slot.update(d=d, i=i, nq=nq, busy=True)
            count += 1
        for m in range(len(slots)):                        # drain in submission order
            slot = slots[(count + m) % len(slots)]
            if slot["busy"]:
                yield collect(slot)
