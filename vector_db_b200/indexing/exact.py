"""Exact brute-force index: batched L2 / cosine top-k over rows resident in HBM.

The reference has no brute-force search API; it has the semantics (the per-pair loop + tuple
sort of tests/indexing/test_hnsw.py:128-134 and IvfIndex.search with n_probe == k,
src/vector_db/indexing/ivf.py:170-199).  This class is that scan as a batched GPU operator:
results are ordered by (distance, id) ascending for L2 (non-squared distances) and by
(-similarity, id) for cosine, padded with (+/-inf, -1).
"""
from __future__ import annotations

from typing import Optional, Sequence, Set, Tuple

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
    def add(self, ids: Sequence[int], vectors: np.ndarray) -> None:
        self.store.append(ids, vectors)

    def add_device(self, ids: torch.Tensor, vecs: torch.Tensor) -> None:
        self.store.append_device(ids, vecs)

    def __len__(self) -> int:
        return self.store.n

    # -- search -----------------------------------------------------------------------------------
    def search_device(self, q: torch.Tensor, k: int, filter_ids: Optional[Set[int]] = None
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
        st = self.store
        mask = st.bitmap_from_ids(filter_ids) if filter_ids is not None else None
        if mask is None and st.n_dead:
            mask = st.alive
        use_tc = (self.engine in ("auto", "tc") and mask is None and ops.tc_supported(q.shape[0], st.n, self.dim, k, self.metric))
        if self.engine == "tc" and not use_tc:
            raise ValueError("tensor-core engine does not support this shape / filter")
        if use_tc:
            mir, max_norm = st.prepared(self.metric)
            d, i, self._last_uncertified = ops.exact_topk_tc(q, st.vecs, mir, max_norm, k, self.metric, ids=st.ids,
                                                            n_rows=st.n, return_uncertified=True)
            return d, i
        return ops.exact_topk_f32(q, st.vecs, k, self.metric, ids=st.ids, row_mask=mask, n_rows=st.n)

    def search(self, queries: np.ndarray, k: int, filter_ids: Optional[Set[int]] = None
               ) -> Tuple[np.ndarray, np.ndarray]:
        """Host in / host out: pinned staging, one H2D copy, kernels, one D2H copy."""
        q_host = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dim)
        nq = q_host.shape[0]
        dev = self.store.device
        if self._pin is None or self._pin.shape[0] < nq:
            self._pin = torch.empty((max(nq, 64), self.dim), dtype=torch.float32).pin_memory()
        self._pin[:nq].copy_(torch.from_numpy(q_host))
        q = self._pin[:nq].to(dev, non_blocking=True)
        d, i = self.search_device(q, k, filter_ids)
        return d.cpu().numpy(), i.cpu().numpy()
