"""Product quantisation with the reference's service API (src/vector_db/inference/pq.py) on the GPU.

``ProductQuantizationService(k, chunks, dim)`` keeps the reference's constructor, validation
messages, ``compress`` contract (int64 ``(N, chunks)`` codes, ``self.centroids`` = list of
``(k, subdim)`` float32 arrays) and the private helpers its tests call by name
(tests/inference/test_pq.py:44,67,89,113).  Training runs Lloyd rounds on the GPU
(``vdb_pq_kmeans_step``); encoding runs ``vdb_pq_encode`` and is bit-exact with the reference's
``argmin(np.linalg.norm(centroids - chunk, axis=1))`` for the same centroids.

``PQIndex`` is the additive part the reference only plumbs (``pq_chunks``, api/models.py:18): uint8
codes resident in HBM, ADC scan top-R, optional exact fp32 rerank of the shortlist.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _lib, ops

_KMEANS_ITERS = 100   # kmeans2(..., iter=100), pq.py:11


def _stack_codebooks(centroids: Sequence[np.ndarray], device) -> torch.Tensor:
    cb = np.stack([np.asarray(c, dtype=np.float32) for c in centroids])
    return torch.from_numpy(np.ascontiguousarray(cb)).to(device)


class ProductQuantizationService:
    def __init__(self, k: int, chunks: int, dim: int):
        if k <= 0:
            raise ValueError("k must be greater than 0")
        if chunks <= 0:
            raise ValueError("chunks must be greater than 0")
        if dim <= 0:
            raise ValueError("dim must be greater than 0")
        if dim % chunks != 0:
            raise ValueError("dim must be divisible by chunks")
        self.k = k
        self.chunks = chunks
        self.dim = dim
        self.subdim = dim // chunks
        self.centroids: Optional[List[np.ndarray]] = None

    # -- helpers with the reference's names ---------------------------------------------------
    def _validate_embeddings(self, embeddings: np.ndarray) -> None:
        if not isinstance(embeddings, np.ndarray):
            raise TypeError("Embeddings must be a numpy array")
        if embeddings.ndim != 2:
            raise ValueError(f"Embeddings must be 2D array, got {embeddings.ndim}D")
        if embeddings.shape[1] != self.dim:
            raise ValueError(f"Embedding dimension must be {self.dim}, got {embeddings.shape[1]}")

    def _chunk_embeddings(self, embeddings: np.ndarray) -> List[List[np.ndarray]]:
        """(N,D) -> ``chunks`` lists of N sub-vectors; sub-space j = columns [j*subdim,(j+1)*subdim)
        (pq.py:48-62).  Host-side view only; the kernels read the (N,D) matrix in place."""
        n, d = embeddings.shape
        assert d == self.subdim * self.chunks
        per_chunk = embeddings.reshape(n, self.chunks, self.subdim).transpose(1, 0, 2)
        return [list(c) for c in per_chunk]

    def _compute_centroids(self, chunks) -> List[np.ndarray]:
        """``chunks`` independent k-means runs, all sub-spaces per launch (pq.py:64-75 ran one
        scipy kmeans2 per sub-space in a process pool).  Random-points init (minit='points'),
        exactly 100 Lloyd rounds, an empty cluster keeps its centroid."""
        x = np.stack([np.asarray(c, dtype=np.float32) for c in chunks], axis=1)     # (N, chunks, subdim)
        x = np.ascontiguousarray(x.reshape(x.shape[0], self.dim))
        return self.fit(x)

    def fit(self, embeddings: np.ndarray, init_rows: Optional[np.ndarray] = None,
            iters: int = _KMEANS_ITERS) -> List[np.ndarray]:
        n = embeddings.shape[0]
        if n < self.k:
            raise ValueError("Cannot take a larger sample than population when 'replace=False'")
        dev = _lib.require_cuda()
        if init_rows is None:
            init_rows = np.random.choice(n, size=int(self.k), replace=False)
        x = torch.from_numpy(np.ascontiguousarray(embeddings, dtype=np.float32)).to(dev)
        init = x[torch.from_numpy(np.asarray(init_rows, dtype=np.int64)).to(dev)]      # (k, D)
        cb = init.view(self.k, self.chunks, self.subdim).permute(1, 0, 2).contiguous()  # (chunks, k, subdim)
        for _ in range(iters):
            ops.pq_kmeans_step(x, cb)
        host = cb.cpu().numpy()
        self.centroids = [host[j].copy() for j in range(self.chunks)]
        return self.centroids

    def _find_nearest_centroid(self, chunk: np.ndarray, centroids: np.ndarray) -> int:
        dev = _lib.require_cuda()
        c = np.ascontiguousarray(centroids, dtype=np.float32)
        x = torch.from_numpy(np.ascontiguousarray(chunk, dtype=np.float32).reshape(1, -1)).to(dev)
        cb = torch.from_numpy(c[None, :, :]).to(dev)
        return int(ops.pq_encode(x, cb)[0, 0].item())

    def _compress_embedding(self, embedding: np.ndarray, centroids: List[np.ndarray]) -> np.ndarray:
        dev = _lib.require_cuda()
        x = torch.from_numpy(np.ascontiguousarray(embedding, dtype=np.float32).reshape(1, -1)).to(dev)
        codes = ops.pq_encode(x, _stack_codebooks(centroids, dev))
        return codes[0].cpu().numpy().astype(np.int64)

    # -- public ---------------------------------------------------------------------------------
    def encode(self, embeddings: np.ndarray) -> np.ndarray:
        """Codes for new rows with the trained centroids, int64 ``(N, chunks)``."""
        if self.centroids is None:
            raise ValueError("centroids not trained; call compress() or fit() first")
        self._validate_embeddings(embeddings)
        dev = _lib.require_cuda()
        x = torch.from_numpy(np.ascontiguousarray(embeddings, dtype=np.float32)).to(dev)
        return ops.pq_encode(x, _stack_codebooks(self.centroids, dev)).cpu().numpy().astype(np.int64)

    def compress(self, embeddings: np.ndarray) -> np.ndarray:
        """Train on ``embeddings`` and encode them (pq.py:91-108)."""
        self._validate_embeddings(embeddings)
        self.fit(np.ascontiguousarray(embeddings, dtype=np.float32))
        return self.encode(embeddings)


class PQIndex:
    """uint8 PQ codes in HBM + ADC search; optionally keeps the fp32 rows for an exact rerank."""

    def __init__(self, centroids: Sequence[np.ndarray], keep_vectors: bool = True, device=None):
        self.device = _lib.require_cuda() if device is None else torch.device(device)
        cb = np.stack([np.asarray(c, dtype=np.float32) for c in centroids])          # (m,K,dsub)
        self.m, self.K, self.dsub = cb.shape
        if self.K > 256:
            raise ValueError("PQIndex supports at most 256 centroids per sub-space (nbits=8)")
        self.dim = self.m * self.dsub
        self.codebooks = torch.from_numpy(np.ascontiguousarray(cb)).to(self.device)
        if self.K < 256:      # pad to 256 entries so LUT rows have the scan kernel's stride; codes never
            # reference the padding, which repeats centroid 0 so it cannot disturb the LUT's fixed-point scale
            pad = self.codebooks[:, :1, :].expand(self.m, 256 - self.K, self.dsub)
            self._cb_lut = torch.cat([self.codebooks, pad], dim=1).contiguous()
        else:
            self._cb_lut = self.codebooks
        self.keep_vectors = keep_vectors
        self.codes = torch.empty((0, self.m), dtype=torch.uint8, device=self.device)
        self.vecs = torch.empty((0, self.dim), dtype=torch.float32, device=self.device) if keep_vectors else None
        self.ids = torch.empty((0,), dtype=torch.int64, device=self.device)
        self.n = 0

    def add_device(self, vecs: torch.Tensor, ids: Optional[torch.Tensor] = None) -> None:
        vecs = vecs.to(self.device, torch.float32).contiguous()
        codes = ops.pq_encode(vecs, self.codebooks)
        if ids is None:
            ids = torch.arange(self.n, self.n + vecs.shape[0], dtype=torch.int64, device=self.device)
        self.codes = torch.cat([self.codes, codes]) if self.n else codes
        self.ids = torch.cat([self.ids, ids.to(self.device)]) if self.n else ids.to(self.device)
        if self.keep_vectors:
            self.vecs = torch.cat([self.vecs, vecs]) if self.n else vecs
        self.n += vecs.shape[0]

    def add(self, vectors: np.ndarray, ids: Optional[Sequence[int]] = None) -> None:
        v = torch.from_numpy(np.ascontiguousarray(vectors, dtype=np.float32)).to(self.device)
        i = None if ids is None else torch.as_tensor(np.asarray(ids, dtype=np.int64), device=self.device)
        self.add_device(v, i)

    def bitmap_from_ids(self, ids) -> torch.Tensor:
        """Row bitmap (int32 words) with the bits of the given external ids set -- the metadata pre-filter
        (filter_ids of src/vector_db/indexing/hnsw.py:96,127) for the ADC scan."""
        want = torch.as_tensor(sorted(int(i) for i in ids), dtype=torch.int64, device=self.device)
        hit = torch.isin(self.ids[:self.n], want)
        bits = torch.zeros(((self.n + 31) // 32) * 32, dtype=torch.int64, device=self.device)
        bits[:self.n] = hit.to(torch.int64)
        words = (bits.view(-1, 32) << torch.arange(32, device=self.device, dtype=torch.int64)).sum(dim=1)
        return torch.where(words >= 2 ** 31, words - 2 ** 32, words).to(torch.int32)

    def search_device(self, q: torch.Tensor, k: int, shortlist: Optional[int] = None, rerank: bool = False,
                      filter_ids=None):
        """ADC top-``shortlist`` (default k); with ``rerank`` the shortlist is re-scored with the exact
        fp32 L2 distance and the k best returned.  Returns (dist, ids) device tensors; ADC
        distances are SQUARED sub-space sums, reranked distances are non-squared L2."""
        q = q.to(self.device, torch.float32).contiguous()
        R = max(k, shortlist or k)
        R = min(R, max(self.n, 1))
        lut = ops.pq_lut(q, self._cb_lut)
        mask = self.bitmap_from_ids(filter_ids) if filter_ids is not None else None
        d, rows = ops.pq_adc_topk(self.codes, lut, R, ids=None, n_rows=self.n, row_mask=mask)
        if rerank:
            if not self.keep_vectors:
                raise ValueError("rerank needs keep_vectors=True")
            return ops.rerank_f32(q, self.vecs, rows, min(k, R), ids=self.ids, n_rows=self.n)
        ids = torch.where(rows >= 0, self.ids[rows.clamp(min=0)], rows)
        return d[:, :k], ids[:, :k]

    def search(self, queries: np.ndarray, k: int, shortlist: Optional[int] = None, rerank: bool = False,
               filter_ids=None) -> Tuple[np.ndarray, np.ndarray]:
        q = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dim)).to(self.device)
        d, i = self.search_device(q, k, shortlist, rerank, filter_ids)
        return d.cpu().numpy(), i.cpu().numpy()
