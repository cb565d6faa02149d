"""IVF index with the reference's API (src/vector_db/indexing/ivf.py), lists and vectors in HBM.

Same constructor, methods, attributes, error messages and pickle schema as the reference class
(ivf.py:13-251): ``build_index / add / delete / search / get_cluster_size / get_cluster_stats /
save_index / load_index`` and ``k, storage, index_file, centroids, inverted_lists``.  What changes is
where the work happens:

  * embeddings live in HBM (device_store.DeviceVectors, row = insertion order) and the inverted lists
    are a CSR pair ``list_ptr`` int32 [k+1] / ``list_rows`` int32 [n] of ROW indices, each list in the
    reference's append order; ``inverted_lists`` (lists of node ids) is materialised on demand;
  * ``search`` ranks the centroids with numpy's exact arithmetic (``vdb_ivf_centroid_dist``), takes the
    ``n_probe`` nearest and scans their lists on the GPU (``vdb_ivf_scan``) in the reference's stable
    candidate order; ``search_batch`` is the many-query form;
  * ``build_index`` trains with exactly 100 Lloyd rounds like ``kmeans2(iter=100, minit='points')``
    (ivf.py:75): initial centroids drawn with the same global-numpy-RNG call scipy makes, assignment by
    the exact GPU scan, mean update by ``vdb_ivf_centroid_mean``; the lists come from the assignment
    made BEFORE the last update, as scipy returns it.  Values are parity-unpinned.
"""
from __future__ import annotations

import pickle
import threading
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from .. import _lib, ops
from ..device_store import DeviceVectors
from ..storage import InMemoryNodeStorage, NodeStorage
from ..types import Node

_ASSIGN_CHUNK = 65536      # rows per assignment call during training


class IvfIndex:
    def __init__(self, k: int, storage: Optional[NodeStorage] = None, index_file: Optional[Union[str, Path]] = None,
                 device=None):
        if k <= 0:
            raise ValueError("k-means parameter should be positive")
        self.k = k
        self.storage = storage or InMemoryNodeStorage()
        self.index_file = Path(index_file) if index_file else None
        self.centroids: Optional[np.ndarray] = None
        self._device = _lib.require_cuda() if device is None else torch.device(device)
        self._lock = threading.RLock()
        self._vectors: Optional[DeviceVectors] = None
        self._centroids_dev: Optional[torch.Tensor] = None
        self._lists: Optional[List[List[int]]] = []      # host lists of ids; None = not materialised from the CSR
        self._list_ptr: Optional[torch.Tensor] = None    # device CSR, valid when not _csr_dirty
        self._list_rows: Optional[torch.Tensor] = None
        self._csr_dirty = True
        self.last_scanned: Optional[np.ndarray] = None   # candidates enumerated per query by the last search
        self.batch_engine = "auto"                       # "auto" | "tc" (list-major tensor-core batch search) | "scan"
        self.last_uncertified = 0
        self._lm = None
        if self.index_file and self.index_file.exists():
            self.load_index()

    # ------------------------------------------------------------------ lists <-> CSR
    @property
    def inverted_lists(self) -> List[List[int]]:
        """cluster_id -> list of node ids (ivf.py:42), materialised from the device CSR when needed."""
        if self._lists is None:
            ptr = self._list_ptr.cpu().numpy()
            ids = self._vectors.ids[self._list_rows.long()].cpu().numpy()
            self._lists = [ids[ptr[c]:ptr[c + 1]].tolist() for c in range(len(ptr) - 1)]
        return self._lists

    @inverted_lists.setter
    def inverted_lists(self, value: List[List[int]]) -> None:
        self._lists = value
        self._csr_dirty = True

    def _csr(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self._csr_dirty:
            lists = self._lists
            row_of = self._vectors.id_to_row if self._vectors is not None else {}
            ptr = np.zeros(len(lists) + 1, dtype=np.int64)
            rows: List[int] = []
            for c, lst in enumerate(lists):
                rows.extend(r for r in (row_of.get(int(i)) for i in lst) if r is not None)   # ivf.py:188-190
                ptr[c + 1] = len(rows)
            self._list_ptr = torch.from_numpy(ptr.astype(np.int32)).to(self._device)
            self._list_rows = torch.tensor(rows if rows else [0], dtype=torch.int32, device=self._device)
            self._csr_dirty = False
            self._lm = None
        return self._list_ptr, self._list_rows

    # ------------------------------------------------------------------ list-major copy for the tensor-core batch search
    def _list_major(self) -> dict:
        """Rows in list-major order, every list padded to whole 256-row tiles, with their L2 operand mirror
        (vdb_ivf_topk_tc).  Built on first use after a change of the lists; costs one more copy of the rows in HBM."""
        ptr, rows = self._csr()
        lm = getattr(self, "_lm", None)
        if lm is not None and lm["ptr"] is ptr:
            return lm
        dev, vec = self._device, self._vectors
        ptr64 = ptr.long()
        counts = ptr64[1:] - ptr64[:-1]
        padded = (counts + 255) // 256 * 256
        start = torch.zeros(self.k + 1, dtype=torch.int64, device=dev)
        start[1:] = torch.cumsum(padded, 0)
        n_pad = int(start[-1].item())
        n = int(ptr64[-1].item())
        lm_row = torch.full((max(n_pad, 256),), -1, dtype=torch.int64, device=dev)
        if n:
            list_of = torch.repeat_interleave(torch.arange(self.k, device=dev), counts)
            pos = start[list_of] + (torch.arange(n, device=dev) - ptr64[list_of])
            lm_row[pos] = rows[:n].long()
        n_pad = lm_row.numel()
        src = lm_row.clamp(min=0)
        lm_vecs = vec.vecs[src].contiguous()
        # the operand mirror holds RESIDUALS against the list's centroid (padding rows: zero, so the measured stats stay real)
        lm_list = torch.searchsorted(start, torch.arange(n_pad, device=dev), right=True) - 1
        res = lm_vecs - self._centroids_dev[lm_list.clamp(max=self.k - 1)]
        res[lm_row < 0] = 0
        mirror, _, stats = ops.tc_prepare_rows(res, "l2")
        # per-list row bounds for the certificate (the global pair `stats` is the maximum over ALL lists: one wide list
        # would loosen every query's allowance); 1.0001 covers the fp32 rounding of the norms computed here
        rb = res.to(torch.bfloat16).to(torch.float32)
        per_row = torch.stack([rb.norm(dim=1), (rb - res).norm(dim=1)], 1) * 1.0001
        del res, rb
        list_stats = torch.zeros((self.k, 2), dtype=torch.float32, device=dev)
        list_stats.scatter_reduce_(0, lm_list.clamp(max=self.k - 1).unsqueeze(1).expand(-1, 2), per_row, "amax")
        pad = (lm_row < 0).nonzero().squeeze(1)
        if pad.numel():                                               # a padding row can never be selected: |x|^2 := 1e30
            off = ops.tc_norm_offset(n_pad, vec.dim)
            mirror[off + pad * 16 + (((pad >> 2) & 1) << 3)] = 1e30
        lm_ids = torch.where(lm_row >= 0, vec.ids[src], torch.full_like(lm_row, -1))
        tile_ptr = (start // 256).to(torch.int32).contiguous()          # first 256-row tile of every list (empty lists: none)
        self._lm = {"ptr": ptr, "vecs": lm_vecs, "mirror": mirror, "stats": stats, "list_stats": list_stats, "ids": lm_ids,
                    "tile_ptr": tile_ptr, "list_of_row": lm_list.clamp(max=self.k - 1).to(torch.int32).contiguous(), "any_rows": n > 0,
                    "start": start,
                    "counts": counts, "padded": padded}
        return self._lm

    def _search_list_major(self, q: torch.Tensor, probes: torch.Tensor, prefix: torch.Tensor, top_k: int):
        """Batch search through the tensor cores: (query, probed list) pairs grouped by list (vdb_ivf_group_pairs), one small
        dense contraction per list (vdb_ivf_topk_tc), selection / exact fp32 rerank / certificate per query, then the
        reference's stable order (distance, enumeration sequence -- ivf.py:177-193) among equal distances
        (vdb_ivf_tc_finish) -- all on the device, one read-back (the list of queries to re-run).  Returns (dist, ids, bad):
        ``bad`` = queries the certificate rejected or whose tie run crosses the cut; the caller re-runs them through the
        query-major scan."""
        lm = self._list_major()
        Q, n_probe = probes.shape
        k2 = top_k + 1                # one extra result: a tie run that crosses the cut is SEEN and sent to the query-major scan
        if not bool(lm["any_rows"]):
            return None
        row_query, row_list, pair_rows, utab, meta = ops.ivf_group_pairs(probes, lm["tile_ptr"])
        d, pos, uncert = ops.ivf_topk_tc(q, lm["vecs"], lm["mirror"], lm["stats"], None, k2, n_probe, pair_rows, row_query,
                                         row_list, self._centroids_dev, utab, lm["list_stats"], meta=meta)
        d, ids, tie = ops.ivf_tc_finish(d, pos, top_k, probes, prefix, lm["list_of_row"], lm["start"], lm["ids"])
        # the ONE read-back: which queries go to the query-major scan (uncert = count + indices, tie = flag per query)
        live = torch.arange(Q, device=q.device, dtype=torch.int32) < uncert[0]
        tie.index_add_(0, torch.where(live, uncert[1:], torch.zeros_like(uncert[1:])).long(), live.to(torch.int32))
        bad = (tie > 0).nonzero().squeeze(1)
        return d, ids, (bad if bad.numel() else None)

    def _probe(self, q: torch.Tensor, n_probe: int, ptr: torch.Tensor):
        """np.argsort(np.linalg.norm(centroids - q, axis=1))[:n_probe] per query (ivf.py:170-171) + candidate prefix sums.
        Batches with few probes go through the tensor cores (candidates re-scored in numpy's order, ops.ivf_probe_tc)."""
        cent = self._centroids_dev
        Q, D = q.shape
        K = cent.shape[0]
        kc = n_probe + 4
        if (Q >= 512 and kc <= 12 and 256 <= K <= 4096 and D % 8 == 0 and D <= 8192
                and ops.tc_shape_supported(Q, K, D, kc, "l2")):
            cm = getattr(self, "_cent_mirror", None)
            if cm is None or cm["cent"] is not cent:
                mirror, _, stats = ops.tc_prepare_rows(cent, "l2")
                cm = self._cent_mirror = {"cent": cent, "mirror": mirror, "stats": stats, "use": None, "calls": 0}
            # Whether the coarse pass can separate the probe boundary depends on the geometry (centroids of unclustered rows
            # sit near the origin: all distances ~ |q|, the bf16 certificate fails and the scan's fp32 rescan costs more than
            # the numpy-order kernels).  Decided by measurement: the first eligible batch (and every 256th) reads back how many
            # queries needed the rescan / the fallback; above 2 % the tensor-core probing is switched off for this index.
            if cm["use"] is not False:
                probes, prefix, flags, n_unc = ops.ivf_probe_tc(q, cent, cm["mirror"], cm["stats"], n_probe, ptr)
                if cm["use"] is None or cm["calls"] % 256 == 0:
                    slow = int(n_unc.item()) + int(flags.sum().item())
                    cm["use"] = slow * 50 <= Q
                cm["calls"] += 1
                return probes, prefix
        return ops.ivf_probe(ops.ivf_centroid_dist(q, cent), n_probe, ptr)

    def _set_centroids(self, cent: torch.Tensor) -> None:
        self._centroids_dev = cent.contiguous()
        self.centroids = cent.cpu().numpy()

    # ------------------------------------------------------------------ build
    def build_index(self, nodes: List[Node]) -> None:
        """ivf.py:49-85: store the nodes, cluster their embeddings, fill the inverted lists."""
        if not nodes:
            raise ValueError("Cannot build index with empty node list")
        for node in nodes:
            self.storage.save(node)
        embeddings = np.array([node.embedding for node in nodes])
        ids = [node.id for node in nodes]
        if embeddings.ndim != 2:
            raise ValueError(f"embeddings must be 2D array, got {embeddings.ndim}D")
        if embeddings.shape[0] < self.k:
            raise ValueError(f"Need at least {self.k} vectors for {self.k} clusters")
        self.build_from_arrays(ids, embeddings)
        if self.index_file:
            self.save_index()

    def build_from_arrays(self, ids: Sequence[int], vectors: np.ndarray, iters: int = 100,
                          init_rows: Optional[np.ndarray] = None) -> None:
        """Bulk form of build_index: no per-node storage writes.  ``init_rows`` replaces the random
        'points' initialisation (rows of ``vectors`` used as the initial centroids)."""
        with self._lock:
            X = np.ascontiguousarray(vectors, dtype=np.float32)
            n, dim = X.shape
            if n < self.k:
                raise ValueError(f"Need at least {self.k} vectors for {self.k} clusters")
            self._vectors = DeviceVectors(dim, self._device, capacity=n)
            rows = self._vectors.append(ids, X)
            row_t = torch.from_numpy(np.asarray(rows, dtype=np.int64)).to(self._device)
            identity = np.array_equal(rows, np.arange(n))          # no id appeared twice
            Xd = self._vectors.vecs[:n] if identity else self._vectors.vecs.index_select(0, row_t)
            if init_rows is None:
                # scipy.cluster.vq._kpoints: rng.choice(n, size=k, replace=False) on the global numpy RandomState
                init_rows = np.random.mtrand._rand.choice(n, size=int(self.k), replace=False)
            cent = Xd.index_select(0, torch.from_numpy(np.asarray(init_rows, dtype=np.int64)).to(self._device)).clone()
            labels = None
            for _ in range(iters):
                labels = self._assign(Xd, cent)
                ptr, order = self._labels_to_csr(labels)
                ops.ivf_centroid_mean(Xd, ptr, order, cent)
            if labels is None:
                labels = self._assign(Xd, cent)
            ptr, order = self._labels_to_csr(labels)          # the assignment made before the last update
            self._list_ptr = ptr
            self._list_rows = order if identity else row_t.index_select(0, order.long()).to(torch.int32)
            self._lists = None
            self._csr_dirty = False
            self._set_centroids(cent)

    def _assign(self, X: torch.Tensor, cent: torch.Tensor) -> torch.Tensor:
        """Nearest centroid of every row -> int32 labels (exact fp32 distances, first minimum wins)."""
        n, dim = X.shape
        K = cent.shape[0]
        out = torch.empty((n,), dtype=torch.int32, device=X.device)
        use_tc = ops.tc_supported(min(n, _ASSIGN_CHUNK), K, dim, 1, "l2")
        if use_tc:
            mirror, _, stats = ops.tc_prepare_rows(cent, "l2")
        for s in range(0, n, _ASSIGN_CHUNK):
            e = min(n, s + _ASSIGN_CHUNK)
            if use_tc and ops.tc_supported(e - s, K, dim, 1, "l2"):
                _, i = ops.exact_topk_tc(X[s:e], cent, mirror, stats, 1, "l2")
            else:
                _, i = ops.exact_topk_f32(X[s:e], cent, 1, "l2")
            out[s:e] = i[:, 0].to(torch.int32)
        return out

    def _labels_to_csr(self, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        order = torch.sort(labels.long(), stable=True).indices.to(torch.int32)      # ascending row inside a label
        counts = torch.bincount(labels.long(), minlength=self.k)
        ptr = torch.zeros((self.k + 1,), dtype=torch.int32, device=labels.device)
        ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
        return ptr, order

    # ------------------------------------------------------------------ mutation
    def add(self, node: Node) -> None:
        """ivf.py:87-117."""
        if self.centroids is None:
            raise ValueError("Index must be built before adding nodes")
        embedding = node.embedding
        if embedding.ndim != 1:
            raise ValueError("embedding must be 1D array")
        if embedding.shape[0] != self.centroids.shape[1]:
            raise ValueError(
                f"embedding dimension {embedding.shape[0]} doesn't match "
                f"centroid dimension {self.centroids.shape[1]}"
            )
        self.storage.save(node)
        with self._lock:
            if self._vectors is None:
                self._vectors = DeviceVectors(self.centroids.shape[1], self._device)
            self._vectors.append([node.id], np.asarray(embedding, dtype=np.float32)[None, :])
            q = torch.from_numpy(np.ascontiguousarray(embedding, dtype=np.float32)[None, :]).to(self._device)
            probes, _ = ops.ivf_probe(ops.ivf_centroid_dist(q, self._centroids_dev), 1)
            self.inverted_lists[int(probes[0, 0].item())].append(node.id)
            self._csr_dirty = True
        if self.index_file:
            self.save_index()

    def delete(self, node_id: int) -> None:
        """ivf.py:119-137."""
        with self._lock:
            for cluster_list in self.inverted_lists:
                if node_id in cluster_list:
                    cluster_list.remove(node_id)
            self._csr_dirty = True
            if self._vectors is not None:
                self._vectors.mark_dead(node_id)
        if hasattr(self.storage, "delete"):
            self.storage.delete(node_id)
        if self.index_file:
            self.save_index()

    # ------------------------------------------------------------------ search
    def search_batch(self, queries: np.ndarray, n_probe: int, top_k: int, filter_ids=None) -> Tuple[np.ndarray, np.ndarray]:
        """Many queries at once: (dist f32 [Q,top_k], ids i64 [Q,top_k]), padded with (+inf, -1).
        ``filter_ids``: optional set of node ids the results are restricted to (metadata pre-filter)."""
        if self.centroids is None:
            raise ValueError("Index must be built before searching")
        queries = np.ascontiguousarray(queries, dtype=np.float32)
        if queries.ndim != 2:
            raise ValueError("queries must be 2D array")
        if queries.shape[1] != self.centroids.shape[1]:
            raise ValueError(
                f"query dimension {queries.shape[1]} doesn't match "
                f"centroid dimension {self.centroids.shape[1]}"
            )
        if n_probe <= 0 or n_probe > self.k:
            raise ValueError(f"n_probe must be between 1 and {self.k}")
        with self._lock:
            q = torch.from_numpy(queries).to(self._device)
            d, i, scanned = self.search_device(q, n_probe, top_k, filter_ids=filter_ids)
            self.last_scanned = scanned.cpu().numpy()
            return d.cpu().numpy(), i.cpu().numpy()

    def search_device(self, q: torch.Tensor, n_probe: int, top_k: int, filter_ids=None):
        """Device in / device out: (dist, ids, candidates enumerated per query)."""
        ptr, rows = self._csr()
        vec = self._vectors
        probes, prefix = self._probe(q, n_probe, ptr)
        alive = vec.alive if vec.n_dead else None
        if filter_ids is not None:
            flt = vec.bitmap_from_ids(filter_ids)
            alive = flt if alive is None else torch.bitwise_and(alive[:flt.numel()], flt)
        Q = q.shape[0]
        # the list-major pass costs ~0.5 ms whatever the batch (grouping, operand rows, select / rerank / finish); the query-major
        # scan streams the probed rows at 1-5 TB/s (L2-served): "auto" takes the tensor cores from ~3 GB of probed rows on
        probed_bytes = Q * n_probe * (vec.n / max(self.k, 1)) * vec.dim * 4
        worth = self.batch_engine == "tc" or (n_probe >= 2 and probed_bytes >= 3e9)
        if (self.batch_engine != "scan" and filter_ids is None and alive is None and n_probe <= 32 and top_k <= 10
                and vec.dim % 8 == 0 and vec.dim <= 512 and worth):
            out = self._search_list_major(q, probes, prefix, top_k)
            if out is not None:
                d, ids, bad = out
                if bad is not None:                                   # the certificate rejected these: query-major exact scan
                    db, ib = ops.ivf_scan(q[bad], vec.vecs, ptr, rows, probes[bad], prefix[bad], top_k, ids=vec.ids)
                    d[bad] = db
                    ids[bad] = ib
                self.last_uncertified = 0 if bad is None else int(bad.numel())
                return d, ids, prefix[:, n_probe].long()
        return ops.ivf_scan(q, vec.vecs, ptr, rows, probes, prefix, top_k, alive=alive, ids=vec.ids, return_scanned=True)

    def search(self, query: np.ndarray, n_probe: int, top_k: int) -> List[Tuple[Node, float]]:
        """ivf.py:139-199: ``[(Node, np.float32 distance)]`` sorted by distance."""
        if self.centroids is None:
            raise ValueError("Index must be built before searching")
        if query.ndim != 1:
            raise ValueError("query must be 1D array")
        if query.shape[0] != self.centroids.shape[1]:
            raise ValueError(
                f"query dimension {query.shape[0]} doesn't match "
                f"centroid dimension {self.centroids.shape[1]}"
            )
        if n_probe <= 0 or n_probe > self.k:
            raise ValueError(f"n_probe must be between 1 and {self.k}")
        if top_k <= 0:
            return []
        d, i = self.search_batch(np.asarray(query)[None, :], n_probe, top_k)
        results = []
        for dist, node_id in zip(d[0], i[0]):
            if node_id < 0:
                break
            node = self.storage.get(int(node_id))
            if node:
                results.append((node, np.float32(dist)))
        return results

    # ------------------------------------------------------------------ stats (ivf.py:201-216)
    def get_cluster_size(self, cluster_id: int) -> int:
        if cluster_id < 0 or cluster_id >= self.k:
            raise ValueError(f"cluster_id must be between 0 and {self.k - 1}")
        return len(self.inverted_lists[cluster_id])

    def get_cluster_stats(self) -> dict:
        sizes = [len(lst) for lst in self.inverted_lists]
        return {
            "min_size": min(sizes),
            "max_size": max(sizes),
            "avg_size": sum(sizes) / len(sizes) if sizes else 0,
            "total_vectors": sum(sizes),
        }

    # ------------------------------------------------------------------ persistence (ivf.py:218-251)
    def save_index(self) -> None:
        if self.index_file is None:
            return
        index_data = {"k": self.k, "centroids": self.centroids, "inverted_lists": self.inverted_lists}
        with open(self.index_file, "wb") as f:
            pickle.dump(index_data, f)

    def load_index(self) -> None:
        if self.index_file is None or not self.index_file.exists():
            return
        with open(self.index_file, "rb") as f:
            index_data = pickle.load(f)
        self.load_state(index_data)

    def load_state(self, index_data: dict, vectors: Optional[Dict[int, np.ndarray]] = None) -> None:
        """Import an index in the reference's pickle schema; embeddings come from ``vectors`` or storage."""
        with self._lock:
            self.k = index_data["k"]
            cent = index_data["centroids"]
            self._lists = [list(lst) for lst in index_data["inverted_lists"]]
            self._csr_dirty = True
            if cent is None:
                self.centroids, self._centroids_dev, self._vectors = None, None, None
                return
            cent = np.ascontiguousarray(cent, dtype=np.float32)
            self._centroids_dev = torch.from_numpy(cent).to(self._device)
            self.centroids = index_data["centroids"]
            ids = [i for lst in self._lists for i in lst]
            get = (lambda i: vectors[i]) if vectors is not None else self.storage.get_embedding
            self._vectors = DeviceVectors(cent.shape[1], self._device, capacity=max(len(ids), 1))
            keep, rows = [], []
            for i in ids:
                try:
                    rows.append(np.asarray(get(i), dtype=np.float32))
                    keep.append(i)
                except KeyError:
                    continue
            if keep:
                self._vectors.append(keep, np.stack(rows))


class IvfPqIndex(IvfIndex):
    """IVF-PQ: the IVF coarse quantiser of :class:`IvfIndex` + product-quantised codes of
    the raw vectors (``ProductQuantizationService`` sub-space layout, src/vector_db/inference/pq.py:48-62),
    stored list-major next to the lists.  ``search_batch`` probes ``n_probe`` lists, takes the ``shortlist``
    best by asymmetric distance (fixed-point ADC, the arithmetic of ``vdb_pq_adc_topk``) and re-scores them with
    the exact fp32 L2 distance.  The reference has the two halves (ivf.py, pq.py) but not their combination."""

    def __init__(self, k: int, chunks: int, pq_k: int = 256, storage: Optional[NodeStorage] = None,
                 index_file: Optional[Union[str, Path]] = None, device=None):
        super().__init__(k, storage=storage, index_file=index_file, device=device)
        if chunks <= 0:
            raise ValueError("k-means parameter and number of chunks should be positive")     # pq.py:24-25
        if pq_k <= 0 or pq_k > 256:
            raise ValueError("pq_k must be between 1 and 256 (uint8 codes)")
        self.chunks = chunks
        self.pq_k = pq_k
        self.codebooks: Optional[torch.Tensor] = None      # [chunks, pq_k, dsub]
        self._cb_lut: Optional[torch.Tensor] = None        # padded to 256 entries per sub-space
        self._codes_lm: Optional[torch.Tensor] = None      # uint8 [n listed rows, chunks], CSR order
        self._codes_for: Optional[torch.Tensor] = None     # the list_rows tensor _codes_lm was built for

    def train_pq(self, iters: int = 25, train_rows: int = 100_000, init_rows: Optional[np.ndarray] = None,
                 seed: int = 777) -> None:
        """Lloyd rounds on a row sample, all sub-spaces per launch (``vdb_pq_kmeans_step``)."""
        vec = self._vectors
        if vec is None or vec.n == 0:
            raise ValueError("Index must be built before training the product quantiser")
        if vec.dim % self.chunks != 0:
            raise ValueError("embedding dimension must be divisible by number of chunks")             # pq.py:28-29
        n = vec.n
        rng = np.random.default_rng(seed)
        sample_rows = rng.choice(n, min(n, train_rows), replace=False)
        sample = vec.vecs.index_select(0, torch.from_numpy(np.sort(sample_rows)).to(self._device)).contiguous()
        if sample.shape[0] < self.pq_k:
            raise ValueError(f"Need at least {self.pq_k} vectors to train {self.pq_k} centroids")
        if init_rows is None:
            init_rows = rng.choice(sample.shape[0], self.pq_k, replace=False)
        dsub = vec.dim // self.chunks
        init = sample.index_select(0, torch.from_numpy(np.asarray(init_rows, dtype=np.int64)).to(self._device))
        cb = init.view(self.pq_k, self.chunks, dsub).permute(1, 0, 2).contiguous()
        for _ in range(iters):
            ops.pq_kmeans_step(sample, cb)
        self.set_codebooks(cb)

    def set_codebooks(self, codebooks) -> None:
        cb = torch.as_tensor(np.asarray(codebooks, dtype=np.float32) if not torch.is_tensor(codebooks) else codebooks)
        cb = cb.to(self._device, torch.float32).contiguous()
        if cb.ndim != 3 or cb.shape[0] != self.chunks or cb.shape[1] != self.pq_k:
            raise ValueError(f"codebooks must have shape ({self.chunks}, {self.pq_k}, dsub)")
        self.codebooks = cb
        if self.pq_k < 256:      # LUT rows have 256 entries; the padding repeats centroid 0 and is never referenced
            pad = cb[:, :1, :].expand(self.chunks, 256 - self.pq_k, cb.shape[2])
            self._cb_lut = torch.cat([cb, pad], dim=1).contiguous()
        else:
            self._cb_lut = cb
        self._codes_lm = None

    def _list_codes(self) -> torch.Tensor:
        ptr, rows = self._csr()
        if self._codes_lm is None or self._codes_for is not rows:
            listed = self._vectors.vecs.index_select(0, rows.long())
            self._codes_lm = ops.pq_encode(listed, self.codebooks)
            self._codes_for = rows
        return self._codes_lm

    def search_device(self, q: torch.Tensor, n_probe: int, top_k: int, shortlist: Optional[int] = None,
                      rerank: bool = True, filter_ids=None):
        """(dist, ids, candidates enumerated).  With ``rerank`` the distances are exact non-squared L2 over the
        ADC shortlist; without it they are the squared ADC estimates and ``top_k`` of them are returned.
        A filtered search (``filter_ids``) takes the exact list scan of :class:`IvfIndex`."""
        if filter_ids is not None:
            return IvfIndex.search_device(self, q, n_probe, top_k, filter_ids=filter_ids)
        if self.codebooks is None:
            raise ValueError("Product quantiser must be trained before searching (train_pq / set_codebooks)")
        ptr, rows = self._csr()
        vec = self._vectors
        codes = self._list_codes()
        cdist = ops.ivf_centroid_dist(q, self._centroids_dev)
        probes, prefix = ops.ivf_probe(cdist, n_probe, ptr)
        R = max(top_k, shortlist or top_k)
        lut = ops.pq_lut(q.to(torch.float32).contiguous(), self._cb_lut)
        d, cand = ops.ivf_adc_scan(codes, lut, ptr, rows, probes, prefix, R)
        scanned = prefix[:, -1].to(torch.int64)
        if rerank:
            dd, ii = ops.rerank_f32(q, vec.vecs, cand, min(top_k, R), "l2", ids=vec.ids, n_rows=vec.n)
            return dd, ii, scanned
        ids = torch.where(cand >= 0, vec.ids[cand.clamp(min=0).long()], cand.long())
        return d[:, :top_k], ids[:, :top_k], scanned

    def search_batch(self, queries: np.ndarray, n_probe: int, top_k: int, shortlist: Optional[int] = None,
                     rerank: bool = True, filter_ids=None) -> Tuple[np.ndarray, np.ndarray]:
        if self.centroids is None:
            raise ValueError("Index must be built before searching")
        queries = np.ascontiguousarray(queries, dtype=np.float32)
        if queries.ndim != 2 or queries.shape[1] != self.centroids.shape[1]:
            raise ValueError(f"queries must have shape (Q, {self.centroids.shape[1]})")
        if n_probe <= 0 or n_probe > self.k:
            raise ValueError(f"n_probe must be between 1 and {self.k}")
        with self._lock:
            q = torch.from_numpy(queries).to(self._device)
            d, i, scanned = self.search_device(q, n_probe, top_k, shortlist, rerank, filter_ids)
            self.last_scanned = scanned.cpu().numpy()
            return d.cpu().numpy(), i.cpu().numpy()
