"""CPU oracle for the HNSW graph path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates the reference's HNSW (src/vector_db/indexing/hnsw.py) over plain containers so the
CUDA traversal / insert kernels can be checked for identical result lists, identical graphs
and identical distance-evaluation counts.  PINNED: ``tests/golden/hnsw_*.npz`` hold graphs and
search results produced by the reference class itself (oracle/gen_golden.py) and
tests/test_oracle_golden.py checks this restatement reproduces them exactly.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
from __future__ import annotations

import heapq
import math
import random
from typing import Dict, List, Optional, Sequence, Set, Tuple

import numpy as np


def _dist(a: np.ndarray, b: np.ndarray):
    # src/vector_db/util/distance.py:3-4
    d = a - b
    return np.sqrt(d.dot(d))


class HnswOracle:
    """Graph = ``links[id][level] -> list of ids``; vectors = ``vec[id]`` (fp32 rows)."""

    def __init__(self, M: int, ef_construction: int, seed: int = 42):
        # hnsw.py:48-60 ; rng = random.Random(42) at services/indexing_service.py:57
        self.M = M
        self.M_max = M
        self.M_max0 = 2 * M
        self.ef_construction = ef_construction
        self.rng = random.Random(seed)
        self.level_mult = 1.0 / math.log(M)
        self.links: Dict[int, Dict[int, List[int]]] = {}
        self.vec: Dict[int, np.ndarray] = {}
        self.entry: Optional[int] = None
        self.max_level = -1
        self.n_evals = 0  # distance evaluations since last reset

    # -- helpers ---------------------------------------------------------------------
    def _d(self, q: np.ndarray, node: int):
        self.n_evals += 1
        return _dist(q, self.vec[node])

    def draw_level(self) -> int:
        # hnsw.py:66-68
        return int(-math.log(self.rng.random()) * self.level_mult)

    # -- a4: best-first beam on one level (hnsw.py:79-134) ----------------------------
    def search_layer(self, q: np.ndarray, entry: int, ef: int, level: int,
                     allowed: Optional[Set[int]] = None) -> List[Tuple[float, int]]:
        d0 = self._d(q, entry)
        frontier = [(d0, entry)]              # min-heap of nodes still to expand
        best: List[Tuple[float, int]] = []    # max-heap through negated distance
        if allowed is None or entry in allowed:
            best.append((-d0, entry))
        seen = {entry}
        while frontier:
            d, node = heapq.heappop(frontier)
            if len(best) >= ef and d > -best[0][0]:
                break
            for nb in self.links[node].get(level, []):
                if nb in seen:
                    continue
                seen.add(nb)
                if nb not in self.vec:        # tombstone: reference swallows the lookup error
                    continue
                dn = self._d(q, nb)
                heapq.heappush(frontier, (dn, nb))
                if allowed is None or nb in allowed:
                    if len(best) < ef or dn < -best[0][0]:
                        heapq.heappush(best, (-dn, nb))
                        if len(best) > ef:
                            heapq.heappop(best)
        return [(-nd, node) for nd, node in sorted(best, reverse=True)]

    # -- a5: strict-improvement hill climb (hnsw.py:289-328) --------------------------
    def greedy_level(self, q: np.ndarray, cur: int, level: int) -> int:
        cur_d = self._d(q, cur)
        while True:
            nxt = cur
            for nb in self.links[cur].get(level, []):
                if nb not in self.vec:
                    continue
                d = self._d(q, nb)
                if d < cur_d:
                    cur_d, nxt = d, nb
            if nxt == cur:
                return cur
            cur = nxt

    # -- a7: insert (hnsw.py:143-225) --------------------------------------------------
    def insert(self, node: int, x: np.ndarray) -> None:
        self.vec[node] = np.asarray(x, dtype=np.float32)
        if node in self.links:
            return
        self.links[node] = {}
        lvl = self.draw_level()
        top = self.max_level
        if self.entry is None:
            self.entry, self.max_level = node, lvl
            for l in range(lvl + 1):
                self.links[node][l] = []
            return
        cur = self.entry
        for l in range(top, lvl, -1):
            cur = self.greedy_level(x, cur, l)
        for l in range(min(lvl, top), -1, -1):
            found = self.search_layer(x, cur, self.ef_construction, l)
            cap = self.M_max0 if l == 0 else self.M_max
            found.sort()
            chosen = [nid for _, nid in found[:self.M]]       # always M, hnsw.py:188
            self.links[node][l] = chosen
            for nb in chosen:
                lst = self.links[nb].setdefault(l, [])
                lst.append(node)
                if len(lst) > cap:                            # hnsw.py:199-213
                    scored = []
                    for other in lst:
                        if other in self.vec:
                            self.n_evals += 1
                            scored.append((_dist(self.vec[nb], self.vec[other]), other))
                    scored.sort()
                    self.links[nb][l] = [o for _, o in scored[:cap]]
            cur = chosen[0]
        if lvl > self.max_level:
            self.max_level, self.entry = lvl, node

    # -- a6: search (hnsw.py:330-374) ---------------------------------------------------
    def search(self, q: np.ndarray, k: int, ef: int = 50,
               allowed: Optional[Set[int]] = None) -> List[Tuple[float, int]]:
        if self.entry is None:
            return []
        q = np.asarray(q, dtype=np.float32)
        cur = self.entry
        for l in range(self.max_level, 0, -1):
            cur = self.greedy_level(q, cur, l)
        found = self.search_layer(q, cur, ef, 0, allowed)
        found.sort()
        return found[:k]

    # -- array export (the device layout of vector_db_b200.indexing.hnsw) --------------
    def export_arrays(self):
        """Dense export: rows are ids in sorted order.  Returns dict with ``ids`` (i64 [N]),
        ``levels`` (i32 [N], top level of each node), ``adj0`` (i32 [N,2M], -1 padded, ROW
        indices), ``up_ptr`` (i32 [N+1]), ``adj_up`` (i32 [sum(levels),M]), ``entry_row``,
        ``max_level``."""
        ids = np.array(sorted(self.links.keys()), dtype=np.int64)
        row = {int(i): r for r, i in enumerate(ids)}
        n = len(ids)
        levels = np.array([max(self.links[int(i)].keys(), default=0) for i in ids], dtype=np.int32)
        adj0 = np.full((n, self.M_max0), -1, dtype=np.int32)
        up_ptr = np.zeros(n + 1, dtype=np.int32)
        up_ptr[1:] = np.cumsum(levels)
        adj_up = np.full((int(up_ptr[-1]), self.M_max), -1, dtype=np.int32)
        for r, i in enumerate(ids):
            nb = self.links[int(i)]
            l0 = [row[x] for x in nb.get(0, [])]
            adj0[r, :len(l0)] = l0
            for l in range(1, int(levels[r]) + 1):
                lu = [row[x] for x in nb.get(l, [])]
                adj_up[up_ptr[r] + l - 1, :len(lu)] = lu
        return dict(ids=ids, levels=levels, adj0=adj0, up_ptr=up_ptr, adj_up=adj_up,
                    entry_row=np.int32(row[self.entry]) if self.entry is not None else np.int32(-1),
                    max_level=np.int32(self.max_level))

    @classmethod
    def from_links(cls, M, ef_construction, links, vec, entry, max_level):
        o = cls(M, ef_construction)
        o.links = {int(k): {int(l): [int(t) for t in v] for l, v in d.items()} for k, d in links.items()}
        o.vec = vec
        o.entry = entry
        o.max_level = max_level
        return o


def search_arrays(X: np.ndarray, adj0: np.ndarray, up_ptr: np.ndarray, adj_up: np.ndarray,
                  entry_row: int, max_level: int, q: np.ndarray, k: int, ef: int,
                  allowed_rows: Optional[np.ndarray] = None):
    """The same search (hnsw.py:330-374) restated over the dense row-indexed export; used to
    run the oracle on graphs that never existed as Python dicts (bench / large tests).
    Returns (list[(dist,row)], n_distance_evals)."""
    evals = 0
    q = np.asarray(q, dtype=np.float32)

    def dist(r):
        nonlocal evals
        evals += 1
        d = X[r] - q
        return np.sqrt(d.dot(d))

    if entry_row < 0:
        return [], 0
    cur = int(entry_row)
    for l in range(int(max_level), 0, -1):
        cur_d = dist(cur)
        while True:
            nxt = cur
            base = up_ptr[cur] + l - 1
            for nb in adj_up[base]:
                if nb < 0:
                    break
                d = dist(int(nb))
                if d < cur_d:
                    cur_d, nxt = d, int(nb)
            if nxt == cur:
                break
            cur = nxt
    d0 = dist(cur)
    frontier = [(d0, cur)]
    best = []
    if allowed_rows is None or allowed_rows[cur]:
        best.append((-d0, cur))
    seen = {cur}
    while frontier:
        d, node = heapq.heappop(frontier)
        if len(best) >= ef and d > -best[0][0]:
            break
        for nb in adj0[node]:
            nb = int(nb)
            if nb < 0:
                break
            if nb in seen:
                continue
            seen.add(nb)
            dn = dist(nb)
            heapq.heappush(frontier, (dn, nb))
            if allowed_rows is None or allowed_rows[nb]:
                if len(best) < ef or dn < -best[0][0]:
                    heapq.heappush(best, (-dn, nb))
                    if len(best) > ef:
                        heapq.heappop(best)
    out = sorted((-nd, r) for nd, r in best)
    return out[:k], evals
