"""CPU oracle for the IVF path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates the reference's IvfIndex (src/vector_db/indexing/ivf.py) over plain containers so the CUDA
probe / list-scan kernels can be checked for identical probe sets, identical result ids and the same
candidate counts.  PINNED: ``tests/golden/ivf.npz`` holds centroids, inverted lists, nearest-cluster
assignments and search results produced by the reference class itself (oracle/gen_golden.py) and
tests/test_oracle_golden.py checks this restatement reproduces them exactly.  K-means VALUES are
parity-unpinned (scipy kmeans2 draws its initial points from the global numpy RNG; the reference tests
pin only shapes, tests/indexing/test_ivf.py:25-45); :func:`oracle.kmeans_lloyd` restates the Lloyd loop
for a given initialisation.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np


def centroid_distances(centroids: np.ndarray, x: np.ndarray) -> np.ndarray:
    # ivf.py:109 and :170
    return np.linalg.norm(centroids - x, axis=1)


def nearest_cluster(centroids: np.ndarray, x: np.ndarray) -> int:
    # ivf.py:109-110 (add): first minimum wins
    return int(np.argmin(centroid_distances(centroids, x)))


def probe_order(centroids: np.ndarray, query: np.ndarray, n_probe: int) -> np.ndarray:
    # ivf.py:170-171
    return np.argsort(centroid_distances(centroids, query))[:n_probe]


def _dist(a: np.ndarray, b: np.ndarray):
    # src/vector_db/util/distance.py:3-4
    d = a - b
    return np.sqrt(d.dot(d))


def search(centroids: np.ndarray, lists: Sequence[Sequence[int]], vec: Dict[int, np.ndarray], query: np.ndarray,
           n_probe: int, top_k: int) -> Tuple[List[Tuple[float, int]], int, np.ndarray]:
    """ivf.py:169-199 without the storage.get of the result nodes.  Returns
    ([(dist, id)] first top_k of the stable sort by dist, candidates enumerated, probed clusters)."""
    probes = probe_order(centroids, query, n_probe)
    cand: List[Tuple[float, int]] = []
    for c in probes:
        for node_id in lists[int(c)]:
            if node_id not in vec:          # deleted from storage but still listed (ivf.py:188-190)
                continue
            cand.append((_dist(query, vec[node_id]), node_id))
    n_cand = len(cand)
    cand.sort(key=lambda t: t[0])           # stable: enumeration order breaks ties (ivf.py:193)
    return cand[:top_k], n_cand, probes


def lists_from_labels(labels: np.ndarray, ids: Sequence[int], k: int) -> List[List[int]]:
    # ivf.py:78-81
    lists: List[List[int]] = [[] for _ in range(k)]
    for i, lab in enumerate(labels):
        lists[int(lab)].append(ids[i])
    return lists


def ivf_adc_search(centroids: np.ndarray, lists: Sequence[Sequence[int]], codes: Dict[int, np.ndarray], lut: np.ndarray,
                   query: np.ndarray, n_probe: int, R: int):
    """PARITY UNPINNED (no reference implementation: the reference has IVF and PQ but not IVF-PQ).  ADC in the
    fixed-point arithmetic of :func:`oracle.adc_scan_topk` over the members of the probed lists (ivf.py:170-191
    enumeration order); the R best by (d_int, enumeration order).  ``lut`` is the query's [m,256] table.
    Returns ([(float32 d, id)], candidates enumerated)."""
    from oracle.oracle import adc_quantize_lut
    lutq, e = adc_quantize_lut(lut[None])
    lutq, e = lutq[0], int(e[0])
    m = lut.shape[0]
    cand = []
    for c in probe_order(centroids, query, n_probe):
        for node_id in lists[int(c)]:
            if node_id not in codes:
                continue
            code = codes[node_id]
            cand.append((int(sum(int(lutq[j, code[j]]) for j in range(m))), len(cand), node_id))
    n_cand = len(cand)
    cand.sort(key=lambda t: (t[0], t[1]))
    return [(np.float32(np.ldexp(np.float32(t[0]), -e)), t[2]) for t in cand[:R]], n_cand
