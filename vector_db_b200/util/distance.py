"""L2 distance on the GPU.  Mirrors src/vector_db/util/distance.py:3-4 (non-squared L2, fp32 in,
np.float32 out) and adds the batched form the hot path actually wants."""
import numpy as np
import torch

from .. import _lib, ops


def euclidean_distances(queries: np.ndarray, vectors: np.ndarray) -> np.ndarray:
    """All-pairs sqrt(sum((q-x)^2)) -> float32 [Q,N], computed by the exact fp32 scan kernel."""
    dev = _lib.require_cuda()
    q = torch.as_tensor(np.ascontiguousarray(queries, dtype=np.float32), device=dev)
    x = torch.as_tensor(np.ascontiguousarray(vectors, dtype=np.float32), device=dev)
    n = x.shape[0]
    d, i = ops.exact_topk_f32(q, x, n)
    out = torch.empty_like(d)
    out.scatter_(1, i, d)
    return out.cpu().numpy()


def euclidean_vector_distance(v1: np.ndarray, v2: np.ndarray):
    """Single pair; same contract as the reference function (returns np.float32)."""
    a = np.asarray(v1, dtype=np.float32).reshape(1, -1)
    b = np.asarray(v2, dtype=np.float32).reshape(1, -1)
    if a.shape != b.shape:
        raise ValueError(f"shape mismatch {a.shape} vs {b.shape}")
    return np.float32(euclidean_distances(a, b)[0, 0])
