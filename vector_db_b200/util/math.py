"""Largest-k selector with the contract of src/vector_db/util/math.py:3-5 (indices of the k largest
scores, descending), run through the GPU merge/select kernel."""
import numpy as np
import torch

from .. import _lib, ops


def top_k_indices_sorted(arr: np.ndarray, k: int) -> np.ndarray:
    dev = _lib.require_cuda()
    a = np.ascontiguousarray(arr, dtype=np.float32).reshape(-1)
    n = a.shape[0]
    if k <= 0 or k > n:
        raise ValueError("k out of range")
    d = torch.as_tensor(a, device=dev)
    # chunk into lists of <= 4096 so the single-CTA merge kernel's envelope is respected
    width = min(n, 4096)
    if k > width:
        raise ValueError("top_k_indices_sorted: k > 4096 is outside the select kernel's envelope")
    lists = (n + width - 1) // width
    pad = lists * width - n
    idx = torch.arange(n, dtype=torch.int64, device=dev)
    if pad:
        d = torch.cat([d, torch.full((pad,), float("-inf"), device=dev)])
        idx = torch.cat([idx, torch.full((pad,), -1, dtype=torch.int64, device=dev)])
    # per-list pre-sort is not required by the kernel (it sorts all L*k keys), only the shape is
    while True:
        L = d.numel() // width
        if L * width <= 16384:
            od, oi = ops.merge_topk(d.view(L, 1, width), idx.view(L, 1, width), descending=True)
            return oi[0, :k].cpu().numpy()
        # reduce 4 lists at a time
        group = 16384 // width
        keep_d, keep_i = [], []
        for s in range(0, L, group):
            dd = d.view(L, 1, width)[s:s + group]
            ii = idx.view(L, 1, width)[s:s + group]
            od, oi = ops.merge_topk(dd, ii, descending=True)
            keep_d.append(od.view(-1))
            keep_i.append(oi.view(-1))
        d, idx = torch.cat(keep_d), torch.cat(keep_i)
