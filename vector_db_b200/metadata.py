"""Columnar metadata in HBM and the equality pre-filter over it.

The reference keeps metadata as a JSON string per node and answers a filter with a Python scan that decodes
every node: ``node.metadata.get(k) != v`` for every key of the filter
(src/vector_db/services/storage_service.py:106-128); the resulting id set reaches the search as ``filter_ids``
(src/vector_db/api/app.py:184-198, src/vector_db/indexing/hnsw.py:96,127).  Here every metadata key is a column
of dictionary codes next to the index rows and the filter is evaluated on the GPU straight into the row bitmap
the search kernels consume (``vdb_filter_eq``, csrc/filter.cu) -- same matches, no id set on the host.

Equality is Python's, as in the reference: the dictionary is a Python dict, so ``1 == 1.0 == True`` share a code;
a missing key and a stored ``None`` are the same code 0 (``dict.get`` cannot tell them apart either); NaN
matches nothing; lists / dicts (JSON containers) are compared by value.
"""
from __future__ import annotations

import math
from typing import Any, Dict, Iterable, Optional, Sequence

import numpy as np
import torch

from . import _lib

_NO_MATCH = -1      # never stored in a column


def _canon(v: Any):
    """Hashable stand-in with Python's == semantics for JSON values."""
    if isinstance(v, dict):
        return ("__dict__", frozenset((k, _canon(x)) for k, x in v.items()))
    if isinstance(v, list):
        return ("__list__", tuple(_canon(x) for x in v))
    if isinstance(v, tuple):
        return ("__tuple__", tuple(_canon(x) for x in v))
    return v


def _is_nan(v: Any) -> bool:
    return isinstance(v, float) and math.isnan(v)


class DeviceMetadata:
    """One int32 code column per metadata key, indexed by index ROW (not node id)."""

    def __init__(self, device=None):
        self.device = _lib.require_cuda() if device is None else torch.device(device)
        self.rows = 0                                   # rows that have been given metadata (prefix of the index rows)
        self._dict: Dict[str, Dict[Any, int]] = {}      # key -> {canonical value -> code >= 1}
        self._host: Dict[str, np.ndarray] = {}          # key -> int32 codes, host mirror (grown geometrically)
        self._dev: Dict[str, torch.Tensor] = {}         # key -> uploaded column (valid while not dirty)
        self._dirty: Dict[str, bool] = {}

    # -- population -------------------------------------------------------------------------------
    def _column(self, key: str, need: int) -> np.ndarray:
        col = self._host.get(key)
        if col is None or col.shape[0] < need:
            cap = max(need, int((col.shape[0] if col is not None else 0) * 1.5) + 1024)
            grown = np.zeros(cap, dtype=np.int32)
            if col is not None:
                grown[:col.shape[0]] = col
            self._host[key] = col = grown
            self._dict.setdefault(key, {})
        return col

    def _code(self, key: str, value: Any, create: bool) -> int:
        if value is None:
            return 0
        if _is_nan(value):
            if not create:
                return _NO_MATCH
            d = self._dict.setdefault(key, {})
            code = len(d) + 1
            d[("__nan__", code)] = code                 # a fresh code per stored NaN: equal to nothing
            return code
        d = self._dict.setdefault(key, {})
        c = _canon(value)
        try:
            code = d.get(c)
        except TypeError:                               # unhashable leaf (a set, ...): compare by repr
            c = ("__repr__", repr(value))
            code = d.get(c)
        if code is None:
            if not create:
                return _NO_MATCH
            code = len(d) + 1
            d[c] = code
        return code

    def set_row(self, row: int, metadata: Optional[Dict[str, Any]]) -> None:
        """(Re)place the metadata of one row; keys the row no longer has fall back to code 0."""
        row = int(row)
        for key, col in self._host.items():
            if row < col.shape[0] and col[row] != 0 and (not metadata or key not in metadata):
                col[row] = 0
                self._dirty[key] = True
        for key, value in (metadata or {}).items():
            col = self._column(key, row + 1)
            col[row] = self._code(key, value, create=True)
            self._dirty[key] = True
        self.rows = max(self.rows, row + 1)

    def set_rows(self, rows: Iterable[int], metadatas: Sequence[Optional[Dict[str, Any]]]) -> None:
        for r, md in zip(rows, metadatas):
            self.set_row(r, md)

    # -- filter ------------------------------------------------------------------------------------
    def _device_column(self, key: str, n: int) -> torch.Tensor:
        col = self._column(key, n)
        dev = self._dev.get(key)
        if dev is None or dev.shape[0] < n or self._dirty.get(key, True):
            src = torch.from_numpy(col)
            self._dev[key] = dev = (src.pin_memory() if src.numel() > (1 << 16) else src).to(self.device, non_blocking=True)
            self._dirty[key] = False
        return dev

    def filter_bitmap(self, filter_dict: Dict[str, Any], n_rows: int, alive: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Row bitmap (uint32 words in an int32 tensor [ceil(n_rows/32)]): bit r set iff row r is alive and
        ``metadata.get(k) == v`` for every (k, v) of the filter -- storage_service.py:118-127 over the rows."""
        lib = _lib.load()
        n = int(n_rows)
        words = (max(n, 1) + 31) // 32
        out = torch.zeros(words, dtype=torch.int32, device=self.device)
        if n == 0:
            return out
        stream = torch.cuda.current_stream(self.device).cuda_stream
        alive_p = alive.data_ptr() if alive is not None else None
        first = 1
        for key, value in filter_dict.items():
            want = self._code(key, value, create=False)
            if want == _NO_MATCH:                       # a value no row holds (or NaN): nothing matches
                return out.zero_()
            col = self._device_column(key, n)
            _lib.check(lib.vdb_filter_eq(col.data_ptr(), n, want, alive_p if first else None, out.data_ptr(), first, stream))
            first = 0
        if first:                                       # empty filter: every live row matches
            if alive is not None:
                out.copy_(alive[:words])
            else:
                ones = np.full(words, 0xFFFFFFFF, dtype=np.uint32)
                if n % 32:
                    ones[-1] = np.uint32((1 << (n % 32)) - 1)
                out.copy_(torch.from_numpy(ones.view(np.int32)))
        return out
