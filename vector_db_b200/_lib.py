"""ctypes binding of libvdb_b200.so -- the thin layer between the Python shim and the C ABI
declared in include/vdb_b200.h.  There is no CPU fallback: if the library is missing, or a call
fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvdb_b200.so")

VDB_OK, VDB_ERR_INVALID, VDB_ERR_CUDA, VDB_ERR_WORKSPACE, VDB_ERR_UNSUPPORTED, VDB_ERR_OVERFLOW = range(6)
METRIC_L2, METRIC_COSINE = 0, 1

_p, _i, _i64, _sz, _f = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float

# name -> (restype, argtypes); mirrors include/vdb_b200.h one to one
SIGNATURES = {
    "vdb_last_error": (C.c_char_p, []),
    "vdb_abi_version": (_i, []),
    "vdb_launch_count": (C.c_uint64, []),
    "vdb_profile_enable": (None, [_i]),
    "vdb_profile_read": (_i, [_i, C.POINTER(C.c_int), C.POINTER(C.c_float)]),
    "vdb_stage_aos_h2d": (_i, [_p, _i64, _i, _p, _p, _p]),
    "vdb_stage_aos_pageable_h2d": (_i, [_p, _i64, _i, _p, _p, _p, _p, _sz, _i, _i, _p]),
    "vdb_prepare_rows": (_i, [_p, _i64, _i, _p, _p, _p]),
    "vdb_exact_topk_f32_workspace": (_sz, [_i, _i64, _i]),
    "vdb_exact_topk_f32": (_i, [_p, _i, _p, _p, _p, _i64, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "vdb_exact_topk_f32_static_max_bytes": (_sz, []),
    "vdb_exact_topk_f32_static": (_i, [_p, _i, _p, _p, _p, _i64, _i, _i, _i, _p, _p, _p]),
    "vdb_rerank_f32": (_i, [_p, _i, _p, _p, _i64, _i, _p, _i, _i, _i, _p, _p, _p]),
    "vdb_exact_topk_tc_supported": (_i, [_i, _i64, _i, _i, _i]),
    "vdb_tc_mirror_elems": (_sz, [_i64, _i, _i]),
    "vdb_tc_prepare_rows": (_i, [_p, _i64, _i, _i, _p, _p, _p, _p]),
    "vdb_exact_topk_tc_workspace": (_sz, [_i, _i64, _i, _i, _i]),
    "vdb_exact_topk_tc": (_i, [_p, _i, _p, _p, _p, _p, _i64, _i, _i, _i, _p, _p, _p, _p, _sz, _p]),
    "vdb_ivf_topk_tc_workspace": (_sz, [_i, _i64, _i64, _i, _i, _i]),
    "vdb_ivf_topk_tc": (_i, [_p, _i, _p, _p, _p, _p, _i64, _i, _i, _i, _p, _p, _p, _p, _i64, _p, _i, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vdb_ivf_probe_refine": (_i, [_p, _i, _i, _p, _i, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p]),
    "vdb_ivf_group_workspace": (_sz, [_i]),
    "vdb_ivf_group_pairs": (_i, [_p, _i, _i, _p, _i, _i64, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vdb_ivf_tc_finish": (_i, [_p, _p, _i, _i, _i, _p, _p, _i, _p, _p, _p, _p, _p, _p, _p]),
    "vdb_debug_set_tc_flags": (None, [_i]),
    "vdb_pq_encode": (_i, [_p, _i64, _i, _i, _i, _p, _p, _p]),
    "vdb_pq_encode_tc_supported": (_i, [_i64, _i, _i, _i]),
    "vdb_pq_encode_tc_workspace": (_sz, [_i]),
    "vdb_pq_encode_tc": (_i, [_p, _i64, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "vdb_debug_set_pq_tc_mode": (None, [_i]),
    "vdb_debug_set_adc_coarse": (None, [_i]),
    "vdb_debug_adc_coarse_tables": (_i, [_p, _i, _i, _p, _p, _p, _i, _p]),
    "vdb_debug_pq_tc_stats": (_i, [_p, _i, _p, _p]),
    "vdb_pq_kmeans_workspace": (_sz, [_i, _i, _i]),
    "vdb_pq_kmeans_step": (_i, [_p, _i64, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "vdb_pq_kmeans_step_coded": (_i, [_p, _i64, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "vdb_pq_lut": (_i, [_p, _i, _i, _i, _i, _p, _p, _p]),
    "vdb_pq_adc_topk_workspace": (_sz, [_i, _i64, _i, _i]),
    "vdb_pq_adc_topk": (_i, [_p, _i64, _i, _p, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "vdb_hnsw_search_workspace": (_sz, [_i, _i, _i]),
    "vdb_hnsw_search": (_i, [_p, _i, _p, _i64, _i, _p, _i, _p, _i, _p, C.c_int32, _i, _i, _i, _p, _p, _p, _i,
                             _p, _p, _p, _p, _p, _sz, _p]),
    "vdb_debug_set_hnsw_smem_hash_limit": (None, [_i]),
    "vdb_debug_set_hnsw_ctas_per_sm": (None, [_i]),
    "vdb_debug_set_hnsw_insert_threads": (None, [_i]),
    "vdb_hnsw_insert_workspace": (_sz, [_i, _i]),
    "vdb_hnsw_insert": (_i, [_p, _i64, _i, _p, _i, _p, _i, _p, _p, C.c_int32, C.c_int32, _i, _i, _p, _p,
                             _p, _sz, _p]),
    "vdb_hnsw_build_search_workspace": (_sz, [_i, _i, _i]),
    "vdb_hnsw_build_search": (_i, [_p, _i64, _i, _p, _i, _p, _i, _p, _p, _i, _i, C.c_int32, _i, C.c_int32, _i, _i,
                                   _p, _p, _p, _sz, _p]),
    "vdb_hnsw_reverse_merge": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _p]),
    "vdb_ivf_centroid_dist": (_i, [_p, _i, _i, _p, _i, _p, _p]),
    "vdb_ivf_probe": (_i, [_p, _i, _i, _i, _p, _p, _p, _p]),
    "vdb_ivf_scan_workspace": (_sz, [_i, _i]),
    "vdb_ivf_scan": (_i, [_p, _i, _p, _i, _p, _p, _p, _p, _i, _p, _p, _i, _p, _p, _p, _p, _sz, _p]),
    "vdb_ivf_centroid_mean": (_i, [_p, _i, _p, _p, _i, _p, _p]),
    "vdb_ivf_adc_workspace": (_sz, [_i, _i, _i]),
    "vdb_ivf_adc_scan": (_i, [_p, _i, _p, _i, _p, _p, _p, _p, _i, _i, _p, _p, _p, _sz, _p]),
    "vdb_filter_eq": (_i, [_p, _i64, C.c_int32, _p, _p, _i, _p]),
    "vdb_merge_topk": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p]),
    "vdb_exchange_bytes": (_sz, [_i, _i64]),
    "vdb_exchange_grid": (_i, [_i]),
    "vdb_exchange_merge_topk": (_i, [_p, _p, _i, _i, _i, _p, _i, _i, _i64, _i, C.c_uint32, _p, _p, _p, _p]),
}

_lib = None
_lock = threading.Lock()


class VdbError(RuntimeError):
    pass


def load():
    """Load (once) and return the ctypes handle; raises if the CUDA library was not built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise VdbError(
                        f"{LIB_PATH} not found: build it with `python -m vector_db_b200.build` "
                        "(there is no CPU fallback)")
                lib = C.CDLL(LIB_PATH)
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(lib, name)
                    fn.restype = res
                    fn.argtypes = args
                _lib = lib
    return _lib


def last_error() -> str:
    return load().vdb_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc == VDB_OK:
        return
    msg = last_error()
    if rc in (VDB_ERR_INVALID, VDB_ERR_UNSUPPORTED):
        raise ValueError(msg)
    raise VdbError(f"vdb error {rc}: {msg}")


def ptr(t) -> int | None:
    """Device/host pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise VdbError("vector_db_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


PROF_SLOTS = {"exact_f32": 0, "exact_tc": 1, "adc_scan": 2, "pq_encode": 3, "hnsw_search": 4, "rerank": 5,
              "finalize": 6, "hnsw_insert": 7, "ivf_scan": 8}


def profile_enable(on: bool) -> None:
    load().vdb_profile_enable(int(on))


def profile_read(slot: str):
    """(launches, total_ms) of a kernel slot since the last read; waits for its events."""
    n, ms = C.c_int(0), C.c_float(0.0)
    check(load().vdb_profile_read(PROF_SLOTS[slot], C.byref(n), C.byref(ms)))
    return n.value, ms.value


def launch_count() -> int:
    return int(load().vdb_launch_count())


class Workspace:
    """Grow-only scratch buffer per (device, stream user); the C ABI never allocates."""

    def __init__(self):
        self._buf = None

    def get(self, nbytes: int, device) -> torch.Tensor:
        nbytes = max(int(nbytes), 256)
        if self._buf is None or self._buf.numel() < nbytes or self._buf.device != device:
            self._buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return self._buf
