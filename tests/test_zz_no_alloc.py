"""GPU: the allocation-free exact scan (vdb_exact_topk_f32_static, scratch in a static __device__ array of the library)
against the existing entry point (vdb_exact_topk_f32, caller workspace), called straight through the C ABI.

Collected last on purpose.  Output buffers are allocated BEFORE the calls; between the
calls torch's allocator statistics must not move, i.e. the entry point itself allocates nothing."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _call_static(lib, q, db, ids, mask, k, metric, out_d, out_i):
    from vector_db_b200 import _lib
    return lib.vdb_exact_topk_f32_static(q.data_ptr(), q.shape[0], db.data_ptr(), ids.data_ptr() if ids is not None else None,
                                         mask.data_ptr() if mask is not None else None, db.shape[0], db.shape[1], k, metric,
                                         out_d.data_ptr(), out_i.data_ptr(), _lib.stream_ptr(torch.device(DEV)))


@pytest.mark.parametrize("metric", [0, 1])
def test_static_scratch_scan_equals_workspace_scan_twice_in_a_row(metric):
    from vector_db_b200 import _lib, ops
    lib = _lib.load()
    assert lib.vdb_exact_topk_f32_static_max_bytes() == 8 << 20
    g = torch.Generator(device=DEV).manual_seed(77 + metric)
    k = 10
    ids = torch.arange(60_000, device=DEV, dtype=torch.int64) * 3 + 1
    # two calls in a row with DIFFERENT shapes, the larger first: the second runs over the first one's leftovers
    for n, d, nq in ((60_000, 96, 200), (9_000, 64, 37)):
        X = torch.randn((n, d), device=DEV, generator=g)
        Q = torch.randn((nq, d), device=DEV, generator=g)
        Q[0] = X[11]
        want_d, want_i = ops.exact_topk_f32(Q, X, k, metric, ids=ids[:n])
        out_d = torch.empty((nq, k), dtype=torch.float32, device=DEV)
        out_i = torch.empty((nq, k), dtype=torch.int64, device=DEV)
        torch.cuda.synchronize()
        before = (torch.cuda.memory_stats(DEV)["allocation.all.allocated"], torch.cuda.memory_allocated(DEV))
        _lib.check(_call_static(lib, Q, X, ids[:n], None, k, metric, out_d, out_i))
        torch.cuda.synchronize()
        after = (torch.cuda.memory_stats(DEV)["allocation.all.allocated"], torch.cuda.memory_allocated(DEV))
        assert before == after, "the static entry point went through torch's allocator"
        assert torch.equal(out_i, want_i), f"{(out_i != want_i).sum().item()} ids differ"
        assert torch.equal(out_d, want_d)                  # same kernels, same arithmetic: bit-equal
        if metric == 0:
            assert out_d[0, 0].item() == 0.0 and out_i[0, 0].item() == 11 * 3 + 1
    # masked variant, same buffers, third call in a row
    words = torch.full(((n + 31) // 32,), 0x55555555, dtype=torch.int32, device=DEV)       # even rows only
    want_d, want_i = ops.exact_topk_f32(Q, X, k, metric, ids=ids[:n], row_mask=words)
    _lib.check(_call_static(lib, Q, X, ids[:n], words, k, metric, out_d, out_i))
    assert torch.equal(out_i, want_i) and torch.equal(out_d, want_d)
    assert ((out_i - 1) // 3 % 2 == 0).all()


def test_static_scratch_scan_rejects_what_does_not_fit():
    from vector_db_b200 import _lib
    lib = _lib.load()
    X = torch.randn((4096, 32), device=DEV)
    nq, k = 3000, 512                                       # 3000 x 512 x 8 B x >= 1 slice = 12 MB > 8 MiB
    Q = torch.randn((nq, 32), device=DEV)
    out_d = torch.empty((nq, k), dtype=torch.float32, device=DEV)
    out_i = torch.empty((nq, k), dtype=torch.int64, device=DEV)
    rc = _call_static(lib, Q, X, None, None, k, 0, out_d, out_i)
    assert rc == _lib.VDB_ERR_UNSUPPORTED
    msg = _lib.last_error()
    assert "bytes of scratch" in msg and "vdb_exact_topk_f32 " in msg
    with pytest.raises(ValueError, match="built-in"):
        _lib.check(rc)
    assert isinstance(ctypes.c_size_t(lib.vdb_exact_topk_f32_static_max_bytes()).value, int)
