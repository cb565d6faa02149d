"""Typed wrappers over the C ABI: torch CUDA tensors in, torch CUDA tensors out.

Every function enqueues on torch's current stream and returns device tensors without
synchronising (unless stated).  All compute happens in libvdb_b200.so; torch is used for
device memory and streams only.
"""
from __future__ import annotations

import threading
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import METRIC_COSINE, METRIC_L2, check, load, ptr

_tls = threading.local()


def _ws(name: str, nbytes: int, device) -> torch.Tensor:
    """Per-thread, per-purpose grow-only workspace (search calls may come from many threads)."""
    pool = getattr(_tls, "pool", None)
    if pool is None:
        pool = _tls.pool = {}
    key = (name, device.index)
    w = pool.get(key)
    if w is None:
        w = pool[key] = _lib.Workspace()
    return w.get(nbytes, device)


def _metric(metric) -> int:
    if metric in (METRIC_L2, "l2", "L2", "euclidean"):
        return METRIC_L2
    if metric in (METRIC_COSINE, "cos", "cosine"):
        return METRIC_COSINE
    raise ValueError(f"unknown metric {metric!r}")


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def prepare_rows(vecs: torch.Tensor, want_norm=True, want_bf16=True):
    vecs = _f32(vecs, "vecs")
    n, d = vecs.shape
    sq = torch.empty(n, dtype=torch.float32, device=vecs.device) if want_norm else None
    mir = torch.empty((n, d), dtype=torch.bfloat16, device=vecs.device) if want_bf16 else None
    check(load().vdb_prepare_rows(ptr(vecs), n, d, ptr(sq), ptr(mir), _lib.stream_ptr(vecs.device)))
    return sq, mir


def exact_topk_f32(q: torch.Tensor, db: torch.Tensor, k: int, metric=METRIC_L2, ids: Optional[torch.Tensor] = None,
                   row_mask: Optional[torch.Tensor] = None, n_rows: Optional[int] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Exact fp32 scan.  Returns (values f32 [Q,k], ids i64 [Q,k])."""
    q = _f32(q, "q")
    db = _f32(db, "db") if db.numel() else db
    Q, D = q.shape
    N = db.shape[0] if n_rows is None else int(n_rows)
    dev = q.device
    out_d = torch.empty((Q, k), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=dev)
    lib = load()
    nbytes = lib.vdb_exact_topk_f32_workspace(Q, N, k)
    ws = _ws("exact", nbytes, dev)
    check(lib.vdb_exact_topk_f32(ptr(q), Q, ptr(db), ptr(ids), ptr(row_mask), N, D, k, _metric(metric), ptr(out_d),
                                 ptr(out_i), ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return out_d, out_i


def rerank_f32(q: torch.Tensor, db: torch.Tensor, cand: torch.Tensor, k: int, metric=METRIC_L2,
               ids: Optional[torch.Tensor] = None, n_rows: Optional[int] = None):
    q = _f32(q, "q")
    Q, D = q.shape
    cand = cand.to(torch.int32).contiguous()
    R = cand.shape[1]
    N = db.shape[0] if n_rows is None else int(n_rows)
    out_d = torch.empty((Q, k), dtype=torch.float32, device=q.device)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=q.device)
    check(load().vdb_rerank_f32(ptr(q), Q, ptr(db), ptr(ids), N, D, ptr(cand), R, k, _metric(metric), ptr(out_d),
                                ptr(out_i), _lib.stream_ptr(q.device)))
    return out_d, out_i


def pq_encode(x: torch.Tensor, codebooks: torch.Tensor, engine: str = "auto") -> torch.Tensor:
    """x [N,D] fp32, codebooks [m,K,dsub] fp32 -> codes uint8 [N,m] (bit-exact with the reference).
    engine: "auto" = tensor-core kernel when the shape allows it (dsub == 8, K <= 256) and the batch
    is large enough to fill the machine, "tc" / "f32" force one of the two kernels."""
    x = _f32(x, "x")
    codebooks = _f32(codebooks, "codebooks")
    n, d = x.shape
    m, K, dsub = codebooks.shape
    if m * dsub != d:
        raise ValueError(f"codebooks {tuple(codebooks.shape)} do not tile dim {d}")
    codes = torch.empty((n, m), dtype=torch.uint8, device=x.device)
    lib = load()
    tc_ok = n > 0 and bool(lib.vdb_pq_encode_tc_supported(n, d, m, K))
    if engine == "tc" and not tc_ok:
        raise ValueError("tensor-core encode needs dsub == 8 and K <= 256")
    if tc_ok and (engine == "tc" or (engine == "auto" and n >= 4096)):
        ws = _ws("pq_encode_tc", lib.vdb_pq_encode_tc_workspace(m), x.device)
        check(lib.vdb_pq_encode_tc(ptr(x), n, d, m, K, ptr(codebooks), ptr(codes), ptr(ws), ws.numel(),
                                   _lib.stream_ptr(x.device)))
    else:
        check(lib.vdb_pq_encode(ptr(x), n, d, m, K, ptr(codebooks), ptr(codes), _lib.stream_ptr(x.device)))
    return codes


def pq_kmeans_step(x: torch.Tensor, codebooks: torch.Tensor, labels: Optional[torch.Tensor] = None) -> None:
    """One Lloyd round for all sub-spaces, in place on `codebooks` [m,K,dsub].  When the tensor-core encoder
    applies (dsub == 8, large batch) and no labels are requested, the assignment is one encode call and the
    round is the mean update over its codes."""
    n, d = x.shape
    m, K, dsub = codebooks.shape
    lib = load()
    nbytes = lib.vdb_pq_kmeans_workspace(m, K, dsub)
    ws = _ws("kmeans", nbytes, x.device)
    if labels is None and n >= 4096 and lib.vdb_pq_encode_tc_supported(n, d, m, K):
        codes = pq_encode(x, codebooks, engine="tc")
        check(lib.vdb_pq_kmeans_step_coded(ptr(x), n, d, m, K, ptr(codes), ptr(codebooks), ptr(ws), ws.numel(),
                                           _lib.stream_ptr(x.device)))
        return
    check(lib.vdb_pq_kmeans_step(ptr(x), n, d, m, K, ptr(codebooks), ptr(labels), ptr(ws), ws.numel(),
                                 _lib.stream_ptr(x.device)))


def pq_lut(q: torch.Tensor, codebooks: torch.Tensor) -> torch.Tensor:
    """LUT [Q,m,K] fp32 of squared sub-space distances."""
    q = _f32(q, "q")
    codebooks = _f32(codebooks, "codebooks")
    Q, d = q.shape
    m, K, dsub = codebooks.shape
    lut = torch.empty((Q, m, K), dtype=torch.float32, device=q.device)
    check(load().vdb_pq_lut(ptr(q), Q, d, m, K, ptr(codebooks), ptr(lut), _lib.stream_ptr(q.device)))
    return lut


def pq_adc_topk(codes: torch.Tensor, lut: torch.Tensor, R: int, ids: Optional[torch.Tensor] = None,
                n_rows: Optional[int] = None, row_mask: Optional[torch.Tensor] = None):
    """codes uint8 [N,m], lut fp32 [Q,m,256] -> (d f32 [Q,R], idx i64 [Q,R]) by (d,row) ascending.
    row_mask: optional int32 bitmap over rows (bit cleared = row skipped)."""
    if lut.shape[2] != 256:
        raise ValueError("ADC scan needs 256-entry LUT rows (nbits=8); pad the codebooks")
    Q, m = lut.shape[0], lut.shape[1]
    N = codes.shape[0] if n_rows is None else int(n_rows)
    dev = lut.device
    out_d = torch.empty((Q, R), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, R), dtype=torch.int64, device=dev)
    lib = load()
    nbytes = lib.vdb_pq_adc_topk_workspace(Q, N, m, R)
    if nbytes == 0 and Q > 0:
        raise ValueError(f"ADC scan: m={m} R={R} do not fit shared memory")
    ws = _ws("adc", nbytes, dev)
    check(lib.vdb_pq_adc_topk(ptr(codes), N, m, ptr(lut), Q, R, ptr(ids), ptr(row_mask), ptr(out_d), ptr(out_i), ptr(ws),
                              ws.numel(), _lib.stream_ptr(dev)))
    return out_d, out_i


def ivf_centroid_dist(q: torch.Tensor, centroids: torch.Tensor) -> torch.Tensor:
    """np.linalg.norm(centroids - q_i, axis=1) for every query, numpy operation order -> f32 [Q,K]."""
    q = _f32(q, "q")
    centroids = _f32(centroids, "centroids")
    Q, D = q.shape
    K = centroids.shape[0]
    out = torch.empty((Q, K), dtype=torch.float32, device=q.device)
    lib = load()
    for s in range(0, Q, 32768):
        e = min(Q, s + 32768)
        check(lib.vdb_ivf_centroid_dist(ptr(q[s:e]), e - s, D, ptr(centroids), K, ptr(out[s:e]), _lib.stream_ptr(q.device)))
    return out


def ivf_probe(cdist: torch.Tensor, n_probe: int, list_ptr: Optional[torch.Tensor] = None):
    """n_probe nearest centroids per query by (distance, centroid index) -> probes i32 [Q,n_probe] and, when
    list_ptr is given, the candidate prefix sums i32 [Q,n_probe+1]."""
    Q, K = cdist.shape
    probes = torch.empty((Q, n_probe), dtype=torch.int32, device=cdist.device)
    prefix = torch.empty((Q, n_probe + 1), dtype=torch.int32, device=cdist.device) if list_ptr is not None else None
    check(load().vdb_ivf_probe(ptr(cdist), Q, K, n_probe, ptr(list_ptr), ptr(probes), ptr(prefix),
                               _lib.stream_ptr(cdist.device)))
    return probes, prefix


def ivf_probe_tc(q: torch.Tensor, centroids: torch.Tensor, cent_mirror: torch.Tensor, cent_stats: torch.Tensor, n_probe: int,
                 list_ptr: Optional[torch.Tensor] = None):
    """The same (probes, prefix) as ivf_probe(ivf_centroid_dist(q, centroids), n_probe, list_ptr) for a batch: the n_probe + 4
    nearest centroids per query from the tensor-core scan, re-scored in numpy's order (vdb_ivf_probe_refine).  Also returns the
    per-query fallback flags and the 0-dim device count of queries the scan's certificate sent through its fp32 rescan."""
    q = _f32(q, "q")
    Q, D = q.shape
    K = centroids.shape[0]
    kc = min(K, n_probe + 4)
    cd, ci, n_unc = exact_topk_tc(q, centroids, cent_mirror, cent_stats, kc, "l2", return_uncertified=True)
    dev = q.device
    probes = torch.empty((Q, n_probe), dtype=torch.int32, device=dev)
    prefix = torch.empty((Q, n_probe + 1), dtype=torch.int32, device=dev) if list_ptr is not None else None
    scratch = _ws("ivf_cdist", Q * K * 4, dev)
    flags = torch.empty((Q,), dtype=torch.int32, device=dev)
    check(load().vdb_ivf_probe_refine(ptr(q), Q, D, ptr(_f32(centroids, "centroids")), K, ptr(cd), ptr(ci), kc, n_probe, ptr(list_ptr),
                                      ptr(probes), ptr(prefix), ptr(scratch), ptr(flags), _lib.stream_ptr(dev)))
    return probes, prefix, flags, n_unc


def ivf_scan(q: torch.Tensor, db: torch.Tensor, list_ptr: torch.Tensor, list_rows: torch.Tensor, probes: torch.Tensor,
             prefix: torch.Tensor, k: int, alive: Optional[torch.Tensor] = None, ids: Optional[torch.Tensor] = None,
             return_scanned: bool = False):
    """Distances to every member of the probed lists -> (d f32 [Q,k], id i64 [Q,k]) in the reference's stable
    candidate order (ivf.py:177-193)."""
    q = _f32(q, "q")
    Q, D = q.shape
    n_probe = probes.shape[1]
    dev = q.device
    out_d = torch.empty((Q, k), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=dev)
    scanned = torch.empty((Q,), dtype=torch.int64, device=dev) if return_scanned else None
    lib = load()
    for s in range(0, Q, 32768):
        e = min(Q, s + 32768)
        ws = _ws("ivf", lib.vdb_ivf_scan_workspace(e - s, k), dev)
        check(lib.vdb_ivf_scan(ptr(q[s:e]), e - s, ptr(db), D, ptr(list_ptr), ptr(list_rows), ptr(probes[s:e]),
                               ptr(prefix[s:e]), n_probe, ptr(alive), ptr(ids), k, ptr(out_d[s:e]), ptr(out_i[s:e]),
                               ptr(scanned[s:e]) if scanned is not None else None, ptr(ws), ws.numel(),
                               _lib.stream_ptr(dev)))
    return (out_d, out_i, scanned) if return_scanned else (out_d, out_i)


def ivf_centroid_mean(x: torch.Tensor, list_ptr: torch.Tensor, list_rows: torch.Tensor, centroids: torch.Tensor) -> None:
    """Lloyd mean update in place: centroid c = mean of rows list_rows[list_ptr[c]:list_ptr[c+1]] (empty: unchanged)."""
    K, D = centroids.shape
    check(load().vdb_ivf_centroid_mean(ptr(x), D, ptr(list_ptr), ptr(list_rows), K, ptr(centroids),
                                       _lib.stream_ptr(x.device)))


def ivf_adc_scan(codes_lm: torch.Tensor, lut: torch.Tensor, list_ptr: torch.Tensor, list_rows: torch.Tensor,
                 probes: torch.Tensor, prefix: torch.Tensor, R: int):
    """IVF-PQ: ADC over the probed lists (codes list-major) -> (d f32 [Q,R] squared estimates, rows i32 [Q,R])."""
    if lut.shape[2] != 256:
        raise ValueError("ADC scan needs 256-entry LUT rows (nbits=8); pad the codebooks")
    Q, m = lut.shape[0], lut.shape[1]
    n_probe = probes.shape[1]
    dev = lut.device
    out_d = torch.empty((Q, R), dtype=torch.float32, device=dev)
    out_r = torch.empty((Q, R), dtype=torch.int32, device=dev)
    lib = load()
    for s in range(0, Q, 32768):
        e = min(Q, s + 32768)
        ws = _ws("ivf_adc", lib.vdb_ivf_adc_workspace(e - s, m, R), dev)
        check(lib.vdb_ivf_adc_scan(ptr(codes_lm), m, ptr(lut[s:e]), e - s, ptr(list_ptr), ptr(list_rows), ptr(probes[s:e]),
                                   ptr(prefix[s:e]), n_probe, R, ptr(out_d[s:e]), ptr(out_r[s:e]), ptr(ws), ws.numel(),
                                   _lib.stream_ptr(dev)))
    return out_d, out_r


def merge_topk(d: torch.Tensor, i: torch.Tensor, descending: bool = False):
    """d,i [L,Q,k] (per-shard sorted lists) -> merged (d [Q,k], i [Q,k])."""
    L, Q, k = d.shape
    d = d.contiguous()
    i = i.contiguous()
    out_d = torch.empty((Q, k), dtype=torch.float32, device=d.device)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=d.device)
    check(load().vdb_merge_topk(ptr(d), ptr(i), L, Q, k, int(descending), ptr(out_d), ptr(out_i),
                                _lib.stream_ptr(d.device)))
    return out_d, out_i


def exchange_bytes(world: int, cap: int) -> int:
    return int(load().vdb_exchange_bytes(int(world), int(cap)))


def exchange_grid(Q: int) -> int:
    return int(load().vdb_exchange_grid(int(Q)))


def exchange_merge_topk(d: torch.Tensor, i: torch.Tensor, peer_bufs: torch.Tensor, rank: int, world: int, cap: int,
                        parity: int, target: int, status: torch.Tensor, descending: bool = False):
    """Fused shard exchange + merge over peer memory (csrc/exchange.cu); d f32 [Q,k], i i64 [Q,k] local lists;
    peer_bufs i64 [world] device array of exchange-buffer addresses.  Returns merged (d [Q,k], i [Q,k])."""
    d = _f32(d, "d")
    if i.dtype != torch.int64 or not i.is_contiguous():
        i = i.to(torch.int64).contiguous()
    Q, k = d.shape
    out_d = torch.empty((Q, k), dtype=torch.float32, device=d.device)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=d.device)
    check(load().vdb_exchange_merge_topk(ptr(d), ptr(i), Q, k, int(descending), ptr(peer_bufs), int(rank), int(world),
                                         int(cap), int(parity), int(target) & 0xFFFFFFFF, ptr(out_d), ptr(out_i),
                                         ptr(status), _lib.stream_ptr(d.device)))
    return out_d, out_i


def _popcount_words(words: torch.Tensor) -> int:
    """Set bits of an int32 bitmap (one small reduction + a synchronising read)."""
    b = words.view(torch.uint8).to(torch.int32)
    b = (b & 0x55) + ((b >> 1) & 0x55)
    b = (b & 0x33) + ((b >> 2) & 0x33)
    b = (b & 0x0F) + (b >> 4)
    return int(b.sum().item())


def hnsw_search(q: torch.Tensor, db: torch.Tensor, n_rows: int, adj0: torch.Tensor, adj_up: torch.Tensor,
                up_ptr: torch.Tensor, entry_row: int, max_level: int, ef: int, k: int,
                alive: Optional[torch.Tensor] = None, flt: Optional[torch.Tensor] = None, want_stats: bool = False,
                levels: Optional[torch.Tensor] = None):
    """Batched graph search.  Returns (dist f32 [Q,k], rows i32 [Q,k], stats u32 [Q,2] | None).
    Synchronises once to read the status word; a full visited table is retried with a larger table.

    Filtered search (``flt``: row bitmap, bit set = allowed).  The reference keeps EVERY evaluated node in its
    candidate heap and walks on until ``ef`` matches are found (hnsw.py:102-131), i.e. through the whole
    component when fewer than ``ef`` rows match.  The kernel's frontier is bounded, so:
      * fewer than ``ef`` allowed rows -> the beam can never fill; the answer is the exact masked scan over the
        allowed rows (what the exhaustive walk returns on a connected graph, a superset otherwise), run directly;
      * a query whose frontier had to drop candidates, or whose visited table filled up, is finished the same way.
    Envelope: every other filtered query is the reference's traversal, step for step."""
    q = _f32(q, "q")
    Q, D = q.shape
    dev = q.device
    out_d = torch.empty((Q, k), dtype=torch.float32, device=dev)
    out_r = torch.empty((Q, k), dtype=torch.int32, device=dev)
    stats = torch.zeros((Q, 2), dtype=torch.int32, device=dev) if want_stats else None
    lib = load()
    deg0 = adj0.shape[1]
    deg_up = adj_up.shape[1] if adj_up is not None and adj_up.dim() == 2 else deg0 // 2

    def masked_scan(sel: Optional[torch.Tensor]):
        words = (max(n_rows, 1) + 31) // 32
        mask = flt[:words]
        if alive is not None:
            mask = torch.bitwise_and(mask, alive[:words])
        qq = q if sel is None else q[sel]
        d, r = exact_topk_f32(qq, db, k, METRIC_L2, row_mask=mask.contiguous(), n_rows=n_rows)
        if sel is None:
            out_d.copy_(d)
            out_r.copy_(r.to(torch.int32))
        else:
            out_d[sel] = d
            out_r[sel] = r.to(torch.int32)

    if flt is not None and Q > 0 and _popcount_words(flt[:(max(n_rows, 1) + 31) // 32]) < ef:
        masked_scan(None)
        return out_d, out_r, stats
    status = torch.zeros((Q,), dtype=torch.int32, device=dev)
    nbytes = lib.vdb_hnsw_search_workspace(Q, ef, deg0)
    hash_slots = 0
    for attempt in range(4):
        ws = _ws("hnsw", nbytes, dev)
        check(lib.vdb_hnsw_search(ptr(q), Q, ptr(db), n_rows, D, ptr(adj0), deg0, ptr(adj_up), deg_up, ptr(up_ptr),
                                  entry_row, max_level, ef, k, ptr(alive), ptr(flt), ptr(levels), hash_slots, ptr(out_d),
                                  ptr(out_r), ptr(stats), ptr(status), ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
        flag = int(ws[:4].view(torch.int32).item()) if Q else 0
        if flag == 0:
            return out_d, out_r, stats
        if flt is not None and (flag & 2 or attempt >= 1):
            masked_scan(torch.nonzero(status).flatten())
            return out_d, out_r, stats
        # a visited table filled up: force a (4x) larger table in global memory and retry
        base = 1024
        want = (ef + 32) * deg0 * 5 // 4
        while base < want:
            base *= 2
        hash_slots = base * 4 if hash_slots == 0 else hash_slots * 4
        nbytes = 256 + 32 * torch.cuda.get_device_properties(dev).multi_processor_count * hash_slots * 4
    raise _lib.VdbError("hnsw_search: visited table overflow persists")


def tc_supported(Q: int, N: int, D: int, k: int, metric=METRIC_L2) -> bool:
    """True when the tcgen05 tensor-core scan covers this shape (see vdb_exact_topk_tc)."""
    if Q <= 0 or N < 256 or (N < 4096 and Q * N < (1 << 24)):
        return False          # tiny problems: the fp32 scan is already launch-bound (k-means assignment of many
                              # rows to a few hundred centroids is not tiny: 7.5 ms against 20-40 ms per 200k rows)
    return bool(load().vdb_exact_topk_tc_supported(int(Q), int(N), int(D), int(k), _metric(metric)))


def tc_shape_supported(Q: int, N: int, D: int, k: int, metric=METRIC_L2) -> bool:
    """vdb_exact_topk_tc_supported alone (tc_supported adds a "worth it" size rule)."""
    return Q > 0 and N >= 256 and bool(load().vdb_exact_topk_tc_supported(int(Q), int(N), int(D), int(k), _metric(metric)))


def tc_prepare_rows(vecs: torch.Tensor, metric=METRIC_L2):
    """Per-metric bf16 operand mirror for the tensor-core scan.  Returns (mirror: flat bf16 buffer of
    vdb_tc_mirror_elems(N, D, metric) elements -- the N x D operand rows, then for L2 the packed norm block --,
    sqnorm f32 [N], stats f32 [2] = max operand norm / max bf16 rounding residual norm over the rows, which the
    certificate of exact_topk_tc needs); the accumulator of q~ . mirror_row~ is the ranking score itself."""
    vecs = _f32(vecs, "vecs")
    n, d = vecs.shape
    m = _metric(metric)
    lib = load()
    mir = torch.empty((lib.vdb_tc_mirror_elems(n, d, m),), dtype=torch.bfloat16, device=vecs.device)
    sq = torch.empty(n, dtype=torch.float32, device=vecs.device)
    stats = torch.empty(2, dtype=torch.float32, device=vecs.device)
    check(lib.vdb_tc_prepare_rows(ptr(vecs), n, d, m, ptr(sq), ptr(mir), ptr(stats), _lib.stream_ptr(vecs.device)))
    return mir, sq, stats


def exact_topk_tc(q: torch.Tensor, db: torch.Tensor, mirror: torch.Tensor, stats: torch.Tensor,
                  k: int, metric=METRIC_L2, ids: Optional[torch.Tensor] = None, n_rows: Optional[int] = None,
                  return_uncertified: bool = False):
    """Exact top-k through the tensor-core coarse scan + fp32 rerank + certificate.  `mirror` and `stats`
    come from tc_prepare_rows(db, metric).  Queries the certificate rejects are re-run through the exact
    fp32 scan inside the same library call (device-side list), so the result is always exact and the
    call never synchronises.  With ``return_uncertified`` the third value is a 0-dim int32 DEVICE
    tensor holding the number of such queries (reading it synchronises)."""
    q = _f32(q, "q")
    Q, D = q.shape
    N = db.shape[0] if n_rows is None else int(n_rows)
    dev = q.device
    m = _metric(metric)
    out_d = torch.empty((Q, k), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=dev)
    uncert = torch.empty((Q + 1,), dtype=torch.int32, device=dev)
    lib = load()
    if stats.dtype != torch.float32 or stats.numel() != 2 or not stats.is_cuda:
        raise ValueError("stats must be the 2-float device tensor of tc_prepare_rows")
    if mirror.numel() != lib.vdb_tc_mirror_elems(N, D, m):
        raise ValueError("operand mirror was built for another metric / shape")
    nbytes = lib.vdb_exact_topk_tc_workspace(Q, N, D, k, m)
    if nbytes == 0:
        raise ValueError(f"tensor-core scan does not support Q={Q} N={N} D={D} k={k}")
    ws = _ws("exact_tc", nbytes, dev)
    check(lib.vdb_exact_topk_tc(ptr(q), Q, ptr(db), ptr(mirror), ptr(stats), ptr(ids), N, D, k, m,
                                ptr(out_d), ptr(out_i), ptr(uncert), ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    if return_uncertified:
        return out_d, out_i, uncert[0]
    return out_d, out_i


def tc_norm_offset(n: int, d: int) -> int:
    """Element offset of the packed norm block inside an L2 operand mirror of n rows (see tc_prepare_rows)."""
    return int(load().vdb_tc_mirror_elems(int(n), int(d), METRIC_L2)) - ((int(n) + 3) & ~3) * 16


def ivf_topk_tc(q: torch.Tensor, lm_vecs: torch.Tensor, lm_mirror: torch.Tensor, lm_stats: torch.Tensor,
                lm_ids: Optional[torch.Tensor], k: int, n_probe: int, pair_rows: torch.Tensor, row_query: torch.Tensor,
                row_list: torch.Tensor, centroids: torch.Tensor, unit_table: torch.Tensor,
                list_stats: Optional[torch.Tensor] = None, meta: Optional[torch.Tensor] = None):
    """List-major IVF search on the tensor-core pipeline (vdb_ivf_topk_tc).  ``list_stats`` f32 [K, 2]: per-list row bounds
    (max |bf16(r)|, max |bf16(r) - r| over the list's residual rows r) that tighten the certificate.  Returns (dist f32 [Q,k], ids-or-positions i64
    [Q,k], uncert i32 [Q+1] = count + indices of the queries the caller must re-run through ivf_scan).  ``meta`` (from
    ivf_group_pairs): the operand-row / unit counts live on the device and the tables are sized for the worst case."""
    q = _f32(q, "q")
    Q, D = q.shape
    N = lm_vecs.shape[0]
    Qa = row_query.numel()
    U = unit_table.shape[0]
    dev = q.device
    out_d = torch.empty((Q, k), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=dev)
    uncert = torch.empty((Q + 1,), dtype=torch.int32, device=dev)
    lib = load()
    nbytes = lib.vdb_ivf_topk_tc_workspace(Q, Qa, N, D, k, U)
    if nbytes == 0:
        raise ValueError(f"list-major IVF scan does not support Q={Q} N={N} D={D} k={k}")
    ws = _ws("ivf_tc", nbytes, dev)
    check(lib.vdb_ivf_topk_tc(ptr(q), Q, ptr(lm_vecs), ptr(lm_mirror), ptr(lm_stats), ptr(lm_ids), N, D, k, n_probe,
                              ptr(pair_rows), ptr(row_query), ptr(row_list), ptr(_f32(centroids, "centroids")), Qa, ptr(unit_table), U,
                              ptr(meta), ptr(list_stats), ptr(out_d), ptr(out_i), ptr(uncert), ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return out_d, out_i, uncert


def ivf_group_pairs(probes: torch.Tensor, list_tile_ptr: torch.Tensor):
    """(query, probe) pairs grouped by list on the device (vdb_ivf_group_pairs) -> (row_query, row_list, pair_rows, unit_table,
    meta), the tables sized for the worst case."""
    Q, n_probe = probes.shape
    K = list_tile_ptr.numel() - 1
    P = Q * n_probe
    qa_max = ((P + 127) // 128 + min(K, P)) * 128
    dev = probes.device
    row_query = torch.empty((qa_max,), dtype=torch.int32, device=dev)
    row_list = torch.empty((qa_max,), dtype=torch.int32, device=dev)
    pair_rows = torch.empty((P,), dtype=torch.int32, device=dev)
    utab = torch.empty((qa_max // 128, 4), dtype=torch.int32, device=dev)
    meta = torch.empty((2,), dtype=torch.int32, device=dev)
    lib = load()
    ws = _ws("ivf_group", lib.vdb_ivf_group_workspace(K), dev)
    check(lib.vdb_ivf_group_pairs(ptr(probes), Q, n_probe, ptr(list_tile_ptr), K, qa_max, ptr(row_query), ptr(row_list), ptr(pair_rows),
                                  ptr(utab), ptr(meta), ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return row_query, row_list, pair_rows, utab, meta


def ivf_tc_finish(d: torch.Tensor, pos: torch.Tensor, top_k: int, probes: torch.Tensor, prefix: torch.Tensor, lm_list: torch.Tensor,
                  lm_start: torch.Tensor, lm_ids: torch.Tensor):
    """vdb_ivf_tc_finish: list-major positions -> (dist, ids) in the reference's stable order + tie flags (int32 [Q])."""
    Q, k2 = d.shape
    dev = d.device
    out_d = torch.empty((Q, top_k), dtype=torch.float32, device=dev)
    out_i = torch.empty((Q, top_k), dtype=torch.int64, device=dev)
    tie = torch.empty((Q,), dtype=torch.int32, device=dev)
    check(load().vdb_ivf_tc_finish(ptr(d), ptr(pos), Q, k2, top_k, ptr(probes), ptr(prefix), probes.shape[1], ptr(lm_list), ptr(lm_start),
                                   ptr(lm_ids), ptr(out_d), ptr(out_i), ptr(tie), _lib.stream_ptr(dev)))
    return out_d, out_i, tie
