"""GPU parity: PQ encode (bit-exact vs the reference's codes), k-means step, LUT and ADC scan."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests._util import check_topk

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def test_encode_bit_exact_vs_reference_codes(golden_dir):
    from vector_db_b200 import ops
    g = np.load(os.path.join(golden_dir, "pq.npz"))           # dsub = 4 (plain loop order)
    cb = _t(g["centroids"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g["Y"]), cb).cpu().numpy(), g["codes_y"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g["X"]), cb).cpu().numpy(), g["codes"])
    g8 = np.load(os.path.join(golden_dir, "pq8.npz"))          # dsub = 8 (numpy's 8-way tree)
    cb8 = _t(g8["centroids"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g8["Y"]), cb8).cpu().numpy(), g8["codes_y"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g8["X"]), cb8).cpu().numpy(), g8["codes"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g8["X12"]), _t(g8["centroids12"])).cpu().numpy(), g8["codes12"])


def test_service_kats_and_compress_contract(golden_dir):
    from vector_db_b200.inference.pq import ProductQuantizationService as PQ
    svc = PQ(k=3, chunks=2, dim=4)
    c = np.array([[0, 0], [1, 1], [2, 2]], dtype=np.float32)
    assert svc._find_nearest_centroid(np.array([0.9, 1.1], np.float32), c) == 1          # test_pq.py:78-90
    out = svc._compress_embedding(np.array([0.9, 1.1, 2.1, 1.9], np.float32), [c, c])   # test_pq.py:92-116
    assert out.dtype == np.int64 and out.tolist() == [1, 2]
    rng = np.random.default_rng(0)
    X = rng.random((500, 32), dtype=np.float32)
    svc = PQ(k=16, chunks=4, dim=32)
    np.random.seed(1)
    codes = svc.compress(X)                                                               # test_pq.py:118-176
    assert codes.shape == (500, 4) and codes.dtype == np.int64 and codes.min() >= 0 and codes.max() < 16
    assert len(svc.centroids) == 4 and all(cc.shape == (16, 8) and cc.dtype == np.float32 for cc in svc.centroids)
    np.testing.assert_array_equal(codes, O.pq_encode(X, svc.centroids))
    with pytest.raises(ValueError):
        PQ(k=600, chunks=4, dim=32).compress(X)


def test_kmeans_matches_lloyd_restatement():
    from vector_db_b200.inference.pq import ProductQuantizationService as PQ
    rng = np.random.default_rng(3)
    centers = rng.standard_normal((8, 8)).astype(np.float32) * 4
    X = (centers[rng.integers(0, 8, 3000)] + 0.3 * rng.standard_normal((3000, 8))).astype(np.float32)
    init = rng.choice(3000, 8, replace=False)
    svc = PQ(k=8, chunks=2, dim=8)
    cents = svc.fit(X, init_rows=init, iters=25)
    for j in range(2):
        ref, _ = O.kmeans_lloyd(X[:, 4 * j:4 * j + 4], X[init, 4 * j:4 * j + 4], 25)
        np.testing.assert_allclose(cents[j], ref, rtol=1e-3, atol=1e-3)


def test_kmeans_round_through_tensor_core_encode_matches_lloyd_restatement():
    """dsub == 8 and >= 4096 rows: the assignment of a Lloyd round is one tensor-core encode call and the round is
    the mean update over its codes (vdb_pq_kmeans_step_coded)."""
    from vector_db_b200.inference.pq import ProductQuantizationService as PQ
    rng = np.random.default_rng(5)
    centers = rng.standard_normal((16, 16)).astype(np.float32) * 4
    X = (centers[rng.integers(0, 16, 6000)] + 0.3 * rng.standard_normal((6000, 16))).astype(np.float32)
    init = rng.choice(6000, 16, replace=False)
    svc = PQ(k=16, chunks=2, dim=16)
    cents = svc.fit(X, init_rows=init, iters=20)
    for j in range(2):
        ref, _ = O.kmeans_lloyd(X[:, 8 * j:8 * j + 8], X[init, 8 * j:8 * j + 8], 20)
        np.testing.assert_allclose(cents[j], ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("n,m,dsub,q,R", [(30000, 16, 4, 9, 10), (20000, 48, 2, 5, 10), (20000, 48, 2, 3, 700),
                                         (5000, 5, 3, 2, 17), (100, 8, 4, 6, 10), (9000, 32, 2, 1, 64)])
def test_lut_and_adc_vs_oracle(n, m, dsub, q, R):
    from vector_db_b200 import ops
    rng = np.random.default_rng(n + m)
    d = m * dsub
    cents = [rng.standard_normal((256, dsub)).astype(np.float32) for _ in range(m)]
    codes = rng.integers(0, 256, (n, m)).astype(np.uint8)
    Q = rng.standard_normal((q, d)).astype(np.float32)
    lut_ref = O.adc_lut(Q, cents)
    lut = ops.pq_lut(_t(Q), _t(np.stack(cents)))
    np.testing.assert_array_equal(lut.cpu().numpy(), lut_ref)          # same fp32 operation order
    R_eff = min(R, n)
    Dref, Iref = O.adc_scan_topk(codes, lut_ref, R_eff)
    D, I = ops.pq_adc_topk(_t(codes), lut, R_eff)
    np.testing.assert_array_equal(D.cpu().numpy(), Dref)
    np.testing.assert_array_equal(I.cpu().numpy(), Iref)


def test_golden_adc_pinned_to_reference_decode_distance(golden_dir):
    """tests/golden/adc.npz: euclidean_vector_distance(q, decode(code)) computed by the reference on its own
    centroids and codes (oracle/gen_golden.py:gen_adc).  LUT + ADC scan on the GPU must give that (dist, id)
    order and the squared distances to fp32 rounding -- through the raw ops and through PQIndex."""
    from vector_db_b200 import ops
    from vector_db_b200.inference.pq import PQIndex
    a = np.load(os.path.join(golden_dir, "adc.npz"))
    for tag in ("pq", "pq8"):
        g = np.load(os.path.join(golden_dir, f"{tag}.npz"))
        cents = g["centroids"]
        m, K, dsub = cents.shape
        cb = np.concatenate([cents, np.repeat(cents[:, :1, :], 256 - K, axis=1)], axis=1)     # pad LUT rows to 256
        lut = ops.pq_lut(_t(a[f"{tag}_Q"]), _t(cb))
        D, I = ops.pq_adc_topk(_t(g["codes"].astype(np.uint8)), lut, 10)
        ref_sq = (a[f"{tag}_D"].astype(np.float32) ** 2).astype(np.float32)
        check_topk(D.cpu().numpy(), I.cpu().numpy(), ref_sq, a[f"{tag}_I"], rtol=1e-5, atol=1e-5)
        assert D[0, 0].item() < 1e-6 and I[0, 0].item() == a[f"{tag}_I"][0, 0]
        idx = PQIndex(list(cents), keep_vectors=False)
        idx.add(g["X"])                                                # GPU encode == the reference's codes (pinned above)
        D2, I2 = idx.search(a[f"{tag}_Q"], 10)
        np.testing.assert_array_equal(I2, I.cpu().numpy())
        np.testing.assert_array_equal(D2, D.cpu().numpy())


@pytest.mark.parametrize("n,m,q,R,dup", [(60_000, 16, 5, 1500, False), (40_000, 48, 3, 600, False),
                                          (40_000, 8, 6, 600, True), (70_000, 24, 2, 4000, False),
                                          (70_000, 16, 4, 300, False), (40_000, 8, 6, 50, True), (40_000, 7, 3, 120, False)])
def test_adc_large_shortlist_threshold_collect_vs_oracle(n, m, q, R, dup):
    """N >= 16 x the collect target takes the threshold-collect path (sampled threshold -> collect -> exact merge, flagged
    queries rescanned by the running-selection kernel).  dup=True makes half the rows identical, so for
    some queries the collected set overflows its buffer and the device-side rescan list is exercised."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(n + m + R)
    codes = rng.integers(0, 256, (n, m)).astype(np.uint8)
    if dup:
        codes[rng.permutation(n)[: n // 2]] = codes[0]
    lut = rng.random((q, m, 256), dtype=np.float32) * np.float32(3.0)
    ids = np.sort(rng.choice(10 * n, n, replace=False)).astype(np.int64)      # increasing: ties break by row == by id
    Dref, Iref = O.adc_scan_topk(codes, lut, R, ids=ids)
    D, I = ops.pq_adc_topk(_t(codes), _t(lut), R, ids=_t(ids))
    np.testing.assert_array_equal(I.cpu().numpy(), Iref)
    np.testing.assert_array_equal(D.cpu().numpy(), Dref)


def test_pq_index_end_to_end_recall_with_rerank(golden_dir):
    from vector_db_b200.inference.pq import PQIndex
    g = np.load(os.path.join(golden_dir, "pq8.npz"))
    cents = list(g["centroids"])                                       # K = 64 < 256: LUT rows padded
    X = g["X"]
    idx = PQIndex(cents, keep_vectors=True)
    idx.add(X, ids=np.arange(1200) + 5000)
    Q = g["Y"][:16]
    D, I = idx.search(Q, 10)
    Dref, Iref = O.adc_scan_topk(g["codes"], O.adc_lut(Q, cents), 10, ids=np.arange(1200) + 5000)
    np.testing.assert_array_equal(I, Iref)
    np.testing.assert_array_equal(D, Dref)
    De, Ie = idx.search(Q, 10, shortlist=1200, rerank=True)            # full shortlist -> exact answer
    Dx, Ix = O.exact_l2_topk(X, np.arange(1200) + 5000, Q, 10)
    np.testing.assert_array_equal(Ie, Ix)
    np.testing.assert_allclose(De, Dx, rtol=1e-5)


@pytest.mark.parametrize("n,m,K,scale", [(20_000, 16, 256, 1.0), (9_001, 48, 256, 1.0), (5_000, 4, 64, 30.0),
                                          (4_100, 6, 200, 1e-3)])
def test_tensor_core_encode_bit_exact_vs_fp32_kernel_and_oracle(n, m, K, scale):
    """tcgen05 bf16-split candidate pass + exact tie resolution == the fp32 kernel (pinned to the reference's
    codes above) on random data, with duplicated centroids (exact ties -> first index wins), rows that
    sit on a centroid and rows between two centroids."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(n + m)
    cb = (rng.standard_normal((m, K, 8)) * scale).astype(np.float32)
    cb[:, K // 2] = cb[:, 3]                                   # duplicate centroid: index 3 must win
    X = (rng.standard_normal((n, 8 * m)) * scale).astype(np.float32)
    X[:50] = cb[:, 7].reshape(-1)                              # rows exactly on centroid 7 of every sub-space
    X[50:100] = (0.5 * (cb[:, 10].astype(np.float64) + cb[:, 11])).astype(np.float32).reshape(-1)   # midpoints
    X[100:150] = cb[:, 3].reshape(-1)                          # on the duplicated centroid
    cbd, Xd = _t(cb), _t(X)
    tc = ops.pq_encode(Xd, cbd, engine="tc").cpu().numpy()
    f32 = ops.pq_encode(Xd, cbd, engine="f32").cpu().numpy()
    np.testing.assert_array_equal(tc, f32)
    assert (tc[:50] == 7).all() and (tc[100:150] == 3).all()
    ref = O.pq_encode(X[:600], list(cb))
    np.testing.assert_array_equal(tc[:600], ref)


@pytest.mark.parametrize("offset,spread", [(100.0, 1.0), (-3000.0, 0.05), (0.0, 1e-18)])
def test_tensor_core_encode_cancellation_and_tiny_inputs(offset, spread):
    """Scores |c|^2 - 2 x.c that cancel catastrophically (rows and centroids far from the origin, close to each
    other) and inputs whose scores underflow the margin test's exponent range: the candidate margin grows / the
    test falls back to the exact path, codes stay equal to the fp32 kernel's and the oracle's."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(99)
    n, m, K = 6_000, 8, 256
    cb = (offset + spread * rng.standard_normal((m, K, 8))).astype(np.float32)
    X = (offset + spread * rng.standard_normal((n, 8 * m))).astype(np.float32)
    tc = ops.pq_encode(_t(X), _t(cb), engine="tc").cpu().numpy()
    f32 = ops.pq_encode(_t(X), _t(cb), engine="f32").cpu().numpy()
    np.testing.assert_array_equal(tc, f32)
    np.testing.assert_array_equal(tc[:300], O.pq_encode(X[:300], list(cb)))


def test_tensor_core_encode_golden_reference_codes(golden_dir):
    from vector_db_b200 import ops
    g8 = np.load(os.path.join(golden_dir, "pq8.npz"))          # dsub = 8, K = 64: codes from the reference itself
    cb8 = _t(g8["centroids"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g8["X"]), cb8, engine="tc").cpu().numpy(), g8["codes"])
    np.testing.assert_array_equal(ops.pq_encode(_t(g8["Y"]), cb8, engine="tc").cpu().numpy(), g8["codes_y"])


@pytest.mark.parametrize("n,m,q,R,keep", [(60_000, 16, 4, 10, 0.5), (60_000, 16, 4, 10, 0.01), (3_000, 8, 3, 20, 0.3),
                                           (50_000, 48, 2, 800, 0.7)])
def test_adc_row_mask_is_exact_over_the_remaining_rows(n, m, q, R, keep):
    """Metadata pre-filter / tombstones: rows whose bit is cleared are skipped; the answer equals the oracle on
    the kept rows.  A selective filter makes the sampled threshold collect fewer than R rows, so the
    device-side rescan path is what answers (keep = 0.01)."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(n + m + int(keep * 100))
    codes = rng.integers(0, 256, (n, m)).astype(np.uint8)
    lut = rng.random((q, m, 256), dtype=np.float32)
    kept = np.flatnonzero(rng.random(n) < keep)
    words = np.zeros((n + 31) // 32, dtype=np.uint32)
    np.bitwise_or.at(words, kept >> 5, np.uint32(1) << (kept & 31).astype(np.uint32))
    R_eff = min(R, kept.size)
    Dref, Iref = O.adc_scan_topk(codes[kept], lut, R_eff, ids=kept)
    D, I = ops.pq_adc_topk(_t(codes), _t(lut), R_eff, row_mask=_t(words.view(np.int32)))
    np.testing.assert_array_equal(I.cpu().numpy(), Iref)
    np.testing.assert_array_equal(D.cpu().numpy(), Dref)


def test_pq_index_filter_ids(golden_dir):
    from vector_db_b200.inference.pq import PQIndex
    g = np.load(os.path.join(golden_dir, "pq8.npz"))
    cents, X = list(g["centroids"]), g["X"]
    ids = np.arange(1200) + 5000
    idx = PQIndex(cents, keep_vectors=True)
    idx.add(X, ids=ids)
    allowed = set(int(i) for i in ids[::3])
    D, I = idx.search(g["Y"][:8], 10, filter_ids=allowed)
    keep = np.array(sorted(allowed)) - 5000
    Dref, Iref = O.adc_scan_topk(g["codes"][keep], O.adc_lut(g["Y"][:8], cents), 10, ids=ids[keep])
    np.testing.assert_array_equal(I, Iref)
    np.testing.assert_array_equal(D, Dref)


@pytest.mark.gpu
def test_f16_accumulator_encode_equals_fp32_accumulator_kernel_and_handles_rows_outside_the_f16_range():
    """The two tensor-core encoders (vdb_debug_set_pq_tc_mode 1 = f16 accumulators, the default; 0 = fp32 accumulators)
    and the fp32 kernel agree code for code -- including rows whose scaled norm overflows f16 (resolver with all-ones
    masks), rows of +-inf / NaN (numpy's argmin over all-NaN distances is 0), a single row, K = 1 and K = 2."""
    from vector_db_b200 import _lib, ops
    lib = _lib.load()
    rng = np.random.default_rng(17)
    try:
        for n, m, K in ((7_000, 12, 256), (1, 5, 256), (300, 3, 1), (513, 7, 2)):
            cb = rng.standard_normal((m, K, 8)).astype(np.float32)
            X = rng.standard_normal((n, 8 * m)).astype(np.float32)
            if n > 100:
                X[3] *= 1e4                                   # far outside the codebook's range: S^2 |x|^2 > 30000
                X[4] *= 1e18
                X[5, :8] = np.nan
                X[6, 8:16] = np.inf
                X[7] = 0.0
            got = {}
            for mode in (1, 0):
                lib.vdb_debug_set_pq_tc_mode(mode)
                got[mode] = ops.pq_encode(_t(X), _t(cb), engine="tc").cpu().numpy()
            f32 = ops.pq_encode(_t(X), _t(cb), engine="f32").cpu().numpy()
            np.testing.assert_array_equal(got[1], f32)
            np.testing.assert_array_equal(got[0], f32)
            with np.errstate(all="ignore"):
                np.testing.assert_array_equal(got[1][:64], O.pq_encode(X[:64], list(cb)))
    finally:
        lib.vdb_debug_set_pq_tc_mode(1)


@pytest.mark.gpu
@pytest.mark.parametrize("n,m,q,R,kind", [(80_000, 48, 19, 10, "plain"), (80_000, 16, 9, 100, "tail"), (60_000, 7, 8, 10, "plain"),
                                           (50_000, 64, 11, 10, "tail"), (80_000, 16, 13, 10, "dup"), (70_000, 24, 6, 900, "mask"),
                                           (80_000, 48, 17, 10, "const")])
def test_adc_collect_on_16bit_lower_bound_tables_equals_the_32bit_collect_and_the_oracle(n, m, q, R, kind):
    """More than four queries: the collect pass tests rows against 15-bit lower-bound tables (eight queries per
    shared-memory lookup, adc_coarse_collect_kernel) and re-evaluates the collected rows on the 32-bit tables.
    Same (d_int, row) keys as the 32-bit collect, hence the oracle's result bit for bit -- on heavy-tailed tables
    (entries clipped at the threshold budget), duplicates (buffer overflow -> device-side rescan), a row mask,
    constant tables (every row ties), ragged query groups and m outside the unrolled cases (7; 64 -> four
    queries per lookup)."""
    from vector_db_b200 import _lib, ops
    lib = _lib.load()
    rng = np.random.default_rng(n + m + R + q)
    codes = rng.integers(0, 256, (n, m)).astype(np.uint8)
    lut = rng.random((q, m, 256), dtype=np.float32) * np.float32(3.0)
    mask = None
    kept = None
    if kind == "tail":
        lut[rng.random(lut.shape) < 0.01] *= np.float32(5000.0)
    elif kind == "dup":
        codes[rng.permutation(n)[: n // 2]] = codes[0]
    elif kind == "const":
        lut[:] = np.float32(1.25)
        lut[1] = rng.random((m, 256), dtype=np.float32)
    elif kind == "mask":
        kept = np.flatnonzero(rng.random(n) < 0.6)
        words = np.zeros((n + 31) // 32, dtype=np.uint32)
        np.bitwise_or.at(words, kept >> 5, np.uint32(1) << (kept & 31).astype(np.uint32))
        mask = _t(words.view(np.int32))
    if kept is None:
        Dref, Iref = O.adc_scan_topk(codes, lut, R)
    else:
        Dref, Iref = O.adc_scan_topk(codes[kept], lut, R, ids=kept)
    out = {}
    try:
        for mode in (1, 0):
            lib.vdb_debug_set_adc_coarse(mode)
            D, I = ops.pq_adc_topk(_t(codes), _t(lut), R, row_mask=mask)
            out[mode] = (D.cpu().numpy(), I.cpu().numpy())
    finally:
        lib.vdb_debug_set_adc_coarse(1)
    np.testing.assert_array_equal(out[1][1], out[0][1])
    np.testing.assert_array_equal(out[1][0], out[0][0])
    np.testing.assert_array_equal(out[1][1], Iref)
    np.testing.assert_array_equal(out[1][0], Dref)


@pytest.mark.gpu
@pytest.mark.parametrize("m,q,G,kind", [(48, 19, 8, "plain"), (16, 8, 8, "tail"), (7, 5, 8, "plain"), (64, 6, 4, "tail"), (24, 3, 2, "wide")])
def test_adc_lower_bound_tables_equal_the_restatement(m, q, G, kind):
    """adc_coarse_lut_kernel == oracle.adc_coarse_tables, entry for entry (the CPU suite proves on the restatement that
    these tables never drop a row under the threshold)."""
    from vector_db_b200 import _lib
    from vector_db_b200._lib import check, ptr
    lib = _lib.load()
    rng = np.random.default_rng(m * 100 + q)
    lut = rng.random((q, m, 256)).astype(np.float32) * np.float32(3.0)
    if kind == "tail":
        lut[rng.random(lut.shape) < 0.01] *= np.float32(5000.0)
    lutq, _ = O.adc_quantize_lut(lut)
    off = lutq.min(axis=2).sum(axis=1)
    tau = off + rng.integers(1, 1 << 22, q) if kind != "wide" else np.full(q, 0xFFFFFFFF, dtype=np.int64)
    ref16, thr, _, _ = O.adc_coarse_tables(lutq, tau)
    mb = (m + 7) // 8
    groups = (q + G - 1) // G
    out16 = torch.full((groups, mb * 2048, G), -1, dtype=torch.int16, device="cuda")
    thr16 = torch.zeros(q, dtype=torch.int32, device="cuda")
    check(lib.vdb_debug_adc_coarse_tables(ptr(_t(lutq.astype(np.uint32).view(np.int32))), m, q,
                                          ptr(_t(tau.astype(np.uint32).view(np.int32))), ptr(out16), ptr(thr16), G,
                                          _lib.stream_ptr(out16.device)))
    got = out16.cpu().numpy().view(np.uint16).astype(np.int64)
    np.testing.assert_array_equal(thr16.cpu().numpy().astype(np.int64), 0x8000 + thr)
    want = np.zeros((groups, mb * 2048, G), dtype=np.int64)        # padded sub-spaces and absent queries read 0
    j, c = np.meshgrid(np.arange(m), np.arange(256), indexing="ij")
    slot = ((j >> 3) * 256 + c) * 8 + (j & 7)
    for qi in range(q):
        want[qi // G, slot, qi % G] = ref16[qi]
    np.testing.assert_array_equal(got, want)
