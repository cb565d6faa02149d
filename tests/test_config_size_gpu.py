"""GPU parity AT THE BASELINE.json CONFIG SIZES: the real launch shapes -- 1M x 384 / Q = 4096
(9 db slices, 288 units, 2 waves, the global threshold word crossing waves), 10M x 48 codes / Q = 1024, 2M x 384
encode calls -- each against a bounded oracle sample (id EQUALITY, tie-aware) and, where a second exact kernel
exists, against it on every query."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests._util import check_topk

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
N, D, Q, K = 1_000_000, 384, 4096, 10


def _randn(shape, seed):
    return torch.randn(shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(seed))


@pytest.fixture(scope="module")
def db():
    X = _randn((N, D), 1234)
    Qd = _randn((Q, D), 4321)
    Qd[5] = X[777]                                   # self query
    yield X, Qd
    del X, Qd
    torch.cuda.empty_cache()


@pytest.mark.parametrize("metric", ["cosine", "l2"])
def test_exact_scan_at_config_2(db, metric):
    """BASELINE configs[1]: 1M x 384, Q = 4096, k = 10 through vdb_exact_topk_tc.  Every query against the exact
    fp32 kernel (ids and values bit-equal: the two share the final fp32 arithmetic), 32 queries -- one per 128-query
    tile -- against the oracle."""
    from vector_db_b200 import ops
    X, Qd = db
    assert ops.tc_supported(Q, N, D, K, metric)
    mir, _, stats = ops.tc_prepare_rows(X, metric)
    Dt, It, n_bad = ops.exact_topk_tc(Qd, X, mir, stats, K, metric, return_uncertified=True)
    Df, If = ops.exact_topk_f32(Qd, X, K, metric)
    assert torch.equal(It, If), f"{(It != If).any(dim=1).sum().item()} of {Q} queries differ from the fp32 scan"
    assert torch.equal(Dt, Df)
    assert int(n_bad) <= Q // 100, f"{int(n_bad)} queries fell back to the fp32 rescan"
    sel = np.arange(0, Q, 128) + 5                   # 32 queries, one per query tile (query 5 is the self query)
    Xh, Qh = X.cpu().numpy(), Qd[torch.from_numpy(sel).to(DEV)].cpu().numpy()
    ids = np.arange(N, dtype=np.int64)
    Vref, Iref = (O.exact_cosine_topk if metric == "cosine" else O.exact_l2_topk)(Xh, ids, Qh, K)
    got_v, got_i = Dt[torch.from_numpy(sel).to(DEV)].cpu().numpy(), It[torch.from_numpy(sel).to(DEV)].cpu().numpy()
    tied = check_topk(got_v, got_i, Vref, Iref, rtol=1e-5, atol=2e-6, max_tied_rows=1)
    assert got_i[0, 0] == 777 and (got_v[0, 0] == 0.0 if metric == "l2" else abs(got_v[0, 0] - 1.0) < 1e-6)
    assert tied <= 1


def test_exact_scan_ties_across_waves_at_1m_rows():
    """1M rows that are copies of a few distinct vectors: whole bins, lists and both waves hold EQUAL scores, so the
    shared threshold (tagged shared-memory word + the global per-query word that crosses waves, exact_tc.cu) is
    exercised on ties.  The answer must be the fp32 scan's, smaller ids first; and a database of identical rows must
    return ids 0..9."""
    from vector_db_b200 import ops
    base = _randn((64, D), 7)
    X = base[torch.arange(N, device=DEV) % 64].contiguous()              # row r == base[r % 64]: 15,625 copies each
    Qd = torch.cat([base[:32] + 0.01 * _randn((32, D), 8), _randn((224, D), 9)])
    for metric in ("l2", "cosine"):
        mir, _, stats = ops.tc_prepare_rows(X, metric)
        Dt, It = ops.exact_topk_tc(Qd, X, mir, stats, K, metric)
        Df, If = ops.exact_topk_f32(Qd, X, K, metric)
        assert torch.equal(It, If) and torch.equal(Dt, Df)
        # the ten best are the ten smallest ids of the nearest vector's copies: j, j + 64, ...
        j = It[0, 0].item()
        assert It[0].tolist() == [j + 64 * t for t in range(K)]
    X[:] = base[3]
    mir, _, stats = ops.tc_prepare_rows(X, "l2")
    Dt, It = ops.exact_topk_tc(Qd[:128], X, mir, stats, K, "l2")
    assert It.tolist() == [list(range(K))] * 128


def _oracle_adc_threaded(codes_h, lut_h, k, threads=16):
    n = codes_h.shape[0]
    step = -(-n // threads)
    bounds = [(s, min(n, s + step)) for s in range(0, n, step)]
    with ThreadPoolExecutor(max_workers=threads) as ex:
        parts = list(ex.map(lambda b: O.adc_scan_topk(codes_h[b[0]:b[1]], lut_h, k, ids=np.arange(b[0], b[1], dtype=np.int64)), bounds))
    Dc = np.concatenate([p[0] for p in parts], axis=1)
    Ic = np.concatenate([p[1] for p in parts], axis=1)
    out_d, out_i = np.empty((lut_h.shape[0], k), np.float32), np.empty((lut_h.shape[0], k), np.int64)
    for qi in range(lut_h.shape[0]):
        order = np.lexsort((Ic[qi], Dc[qi]))[:k]
        out_d[qi], out_i[qi] = Dc[qi][order], Ic[qi][order]
    return out_d, out_i


def test_pq_encode_and_adc_scan_at_config_3():
    """BASELINE configs[2]: 2M x 384 rows per encode call (m = 48, K = 256) -- 2000 sampled rows over the first /
    middle / last row tiles against oracle.pq_encode, bit-exact; then the ADC scan of 10M x 48 codes with Q = 1024
    against oracle.adc_scan_topk over ALL rows on 4 queries, bit-equal ids and distances."""
    from vector_db_b200 import ops
    m, CH = 48, 2_000_000
    cb = _randn((m, 256, 8), 11)
    cents = list(cb.cpu().numpy())
    codes = torch.empty((10_000_000, m), dtype=torch.uint8, device=DEV)
    for c in range(5):
        X = _randn((CH, D), 100 + c)
        codes[c * CH:(c + 1) * CH] = ops.pq_encode(X, cb, engine="tc")
        if c in (0, 4):
            rows = np.unique(np.concatenate([np.arange(0, 400), np.arange(CH // 2 - 100, CH // 2 + 100), np.arange(CH - 400, CH)]))
            ref = O.pq_encode(X[torch.from_numpy(rows).to(DEV)].cpu().numpy(), cents)
            got = codes[c * CH + torch.from_numpy(rows).to(DEV)].cpu().numpy().astype(np.int64)
            np.testing.assert_array_equal(got, ref)
        del X
    Qd = _randn((1024, D), 4321)
    lut = ops.pq_lut(Qd, cb)
    Dg, Ig = ops.pq_adc_topk(codes, lut, K)
    sel = [0, 341, 682, 1023]
    lut_ref = O.adc_lut(Qd[sel].cpu().numpy(), cents)
    np.testing.assert_array_equal(lut[sel].cpu().numpy(), lut_ref)
    Dr, Ir = _oracle_adc_threaded(codes.cpu().numpy(), lut_ref, K)
    np.testing.assert_array_equal(Ig[sel].cpu().numpy(), Ir)
    np.testing.assert_array_equal(Dg[sel].cpu().numpy(), Dr)
    assert (Ig >= 0).all() and (torch.diff(Dg, dim=1) >= 0).all()
