"""GPU parity: tensor-core exact scan (tcgen05 coarse pass + fp32 rerank + certificate) vs the fp32
scan kernel (itself pinned to the oracle / golden vectors in test_exact_gpu.py) and vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests._util import check_topk

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _prep(X, metric="l2"):
    from vector_db_b200 import ops
    mir, _, stats = ops.tc_prepare_rows(X, metric)
    return mir, stats


@pytest.mark.parametrize("n,d,q,k,metric", [(50_000, 384, 300, 10, "l2"), (20_011, 128, 129, 10, "l2"),
                                             (33_000, 64, 7, 5, "l2"), (50_000, 384, 260, 10, "cosine"),
                                             (8_200, 256, 128, 16, "cosine")])
def test_tc_matches_fp32_scan(n, d, q, k, metric):
    from vector_db_b200 import ops
    g = torch.Generator(device=DEV).manual_seed(n + d)
    X = torch.randn((n, d), device=DEV, generator=g)
    Q = torch.randn((q, d), device=DEV, generator=g)
    Q[0] = X[17]                                       # self query
    if metric == "l2":
        X[40] = X[17]                                  # duplicate row -> (dist,id) tie
    assert ops.tc_supported(q, n, d, k, metric)
    mir, mx = _prep(X, metric)
    D, I, n_bad = ops.exact_topk_tc(Q, X, mir, mx, k, metric, return_uncertified=True)
    Dr, Ir = ops.exact_topk_f32(Q, X, k, metric)
    assert torch.equal(I, Ir), f"{(I != Ir).sum().item()} ids differ"
    assert torch.equal(D, Dr)                          # final values come from the same fp32 arithmetic
    assert n_bad <= max(2, q // 20), f"{n_bad} of {q} queries fell back to the fp32 scan"
    if metric == "l2":
        assert D[0, 0].item() == 0.0 and I[0, 0].item() == 17 and I[0, 1].item() == 40


def test_tc_vs_oracle_and_ids():
    from vector_db_b200 import ops
    rng = np.random.default_rng(5)
    X = rng.standard_normal((12_000, 192)).astype(np.float32)
    Q = rng.standard_normal((40, 192)).astype(np.float32)
    ids = (np.arange(12_000, dtype=np.int64) * 5 + 3)
    Xd, Qd = torch.from_numpy(X).to(DEV), torch.from_numpy(Q).to(DEV)
    mir, mx = _prep(Xd)
    D, I = ops.exact_topk_tc(Qd, Xd, mir, mx, 10, "l2", ids=torch.from_numpy(ids).to(DEV))
    Dref, Iref = O.exact_l2_topk(X, ids, Q, 10)
    np.testing.assert_array_equal(I.cpu().numpy(), Iref)
    np.testing.assert_allclose(D.cpu().numpy(), Dref, rtol=1e-5)


def test_tc_adversarial_near_duplicates_fall_back_not_fail():
    """Rows that differ from the query by less than bf16 resolution: the coarse pass cannot rank them,
    the certificate must notice (or the shortlist must contain them) -- the answer stays exact."""
    from vector_db_b200 import ops
    g = torch.Generator(device=DEV).manual_seed(1)
    base = torch.randn((1, 128), device=DEV, generator=g)
    X = base + 1e-3 * torch.randn((20_000, 128), device=DEV, generator=g)      # a tight cluster
    Q = base + 1e-3 * torch.randn((64, 128), device=DEV, generator=g)
    mir, mx = _prep(X)
    D, I, n_bad = ops.exact_topk_tc(Q, X, mir, mx, 10, "l2", return_uncertified=True)
    Dr, Ir = ops.exact_topk_f32(Q, X, 10, "l2")
    assert torch.equal(I, Ir) and torch.equal(D, Dr)
    assert n_bad > 0          # this data cannot be certified from a bf16 pass


@pytest.mark.parametrize("metric", ["l2", "cosine"])
def test_tc_certificate_covers_both_bf16_roundings(metric):
    """ADVICE r1 (exact_tc.cu:487): the coarse score carries TWO bf16 roundings (query and row), so its error can
    reach (4u + 2u^2)|q||x|, u = 2^-8 -- attained by q = x on bf16 half-way points, where both operands round the
    same way.  Rows are 9-bit integers (257 = 0b100000001 needs nine significand bits: every odd value is a
    half-way point), queries are rows of the database and near-copies of them.  The answer must be the fp32 scan's,
    whatever the certificate decides, and the measured residual stats must show the full half-ulp residual."""
    from vector_db_b200 import ops
    g = torch.Generator(device=DEV).manual_seed(9)
    n, d, nq = 30_000, 128, 192
    X = (2 * torch.randint(128, 256, (n, d), device=DEV, generator=g) + 1).float()          # odd integers in [257, 511]
    X *= (torch.randint(0, 2, (n, d), device=DEV, generator=g) * 2 - 1).float()
    Q = X[torch.randint(0, n, (nq,), device=DEV, generator=g)].clone()
    Q[nq // 2:] += torch.randint(-2, 3, (nq - nq // 2, d), device=DEV, generator=g).float() * 2
    mir, stats = _prep(X, metric)
    A, B = stats.tolist()
    if metric == "l2":
        rn = X.norm(dim=1).max().item()
        assert abs(B - d ** 0.5) < 0.01 * d ** 0.5        # |x~ - x| = 1 per element on this grid (ulp = 2 in [256, 512))
        assert 0.99 * rn * (1 - 2.0 ** -8) <= A <= 1.01 * rn * (1 + 2.0 ** -8)
    D, I, n_bad = ops.exact_topk_tc(Q, X, mir, stats, 10, metric, return_uncertified=True)
    Dr, Ir = ops.exact_topk_f32(Q, X, 10, metric)
    assert torch.equal(I, Ir) and torch.equal(D, Dr), f"{(I != Ir).sum().item()} ids differ"
    Dref, Iref = (O.exact_l2_topk if metric == "l2" else O.exact_cosine_topk)(
        X.cpu().numpy(), np.arange(n, dtype=np.int64), Q[:8].cpu().numpy(), 10)
    np.testing.assert_array_equal(I[:8].cpu().numpy(), Iref)


def test_tc_cosine_vs_oracle():
    """The tensor-core cosine path against the oracle itself (not only against the fp32 kernel)."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(15)
    X = rng.standard_normal((16_000, 384)).astype(np.float32)
    Q = rng.standard_normal((48, 384)).astype(np.float32)
    ids = np.arange(16_000, dtype=np.int64) * 3 + 1
    Xd, Qd = torch.from_numpy(X).to(DEV), torch.from_numpy(Q).to(DEV)
    mir, stats = _prep(Xd, "cosine")
    S, I = ops.exact_topk_tc(Qd, Xd, mir, stats, 10, "cosine", ids=torch.from_numpy(ids).to(DEV))
    Sref, Iref = O.exact_cosine_topk(X, ids, Q, 10)
    check_topk(S.cpu().numpy(), I.cpu().numpy(), Sref, Iref, rtol=1e-5, atol=2e-6, max_tied_rows=1)


def test_exact_index_uses_tensor_cores_when_it_can():
    from vector_db_b200 import _lib
    from vector_db_b200.indexing import ExactIndex
    rng = np.random.default_rng(3)
    X = rng.standard_normal((30_000, 128)).astype(np.float32)
    idx = ExactIndex(128, "l2")
    idx.add(range(30_000), X)
    _lib.profile_enable(True)
    _lib.profile_read("exact_tc")
    D, I = idx.search(X[:50], 10)
    n, ms = _lib.profile_read("exact_tc")
    _lib.profile_enable(False)
    assert n == 1 and I[:, 0].tolist() == list(range(50)) and np.all(D[:, 0] == 0)
    Dref, Iref = O.exact_l2_topk(X, np.arange(30_000), X[:50], 10)
    np.testing.assert_array_equal(I, Iref)


@pytest.mark.gpu
def test_cta_pair_kernel_returns_the_single_cta_kernels_results():
    """exact_tc2_kernel (tcgen05 cta_group::2, selected by debug bit 5) is kept although it measured slower: ids and
    distances must equal the default kernel's, for an odd number of query tiles, a ragged last row tile and k = 1 / 10."""
    import torch
    from vector_db_b200 import _lib, ops
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(3)
    X = torch.randn((70_001, 192), device="cuda", generator=g)
    mir, _, stats = ops.tc_prepare_rows(X, "cosine")
    try:
        for nq, k in ((700, 10), (129, 1), (4096, 10)):
            q = torch.randn((nq, 192), device="cuda", generator=g)
            lib.vdb_debug_set_tc_flags(0)
            d0, i0 = ops.exact_topk_tc(q, X, mir, stats, k, "cosine")
            lib.vdb_debug_set_tc_flags(32)
            d1, i1 = ops.exact_topk_tc(q, X, mir, stats, k, "cosine")
            assert torch.equal(i0, i1) and torch.equal(d0, d1)
    finally:
        lib.vdb_debug_set_tc_flags(0)
