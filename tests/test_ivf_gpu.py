"""GPU parity: IVF probe selection, list scan, training, and the reference's IvfIndex contract
(src/vector_db/indexing/ivf.py, tests/indexing/test_ivf.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import ivf_oracle as IO
from oracle import oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _golden_index(golden_dir, with_adds=True):
    from vector_db_b200.indexing import IvfIndex
    from vector_db_b200.types import Node
    g = np.load(os.path.join(golden_dir, "ivf.npz"))
    k, ptr = int(g["k"]), g["list_ptr"]
    lists = [[int(i) for i in g["list_ids"][ptr[c]:ptr[c + 1]]] for c in range(k)]
    vec = {int(i): g["X"][j] for j, i in enumerate(g["ids"])}
    idx = IvfIndex(k=k)
    for i, v in vec.items():
        idx.storage.save(Node(id=i, embedding=v))
    idx.load_state({"k": k, "centroids": g["centroids"], "inverted_lists": lists}, vectors=vec)
    if with_adds:
        for j in range(g["A"].shape[0]):
            idx.add(Node(id=int(g["add_ids"][j]), embedding=g["A"][j]))
    return g, idx


@pytest.mark.parametrize("K,D", [(12, 24), (50, 7), (33, 130), (64, 384), (10, 777), (5, 1024)])
def test_centroid_distances_bit_equal_numpy(K, D):
    """np.linalg.norm(centroids - q, axis=1) (ivf.py:109,170): numpy's pairwise summation order reproduced."""
    from vector_db_b200 import ops
    rng = np.random.default_rng(K * 1000 + D)
    cent = rng.standard_normal((K, D)).astype(np.float32)
    Q = rng.standard_normal((9, D)).astype(np.float32)
    got = ops.ivf_centroid_dist(_t(Q), _t(cent)).cpu().numpy()
    ref = np.stack([IO.centroid_distances(cent, q) for q in Q])
    np.testing.assert_array_equal(got, ref)
    probes, _ = ops.ivf_probe(_t(ref), min(K, 4))
    np.testing.assert_array_equal(probes.cpu().numpy(), np.stack([np.argsort(r, kind="stable")[:min(K, 4)] for r in ref]))


def test_search_add_delete_match_reference_run(golden_dir):
    g, idx = _golden_index(golden_dir)
    k = int(g["k"])
    for j in range(g["A"].shape[0]):                                   # add(): nearest cluster, appended last
        assert int(g["add_ids"][j]) in idx.inverted_lists[int(g["add_cluster"][j])]
    for n_probe in (1, 3, k):
        D, I = idx.search_batch(g["Q"], n_probe, 10)
        np.testing.assert_array_equal(I, g[f"I{n_probe}"])
        np.testing.assert_allclose(D, g[f"D{n_probe}"], rtol=1e-6, atol=1e-7)
    res = idx.search(g["Q"][0], n_probe=3, top_k=5)                    # duplicate embeddings: list order breaks the tie
    assert [n.id for n, _ in res[:2]] == [int(g["ids"][7]), int(g["ids"][40])] and res[0][1] < 1e-6
    assert isinstance(res[0][1], np.float32) and np.array_equal(res[0][0].embedding, g["X"][7])
    for i in g["dead"]:
        idx.delete(int(i))
    D, I = idx.search_batch(g["Q"], 3, 10)
    np.testing.assert_array_equal(I, g["Id"])
    np.testing.assert_allclose(D, g["Dd"], rtol=1e-6, atol=1e-7)
    assert idx.get_cluster_stats()["total_vectors"] == 1500 + 30 - len(set(int(i) for i in g["dead"]))


def test_training_matches_reference_kmeans_given_its_initial_rows(golden_dir):
    """kmeans2(iter=100, minit='points') (ivf.py:75) from the rows scipy drew: same lists, same centroids."""
    from vector_db_b200.indexing import IvfIndex
    g = np.load(os.path.join(golden_dir, "ivf.npz"))
    k, ptr = int(g["k"]), g["list_ptr"]
    idx = IvfIndex(k=k)
    idx.build_from_arrays([int(i) for i in g["ids"]], g["X"], init_rows=g["init_rows"])
    np.testing.assert_allclose(idx.centroids, g["centroids"], rtol=1e-4, atol=1e-5)
    assert idx.inverted_lists == [[int(i) for i in g["list_ids"][ptr[c]:ptr[c + 1]]] for c in range(k)]
    # the same draw through the global numpy RNG, as scipy makes it
    np.random.seed(5)
    idx2 = IvfIndex(k=k)
    idx2.build_from_arrays([int(i) for i in g["ids"]], g["X"])
    assert idx2.inverted_lists == idx.inverted_lists


@pytest.mark.parametrize("n,dim,K,n_probe,nq,top_k", [(6000, 48, 32, 5, 64, 10), (3000, 30, 16, 16, 17, 25),
                                                     (20000, 128, 100, 8, 300, 10), (500, 8, 7, 2, 5, 600)])
def test_batched_search_vs_oracle(n, dim, K, n_probe, nq, top_k):
    from vector_db_b200.indexing import IvfIndex
    rng = np.random.default_rng(n + K)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    ids = np.arange(n, dtype=np.int64) * 3 + 1
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    idx = IvfIndex(k=K)
    idx.build_from_arrays([int(i) for i in ids], X, iters=5, init_rows=rng.choice(n, K, replace=False))
    D, I = idx.search_batch(Q, n_probe, top_k)
    vec = {int(i): X[j] for j, i in enumerate(ids)}
    lists = idx.inverted_lists
    assert sum(len(l) for l in lists) == n
    for qi in range(0, nq, max(1, nq // 16)):
        res, n_cand, _ = IO.search(idx.centroids, lists, vec, Q[qi], n_probe, top_k)
        assert int(idx.last_scanned[qi]) == n_cand
        assert [r[1] for r in res] == [int(i) for i in I[qi] if i >= 0]
        np.testing.assert_allclose(D[qi][:len(res)], np.array([r[0] for r in res], np.float32), rtol=1e-5)
        assert np.all(I[qi][len(res):] == -1) and np.all(np.isinf(D[qi][len(res):]))


def test_full_probe_equals_exact_scan():
    """n_probe == k visits every list: the values are the exact fp32 scan's, bit for bit (property at a size the
    oracle does not reach)."""
    from vector_db_b200 import ops
    from vector_db_b200.indexing import IvfIndex
    rng = np.random.default_rng(99)
    n, dim, K = 100_000, 64, 64
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((128, dim)).astype(np.float32)
    idx = IvfIndex(k=K)
    idx.build_from_arrays(range(n), X, iters=3, init_rows=np.arange(K))
    D, I = idx.search_batch(Q, K, 10)
    De, Ie = ops.exact_topk_f32(_t(Q), _t(X), 10, "l2")
    np.testing.assert_array_equal(D, De.cpu().numpy())
    np.testing.assert_array_equal(I, Ie.cpu().numpy())
    assert np.all(idx.last_scanned == n)


def test_reference_contract_errors_stats_persistence(tmp_path):
    """tests/indexing/test_ivf.py:17-23,47-64,96-102,169-195,231-313 restated against the GPU class."""
    from vector_db_b200.indexing import IvfIndex
    from vector_db_b200.types import Node
    with pytest.raises(ValueError, match="k-means parameter should be positive"):
        IvfIndex(k=0)
    ivf = IvfIndex(k=4)
    assert ivf.k == 4 and ivf.centroids is None and ivf.inverted_lists == []
    with pytest.raises(ValueError, match="Cannot build index with empty node list"):
        ivf.build_index([])
    with pytest.raises(ValueError, match="Need at least 4 vectors for 4 clusters"):
        ivf.build_index([Node(id=i, embedding=np.random.rand(8).astype(np.float32)) for i in range(2)])
    with pytest.raises(ValueError, match="Index must be built before adding nodes"):
        IvfIndex(k=4).add(Node(id=1, embedding=np.zeros(8, np.float32)))
    with pytest.raises(ValueError, match="Index must be built before searching"):
        IvfIndex(k=4).search(np.zeros(8, np.float32), n_probe=2, top_k=5)
    index_file = tmp_path / "ivf.pkl"
    np.random.seed(3)
    ivf = IvfIndex(k=4, index_file=index_file)
    nodes = [Node(id=i, embedding=np.random.rand(8).astype(np.float32), metadata={"i": i}) for i in range(40)]
    ivf.build_index(nodes)
    assert ivf.centroids.shape == (4, 8) and len(ivf.inverted_lists) == 4
    assert sum(len(l) for l in ivf.inverted_lists) == 40
    for bad in (0, -1, 5):
        with pytest.raises(ValueError, match="n_probe must be between 1 and 4"):
            ivf.search(nodes[0].embedding, n_probe=bad, top_k=5)
    with pytest.raises(ValueError, match="query must be 1D array"):
        ivf.search(np.zeros((2, 8), np.float32), n_probe=2, top_k=5)
    with pytest.raises(ValueError, match="doesn't match"):
        ivf.search(np.zeros(9, np.float32), n_probe=2, top_k=5)
    first, dist = ivf.search(nodes[0].embedding, n_probe=4, top_k=5)[0]
    assert first.id == 0 and dist < 1e-6 and first.metadata == {"i": 0}
    stats = ivf.get_cluster_stats()
    assert stats["total_vectors"] == 40 and stats["min_size"] <= stats["avg_size"] <= stats["max_size"]
    assert ivf.get_cluster_size(0) == len(ivf.inverted_lists[0])
    with pytest.raises(ValueError, match="cluster_id must be between 0 and 3"):
        ivf.get_cluster_size(4)
    # persistence: a second instance on the same file + storage answers the same (test_ivf.py:277-313)
    assert index_file.exists()
    ivf2 = IvfIndex(k=4, storage=ivf.storage, index_file=index_file)
    np.testing.assert_array_equal(ivf2.centroids, ivf.centroids)
    assert ivf2.inverted_lists == ivf.inverted_lists
    a = [(n.id, d) for n, d in ivf.search(nodes[3].embedding, n_probe=2, top_k=5)]
    b = [(n.id, d) for n, d in ivf2.search(nodes[3].embedding, n_probe=2, top_k=5)]
    assert a == b
    ivf.delete(0)
    assert all(0 not in l for l in ivf.inverted_lists)
    assert all(n.id != 0 for n, _ in ivf.search(nodes[0].embedding, n_probe=4, top_k=5))


@pytest.mark.parametrize("n,dim,K,m,n_probe,nq,R", [(6000, 64, 24, 8, 4, 40, 50), (4000, 48, 10, 16, 10, 9, 300),
                                                   (3000, 40, 8, 5, 3, 7, 20),
                                                   (40000, 32, 6, 4, 6, 5, 1000),   # several radix compactions per CTA
                                                   (3000, 128, 6, 64, 3, 6, 40),    # m = 64: one LUT copy, 256 threads
                                                   (300, 32, 8, 4, 2, 5, 200)])     # fewer candidates than the shortlist
def test_ivf_pq_scan_vs_oracle(n, dim, K, m, n_probe, nq, R):
    """IVF-PQ shortlist = fixed-point ADC over the probed lists in enumeration order; rerank = exact L2."""
    from vector_db_b200 import ops
    from vector_db_b200.indexing import IvfPqIndex
    rng = np.random.default_rng(n + m)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    ids = np.arange(n, dtype=np.int64) * 2 + 5
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    idx = IvfPqIndex(k=K, chunks=m, pq_k=64 if m == 5 else 256)
    idx.build_from_arrays([int(i) for i in ids], X, iters=4, init_rows=rng.choice(n, K, replace=False))
    idx.train_pq(iters=3, train_rows=2000)
    cb = idx.codebooks.cpu().numpy()
    codes_all = O.pq_encode(X, list(cb))
    codes = {int(i): codes_all[j] for j, i in enumerate(ids)}
    lut = O.adc_lut(Q, list(idx._cb_lut.cpu().numpy()))
    D, I = idx.search_batch(Q, n_probe, R, rerank=False)
    lists = idx.inverted_lists
    for qi in range(nq):
        ref, n_cand = IO.ivf_adc_search(idx.centroids, lists, codes, lut[qi], Q[qi], n_probe, R)
        assert int(idx.last_scanned[qi]) == n_cand
        assert [r[1] for r in ref] == [int(i) for i in I[qi] if i >= 0]
        np.testing.assert_array_equal(D[qi][:len(ref)], np.array([r[0] for r in ref], np.float32))
    # rerank over the full candidate set == plain IVF search
    Dr, Ir = idx.search_batch(Q, n_probe, 10, shortlist=min(4096, n), rerank=True)
    De, Ie = IvfIndex_search(idx, Q, n_probe, 10)
    full = idx.last_scanned <= min(4096, n)
    np.testing.assert_array_equal(Ir[full], Ie[full])
    np.testing.assert_array_equal(Dr[full], De[full])


def IvfIndex_search(idx, Q, n_probe, k):
    """Plain IVF search (exact distances over all probed members) on an IvfPqIndex."""
    from vector_db_b200.indexing import IvfIndex
    d, i, _ = IvfIndex.search_device(idx, _t(Q), n_probe, k)
    return d.cpu().numpy(), i.cpu().numpy()


def test_ivf_filter_ids_restricts_results(golden_dir):
    """filter_ids (the metadata pre-filter of services/storage_service.py:106-128 -> hnsw.py:96,127, offered for
    IVF as an extension): results equal the oracle run on the allowed nodes only."""
    g, idx = _golden_index(golden_dir, with_adds=False)
    k, ptr = int(g["k"]), g["list_ptr"]
    lists = [[int(i) for i in g["list_ids"][ptr[c]:ptr[c + 1]]] for c in range(k)]
    allowed = set(int(i) for i in g["ids"][::4])
    vec = {int(i): g["X"][j] for j, i in enumerate(g["ids"]) if int(i) in allowed}
    D, I = idx.search_batch(g["Q"], 3, 10, filter_ids=allowed)
    for qi in range(g["Q"].shape[0]):
        res, _, _ = IO.search(g["centroids"], lists, vec, g["Q"][qi], 3, 10)
        assert [r[1] for r in res] == [int(i) for i in I[qi] if i >= 0]


def test_list_major_tensor_core_batch_search_matches_reference_run(golden_dir):
    """The list-major tensor-core batch path (vdb_ivf_topk_tc, forced with batch_engine = "tc") on the reference-run
    fixture: ids and distances of the reference's own IvfIndex.search at n_probe 1 / 3 / k, duplicate embeddings in list
    order, and after delete()s -- lists far smaller than a 256-row tile, so padding rows and pair-tile padding dominate."""
    g, idx = _golden_index(golden_dir)
    idx.batch_engine = "tc"
    k = int(g["k"])
    for n_probe in (1, 3, k):
        D, I = idx.search_batch(g["Q"], n_probe, 10)
        np.testing.assert_array_equal(I, g[f"I{n_probe}"])
        np.testing.assert_allclose(D, g[f"D{n_probe}"], rtol=1e-6, atol=1e-7)
    D, I = idx.search_batch(g["Q"][:1], 3, 5)
    assert I[0, :2].tolist() == [int(g["ids"][7]), int(g["ids"][40])] and D[0, 0] < 1e-6
    for i in g["dead"]:
        idx.delete(int(i))
    D, I = idx.search_batch(g["Q"], 3, 10)
    np.testing.assert_array_equal(I, g["Id"])
    np.testing.assert_allclose(D, g["Dd"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("n,dim,K,nq,n_probe,top_k,max_bad", [
    (120_000, 64, 256, 700, 4, 10, 0.05), (120_000, 64, 256, 700, 16, 5, 0.05), (60_000, 128, 40, 300, 32, 10, None),
    (50_000, 384, 512, 2100, 1, 10, None), (200_000, 128, 256, 1500, 8, 10, 0.01)])
def test_list_major_tensor_core_batch_search_equals_query_major_scan(n, dim, K, nq, n_probe, top_k, max_bad):
    """Same index, same queries: the tensor-core batch path must return what the query-major fp32 list scan returns --
    ids equal, distances bit-equal (both re-score with the reference's arithmetic), candidates-enumerated counts equal.
    ``max_bad``: largest share of queries the certificate may send back to the query-major scan (None: clustered rows
    under 4 Lloyd rounds leave lists that merge several clusters, whose residuals are too wide for a bf16 coarse pass to
    separate the k-th from the next neighbours -- still exact through the fallback, so only the results are checked);
    the last case is unclustered Gaussian rows, where nothing should come back."""
    from vector_db_b200.indexing import IvfIndex
    rng = np.random.default_rng(n + K + n_probe)
    if max_bad is not None and max_bad < 0.05:
        X = rng.standard_normal((n, dim), dtype=np.float32)
        Q = rng.standard_normal((nq, dim), dtype=np.float32)
    else:
        centres = rng.standard_normal((K, dim)).astype(np.float32) * 3
        X = (centres[rng.integers(0, K, n)] + rng.standard_normal((n, dim))).astype(np.float32)
        Q = (centres[rng.integers(0, K, nq)] + rng.standard_normal((nq, dim))).astype(np.float32)
    X[5000] = X[17]                                                    # duplicate rows in one list: list order breaks the tie
    Q[0] = X[17]
    idx = IvfIndex(k=K)
    idx.build_from_arrays(range(n), X, iters=4, init_rows=rng.choice(n, K, replace=False))
    idx.batch_engine = "scan"
    Ds, Is = idx.search_batch(Q, n_probe, top_k)
    scanned = idx.last_scanned.copy()
    idx.batch_engine = "tc"
    Dt, It = idx.search_batch(Q, n_probe, top_k)
    np.testing.assert_array_equal(It, Is)
    np.testing.assert_array_equal(Dt, Ds)
    np.testing.assert_array_equal(idx.last_scanned, scanned)
    assert max_bad is None or idx.last_uncertified <= int(nq * max_bad)
    assert Is[0, 0] == 17 and Ds[0, 0] == 0.0


@pytest.mark.gpu
def test_probe_selection_through_the_tensor_cores_equals_the_numpy_order_kernels():
    """ops.ivf_probe_tc (n_probe + 4 nearest centroids from the tensor-core scan, re-scored in numpy's summation order, flagged
    queries through the full kernels) must return the probes and prefix sums of ivf_probe(ivf_centroid_dist(...)) -- on
    well-separated centroids (nothing flagged) and on centroids so close to one another that the certificate gives up."""
    import torch
    from vector_db_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    for K, D, spread in ((1024, 384, 0.3), (300, 64, 1.0), (2048, 128, 1e-3)):
        cent = torch.randn((K, D), device="cuda", generator=g) * spread
        cent[7] = cent[3]                                           # duplicate centroid: the lower index must come first
        q = torch.randn((2000, D), device="cuda", generator=g)
        q[:50] = cent[torch.arange(50, device="cuda") % K] + 1e-4 * torch.randn((50, D), device="cuda", generator=g)
        ptr = (torch.arange(K + 1, device="cuda", dtype=torch.int32) * 37).contiguous()
        mirror, _, stats = ops.tc_prepare_rows(cent, "l2")
        for n_probe in (1, 4, 8):
            p0, x0 = ops.ivf_probe(ops.ivf_centroid_dist(q, cent), n_probe, ptr)
            p1, x1, flags, _ = ops.ivf_probe_tc(q, cent, mirror, stats, n_probe, ptr)
            assert torch.equal(p0, p1) and torch.equal(x0, x1), (K, D, spread, n_probe, int(flags.sum()))
