"""Pins the CPU oracle (oracle/*.py) against fixtures produced by the reference itself
(oracle/gen_golden.py) and against the reference tests' known answers."""
import os

import numpy as np

from oracle import oracle as O
from oracle.hnsw_oracle import HnswOracle, search_arrays
from tests._util import check_topk


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_l2_distance_and_exact_topk_match_reference(golden_dir):
    g = _load(golden_dir, "exact_l2.npz")
    X, ids, Q = g["X"], g["ids"], g["Q"]
    for qi in range(Q.shape[0]):
        for i in (0, 5, 17, 299):
            assert O.l2_distance(Q[qi], X[i]) == g["full"][qi, i]
    D, I = O.exact_l2_topk(X, ids, Q, 10)
    np.testing.assert_array_equal(I, g["I"])
    np.testing.assert_allclose(D, g["D"], rtol=1e-6, atol=0)
    assert D[0, 0] == 0.0 and I[0, 0] == ids[5] and I[0, 1] == ids[17]   # duplicate row: smaller id first
    for qi in range(3):
        pw = O.exact_l2_topk_pairwise(X, ids, Q[qi], 10)
        assert [p[1] for p in pw] == list(g["I"][qi])
    Dg, Ig = O.exact_l2_topk_gemm(X, ids, Q, 10)
    np.testing.assert_array_equal(Ig, g["I"])


def test_exact_topk_short_database_pads():
    X = np.eye(3, 4, dtype=np.float32)
    D, I = O.exact_l2_topk(X, np.array([4, 2, 9]), X[:1], 5)
    assert list(I[0]) == [4, 2, 9, -1, -1] and D[0, 0] == 0 and np.isinf(D[0, 3])


def test_topk_largest_sorted_matches_reference(golden_dir):
    g = _load(golden_dir, "exact_l2.npz")
    np.testing.assert_array_equal(O.topk_largest_sorted(g["scores"], 7), g["top7"])


def test_cosine_topk_is_largest_first():
    rng = np.random.default_rng(0)
    X = rng.standard_normal((200, 8)).astype(np.float32)
    S, I = O.exact_cosine_topk(X, np.arange(200), X[:4] * 3.0, 5)
    assert list(I[:, 0]) == [0, 1, 2, 3]
    assert np.all(np.diff(S, axis=1) <= 0) and np.allclose(S[:, 0], 1.0, atol=1e-6)


def test_cosine_order_pinned_to_reference_arithmetic(golden_dir):
    """cosine.npz (oracle/gen_golden.py:gen_cosine): the reference's euclidean_vector_distance + (dist, id) sort on
    L2-normalised vectors, and top_k_indices_sorted on their dot products.  The oracle's cosine top-k must give
    the same ids except inside fp32 value ties, sims equal to the reference's dot products, and
    |q^ - x^|^2 == 2 - 2 sim."""
    g = _load(golden_dir, "cosine.npz")
    S, I = O.exact_cosine_topk(g["X"], g["ids"], g["Q"], 10)
    check_topk(S, I, g["S_dot"], g["I_dot"], rtol=1e-5, atol=2e-6)
    # unit-vector L2 order == cosine order: compare through the monotone map sim -> sqrt(2 - 2 sim)
    d_from_s = np.sqrt(np.maximum(0.0, 2.0 - 2.0 * S.astype(np.float64))).astype(np.float32)
    check_topk(d_from_s, I, g["D_unit_l2"], g["I_unit_l2"], rtol=2e-4, atol=1e-3)
    assert I[0, 0] in (g["ids"][9], g["ids"][33]) and I[0, 1] in (g["ids"][9], g["ids"][33])   # same direction twice
    assert abs(S[0, 0] - 1.0) < 1e-6


def test_adc_pinned_to_reference_decode_distance(golden_dir):
    """adc.npz (oracle/gen_golden.py:gen_adc): euclidean_vector_distance(q, decode(code)) from the reference on
    the reference's own centroids / codes.  The ADC restatement (fixed-point table) must give that order and the
    squared distances to fp32 rounding."""
    a = _load(golden_dir, "adc.npz")
    for tag in ("pq", "pq8"):
        g = _load(golden_dir, f"{tag}.npz")
        cents = list(g["centroids"])
        lut = O.adc_lut(a[f"{tag}_Q"], cents)
        D, I = O.adc_scan_topk(g["codes"], lut, 10)
        ref_sq = (a[f"{tag}_D"].astype(np.float32) ** 2).astype(np.float32)
        check_topk(D, I, ref_sq, a[f"{tag}_I"], rtol=1e-5, atol=1e-5)
        assert I[0, 0] == a[f"{tag}_I"][0, 0] and D[0, 0] < 1e-6       # query on a decoded point


def test_pq_kats_and_codes_match_reference(golden_dir):
    g = _load(golden_dir, "pq.npz")
    # tests/inference/test_pq.py:78-90 and :92-116
    c = np.array([[0, 0], [1, 1], [2, 2]], dtype=np.float32)
    assert O.pq_nearest_centroid(np.array([0.9, 1.1], np.float32), c) == 1 == int(g["kat1"])
    assert list(O.pq_encode_loop(np.array([[0.9, 1.1, 2.1, 1.9]], np.float32), [c, c])[0]) == [1, 2]
    assert list(g["kat2"]) == [1, 2]
    cents = list(g["centroids"])
    np.testing.assert_array_equal(O.pq_chunk(g["X"][:5], 8), g["chunk_kat"])
    np.testing.assert_array_equal(O.pq_encode(g["Y"], cents), g["codes_y"])
    np.testing.assert_array_equal(O.pq_encode_loop(g["Y"][:50], cents), g["codes_y"][:50])
    # compress() labels the training rows with the FINAL centroids (pq.py:103-106)
    np.testing.assert_array_equal(O.pq_encode(g["X"], cents), g["codes"])


def test_adc_restatement_consistent_with_decode(golden_dir):
    g = _load(golden_dir, "pq.npz")
    cents = list(g["centroids"])
    Q = g["Y"][:6]
    codes = g["codes"]
    lut = O.adc_lut(Q, cents)
    D, I = O.adc_scan_topk(codes, lut, 8)
    recon = np.concatenate([cents[j][codes[:, j]] for j in range(len(cents))], axis=1)
    ref = ((Q[:, None, :] - recon[None, :, :]) ** 2).sum(axis=2)
    np.testing.assert_allclose(D, np.sort(ref, axis=1)[:, :8], rtol=2e-5)
    assert np.all(np.diff(D, axis=1) >= 0)


def _rebuild(g):
    M, efc = int(g["M"]), int(g["efc"])
    X = g["X"]
    o = HnswOracle(M, efc)
    for i in range(X.shape[0]):
        o.insert(i, X[i])
    return o


def _graph_rows(o):
    rows = []
    for nid, d in o.links.items():
        for lvl, lst in d.items():
            for pos, nb in enumerate(lst):
                rows.append((nid, lvl, pos, nb))
    return np.array(sorted(rows), dtype=np.int64)


def test_adc_lower_bound_tables_never_drop_a_row_under_the_threshold():
    """The collect pass of the ADC scan tests rows against 15-bit lower-bound tables (eight queries per shared-memory
    lookup) and re-evaluates the rows that pass.  What makes that exact: a row with d_int <= tau always passes, and no
    sum of table entries reaches 2^15 (two 16-bit fields share a word; bit 15 carries the packed compare).  Checked
    on the restatement of the table construction for plain, heavy-tailed and constant tables, thresholds from the
    5e-4 quantile up to "collect everything"."""
    rng = np.random.default_rng(99)
    n, m = 20000, 24
    codes = rng.integers(0, 256, (n, m))
    for kind in ("plain", "tail", "const", "wide"):
        lut = rng.random((6, m, 256)).astype(np.float32) * np.float32(3.0)
        if kind == "tail":
            lut[rng.random(lut.shape) < 0.01] *= np.float32(5000.0)
        elif kind == "const":
            lut[:] = np.float32(0.75)
        lutq, _ = O.adc_quantize_lut(lut)
        d = np.zeros((6, n), dtype=np.int64)
        for j in range(m):
            d += lutq[:, j, codes[:, j]]
        tau = np.sort(d, axis=1)[:, 10] if kind != "wide" else np.full(6, 0xFFFFFFFF, dtype=np.int64)
        tau[0] = d[0].min()                                   # the tightest useful threshold
        lut16, thr, off, shift = O.adc_coarse_tables(lutq, tau)
        assert lut16.min() >= 0 and int(lut16.max(axis=2).sum(axis=1).max()) <= 32767
        cs = np.zeros((6, n), dtype=np.int64)
        for j in range(m):
            cs += lut16[:, j, codes[:, j]]
        assert np.all((cs << shift[:, None]) + off[:, None] <= d)              # a lower bound of every row's distance
        passed = cs <= thr[:, None]
        assert np.all(passed[d <= tau[:, None]])                                # no row under the threshold is dropped
        if kind == "plain":                                                     # and the filter is worth having
            assert passed.sum(axis=1).max() < 40


def test_hnsw_oracle_reproduces_reference_graph_and_search(golden_dir):
    for name in ("hnsw_small.npz", "hnsw_medium.npz"):
        g = _load(golden_dir, name)
        o = _rebuild(g)
        assert o.entry == int(g["entry"]) and o.max_level == int(g["max_level"])
        assert o.n_evals == int(g["build_evals"])
        np.testing.assert_array_equal(_graph_rows(o), np.array(sorted(map(tuple, g["graph"])), dtype=np.int64))
        k, ef = int(g["k"]), int(g["ef"])
        allowed = {i for i in range(g["X"].shape[0]) if i % 3 == 0}
        arr = o.export_arrays()
        for qi in range(g["Q"].shape[0]):
            o.n_evals = 0
            res = o.search(g["Q"][qi], k, ef)
            assert o.n_evals == int(g["evals"][qi])
            assert [r[1] for r in res] == [int(x) for x in g["I"][qi] if x >= 0]
            np.testing.assert_array_equal(np.array([r[0] for r in res], np.float32), g["D"][qi][:len(res)])
            resf = o.search(g["Q"][qi], k, ef, allowed)
            assert [r[1] for r in resf] == [int(x) for x in g["If"][qi] if x >= 0]
            ra, ev = search_arrays(g["X"], arr["adj0"], arr["up_ptr"], arr["adj_up"], int(arr["entry_row"]),
                                   int(arr["max_level"]), g["Q"][qi], k, ef)
            assert [int(arr["ids"][r[1]]) for r in ra] == [r[1] for r in res] and ev == int(g["evals"][qi])
        assert g["D"][0, 0] < 1e-6 and g["I"][0, 0] == 3     # self query, test_hnsw.py:103-111


def test_pq_dsub8_and_dsub12_codes_match_reference(golden_dir):
    g = _load(golden_dir, "pq8.npz")
    np.testing.assert_array_equal(O.pq_encode(g["Y"], list(g["centroids"])), g["codes_y"])
    np.testing.assert_array_equal(O.pq_encode(g["X"], list(g["centroids"])), g["codes"])
    np.testing.assert_array_equal(O.pq_encode(g["X12"], list(g["centroids12"])), g["codes12"])


def test_ivf_oracle_matches_reference(golden_dir):
    """oracle/ivf_oracle.py against the reference IvfIndex run (oracle/gen_golden.py:gen_ivf): k-means given
    scipy's initial rows, list construction, add(), search at n_probe 1/3/k, delete() + search."""
    from oracle import ivf_oracle as IO
    g = _load(golden_dir, "ivf.npz")
    X, ids, Q, k = g["X"], g["ids"], g["Q"], int(g["k"])
    cent, ptr = g["centroids"], g["list_ptr"]
    lists = [[int(i) for i in g["list_ids"][ptr[c]:ptr[c + 1]]] for c in range(k)]
    C, labels = O.kmeans_lloyd(X, X[g["init_rows"]], 100)                 # ivf.py:75
    np.testing.assert_allclose(C, cent, rtol=1e-5, atol=1e-6)
    assert IO.lists_from_labels(labels, [int(i) for i in ids], k) == lists   # ivf.py:78-81
    vec = {int(i): X[j] for j, i in enumerate(ids)}
    for j in range(g["A"].shape[0]):                                       # ivf.py:87-117
        c = IO.nearest_cluster(cent, g["A"][j])
        assert c == int(g["add_cluster"][j])
        lists[c].append(int(g["add_ids"][j]))
        vec[int(g["add_ids"][j])] = g["A"][j]
    np.testing.assert_array_equal(np.stack([IO.centroid_distances(cent, q) for q in Q]), g["cdist"])
    for n_probe in (1, 3, k):
        for qi in range(Q.shape[0]):
            res, _, _ = IO.search(cent, lists, vec, Q[qi], n_probe, 10)
            assert [r[1] for r in res] == [int(i) for i in g[f"I{n_probe}"][qi] if i >= 0]
            np.testing.assert_array_equal(np.array([r[0] for r in res], np.float32), g[f"D{n_probe}"][qi][:len(res)])
    for i in g["dead"]:                                                    # ivf.py:119-137
        for lst in lists:
            if int(i) in lst:
                lst.remove(int(i))
        vec.pop(int(i), None)
    for qi in range(Q.shape[0]):
        res, _, _ = IO.search(cent, lists, vec, Q[qi], 3, 10)
        assert [r[1] for r in res] == [int(i) for i in g["Id"][qi] if i >= 0]
