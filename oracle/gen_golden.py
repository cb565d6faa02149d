"""Generate tests/golden/*.npz by running the REFERENCE ITSELF (imported read-only from
/root/reference/src) on small seeded inputs.  Run in the build container only:

    python oracle/gen_golden.py

The GPU box has no /root/reference; tests read the committed .npz files.  TEST INFRASTRUCTURE.
"""
from __future__ import annotations

import os
import random
import sys

import numpy as np

REF_SRC = "/root/reference/src"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _import_reference():
    sys.path.insert(0, REF_SRC)
    import vector_db.indexing.hnsw as ref_hnsw
    import vector_db.inference.pq as ref_pq
    from vector_db.storage import InMemoryNodeStorage
    from vector_db.types import Node
    from vector_db.util.distance import euclidean_vector_distance
    from vector_db.util.math import top_k_indices_sorted
    return ref_hnsw, ref_pq, InMemoryNodeStorage, Node, euclidean_vector_distance, top_k_indices_sorted


def gen_exact(ref_dist, topk_sorted):
    rng = np.random.default_rng(1234)
    X = rng.standard_normal((300, 24)).astype(np.float32)
    X[17] = X[5]                      # a duplicate row: (dist,id) tie-break must pick the smaller id
    ids = np.arange(300, dtype=np.int64) * 3 + 7
    Q = rng.standard_normal((12, 24)).astype(np.float32)
    Q[0] = X[5]                       # self query -> distance exactly 0 (test_hnsw.py:103-111)
    k = 10
    D = np.empty((12, k), np.float32)
    I = np.empty((12, k), np.int64)
    full = np.empty((12, 300), np.float32)
    for qi in range(12):
        pairs = [(ref_dist(Q[qi], X[i]), int(ids[i])) for i in range(300)]   # test_hnsw.py:128-134
        full[qi] = [p[0] for p in pairs]
        pairs.sort()
        D[qi] = [p[0] for p in pairs[:k]]
        I[qi] = [p[1] for p in pairs[:k]]
    scores = rng.standard_normal(50).astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "exact_l2.npz"), X=X, ids=ids, Q=Q, D=D, I=I, full=full,
                        scores=scores, top7=topk_sorted(scores, 7))


def gen_cosine(ref_dist, topk_sorted):
    """Pins the cosine order to the reference's OWN arithmetic (the reference has no cosine function, VERDICT r1
    next #3): rows and queries are L2-normalised in fp32, then ranked (a) by the reference's
    euclidean_vector_distance + (dist, id) tuple sort (test_hnsw.py:128-134) -- |q^ - x^|^2 = 2 - 2 cos, so this
    order IS the cosine order up to fp32 ties -- and (b) by top_k_indices_sorted (util/math.py:3-5) applied to
    the dot products q^ . x^."""
    rng = np.random.default_rng(2468)
    n, dim, nq, k = 600, 48, 16, 10
    X = (rng.standard_normal((n, dim)) * rng.uniform(0.2, 4.0, (n, 1))).astype(np.float32)
    X[33] = 2.0 * X[9]                # same direction, different length: a cosine tie, distinct L2
    ids = np.arange(n, dtype=np.int64) * 2 + 5
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    Q[0] = 3.0 * X[9]
    Xn = (X / np.linalg.norm(X, axis=1, keepdims=True)).astype(np.float32)
    Qn = (Q / np.linalg.norm(Q, axis=1, keepdims=True)).astype(np.float32)
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    S = np.empty((nq, k), np.float32)
    Is = np.empty((nq, k), np.int64)
    for qi in range(nq):
        pairs = [(ref_dist(Qn[qi], Xn[i]), int(ids[i])) for i in range(n)]
        pairs.sort()
        D[qi] = [p_[0] for p_ in pairs[:k]]
        I[qi] = [p_[1] for p_ in pairs[:k]]
        sims = np.array([np.dot(Qn[qi], Xn[i]) for i in range(n)], dtype=np.float32)
        top = topk_sorted(sims, k)
        S[qi] = sims[top]
        Is[qi] = ids[top]
    np.savez_compressed(os.path.join(OUT, "cosine.npz"), X=X, ids=ids, Q=Q, D_unit_l2=D, I_unit_l2=I, S_dot=S, I_dot=Is)


def gen_adc(ref_dist):
    """Pins the ADC scan to the reference's arithmetic: for the centroids / codes the
    reference's compress() produced (pq.npz, pq8.npz), the distance of a query to a row's DECODED vector
    (concatenated centroids, pq.py:48-62 layout) by euclidean_vector_distance, squared, and its (dist, id) top-k.
    sum_j |q_j - C_j[c_j]|^2 == |q - decode(c)|^2, so the ADC order is this order up to fp32 rounding."""
    out = {}
    for tag, seed in (("pq", 31), ("pq8", 32)):
        g = np.load(os.path.join(OUT, f"{tag}.npz"))
        cents, codes = g["centroids"], g["codes"]                  # (m,K,dsub), (N,m)
        m = cents.shape[0]
        dec = np.concatenate([cents[j][codes[:, j]] for j in range(m)], axis=1).astype(np.float32)
        rng = np.random.default_rng(seed)
        nq, k = 12, 10
        Q = rng.standard_normal((nq, dec.shape[1])).astype(np.float32)
        Q[0] = dec[4]                                              # a query ON a decoded point: ADC distance 0
        D = np.empty((nq, k), np.float32)
        I = np.empty((nq, k), np.int64)
        full = np.empty((nq, dec.shape[0]), np.float32)
        for qi in range(nq):
            pairs = [(ref_dist(Q[qi], dec[i]), i) for i in range(dec.shape[0])]
            full[qi] = [p_[0] for p_ in pairs]
            pairs.sort()
            D[qi] = [p_[0] for p_ in pairs[:k]]
            I[qi] = [p_[1] for p_ in pairs[:k]]
        out.update({f"{tag}_Q": Q, f"{tag}_D": D, f"{tag}_I": I, f"{tag}_full": full})
    np.savez_compressed(os.path.join(OUT, "adc.npz"), **out)


def gen_pq(ref_pq):
    # KATs restated from tests/inference/test_pq.py:78-116
    svc = ref_pq.ProductQuantizationService(k=3, chunks=2, dim=4)
    kat_c = np.array([[0.0, 0.0], [1.0, 1.0], [2.0, 2.0]], dtype=np.float32)
    kat1 = svc._find_nearest_centroid(np.array([0.9, 1.1], dtype=np.float32), kat_c)
    cents = [np.array([[0.0, 0.0], [1.0, 1.0], [2.0, 2.0]], dtype=np.float32),
             np.array([[0.0, 0.0], [1.0, 1.0], [2.0, 2.0]], dtype=np.float32)]
    kat2 = svc._compress_embedding(np.array([0.9, 1.1, 2.1, 1.9], dtype=np.float32), cents)

    # full compress(): centroids + codes from ONE reference run (SURVEY.md 8c determinism note)
    np.random.seed(777)
    rng = np.random.default_rng(777)
    X = rng.standard_normal((1500, 32)).astype(np.float32)
    svc = ref_pq.ProductQuantizationService(k=32, chunks=8, dim=32)
    codes = svc.compress(X)
    cents = np.stack([np.asarray(c, dtype=np.float32) for c in svc.centroids])    # (m,K,dsub)
    chunked = svc._chunk_embeddings(X[:5])
    chunk_kat = np.array([[np.asarray(v) for v in ch] for ch in chunked], dtype=np.float32)
    # encode of held-out rows with the SAME centroids through the reference's per-row path
    Y = rng.standard_normal((400, 32)).astype(np.float32)
    codes_y = np.array([svc._compress_embedding(y, svc.centroids) for y in Y], dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "pq.npz"), X=X, codes=codes, centroids=cents, Y=Y,
                        codes_y=codes_y, chunk_kat=chunk_kat, kat1=np.int64(kat1), kat2=kat2)


def gen_pq8(ref_pq):
    """dsub = 8: numpy's pairwise_sum switches to its 8-accumulator tree at n >= 8, the case the CUDA
    encode kernel has to reproduce bit for bit (BASELINE configs use dsub = 8)."""
    np.random.seed(4242)
    rng = np.random.default_rng(4242)
    X = rng.standard_normal((1200, 64)).astype(np.float32)
    svc = ref_pq.ProductQuantizationService(k=64, chunks=8, dim=64)
    codes = svc.compress(X)
    cents = np.stack([np.asarray(c, dtype=np.float32) for c in svc.centroids])
    Y = rng.standard_normal((600, 64)).astype(np.float32)
    codes_y = np.array([svc._compress_embedding(y, svc.centroids) for y in Y], dtype=np.int64)
    # dsub = 12 exercises the tree + remainder loop
    svc12 = ref_pq.ProductQuantizationService(k=16, chunks=2, dim=24)
    X12 = rng.standard_normal((300, 24)).astype(np.float32)
    codes12 = svc12.compress(X12)
    cents12 = np.stack([np.asarray(c, dtype=np.float32) for c in svc12.centroids])
    np.savez_compressed(os.path.join(OUT, "pq8.npz"), X=X, codes=codes, centroids=cents, Y=Y, codes_y=codes_y,
                        X12=X12, codes12=codes12, centroids12=cents12)


def gen_hnsw(ref_hnsw, Storage, Node, tag, n, dim, M, efc, nq, ef, k, seed):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    Q[0] = X[3]
    counter = {"n": 0}
    real = ref_hnsw.euclidean_vector_distance

    def counting(a, b):
        counter["n"] += 1
        return real(a, b)

    ref_hnsw.euclidean_vector_distance = counting
    try:
        idx = ref_hnsw.HNSW(M=M, ef_construction=efc, rng=random.Random(42), storage=Storage())
        for i in range(n):
            idx.insert_node(Node(id=i, embedding=X[i]))
        build_evals = counter["n"]
        # flatten graph: (node, level, pos) -> neighbour
        rows = []
        for nid, node in idx.node_store.items():
            for lvl, lst in node.neighbors.items():
                for pos, nb in enumerate(lst):
                    rows.append((nid, lvl, pos, nb))
        graph = np.array(rows, dtype=np.int64)
        node_levels = np.array([max(idx.node_store[i].neighbors.keys()) for i in range(n)], np.int32)
        D = np.full((nq, k), np.inf, np.float32)
        I = np.full((nq, k), -1, np.int64)
        evals = np.zeros(nq, np.int64)
        for qi in range(nq):
            counter["n"] = 0
            res = idx.search(Q[qi], k=k, ef=ef)
            evals[qi] = counter["n"]
            for j, (node, d) in enumerate(res):
                D[qi, j], I[qi, j] = d, node.id
        # filtered search (hnsw.py:96,127): allow only ids divisible by 3
        allowed = {i for i in range(n) if i % 3 == 0}
        Df = np.full((nq, k), np.inf, np.float32)
        If = np.full((nq, k), -1, np.int64)
        for qi in range(nq):
            res = idx.search(Q[qi], k=k, ef=ef, filter_ids=allowed)
            for j, (node, d) in enumerate(res):
                Df[qi, j], If[qi, j] = d, node.id
    finally:
        ref_hnsw.euclidean_vector_distance = real
    np.savez_compressed(os.path.join(OUT, f"hnsw_{tag}.npz"), X=X, Q=Q, graph=graph,
                        node_levels=node_levels, entry=np.int64(idx.entry_node_id),
                        max_level=np.int64(idx.max_level), M=np.int64(M), efc=np.int64(efc),
                        ef=np.int64(ef), k=np.int64(k), D=D, I=I, evals=evals, Df=Df, If=If,
                        build_evals=np.int64(build_evals))


def gen_hnsw_toplevel(ref_hnsw, Storage, Node):
    """a graph whose ENTRY POINT lacks its top-level keys.  random.Random(7), M=4, 500 nodes:
    the entry is inserted at level 6 while the top was 4, so its `neighbors` has keys 0..4 only (hnsw.py:179-191)
    and max_level = 6 (hnsw.py:215-217).  Records the key sets (empty lists included), searches on that graph, then
    60 more inserts through the same rng (some of them link into the entry's missing levels) and searches again."""
    rng = np.random.default_rng(3)
    n, extra, dim, nq, k, ef = 500, 60, 8, 16, 5, 30
    X = rng.standard_normal((n + extra, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    idx = ref_hnsw.HNSW(M=4, ef_construction=20, rng=random.Random(7), storage=Storage())
    for i in range(n):
        idx.insert_node(Node(id=i, embedding=X[i]))

    def snap():
        rows, keys = [], []
        for nid, node in idx.node_store.items():
            for lvl, lst in node.neighbors.items():
                keys.append((nid, lvl))
                for pos, nb in enumerate(lst):
                    rows.append((nid, lvl, pos, nb))
        return np.array(rows, dtype=np.int64), np.array(keys, dtype=np.int64)

    def run():
        D = np.full((nq, k), np.inf, np.float32)
        I = np.full((nq, k), -1, np.int64)
        for qi in range(nq):
            for j, (node, d) in enumerate(idx.search(Q[qi], k=k, ef=ef)):
                D[qi, j], I[qi, j] = d, node.id
        return D, I

    graph0, keys0 = snap()
    entry0, top0 = idx.entry_node_id, idx.max_level
    assert top0 > max(idx.node_store[entry0].neighbors.keys()), "fixture lost its point: entry has all its keys"
    D0, I0 = run()
    for i in range(n, n + extra):
        idx.insert_node(Node(id=i, embedding=X[i]))
    graph1, keys1 = snap()
    D1, I1 = run()
    np.savez_compressed(os.path.join(OUT, "hnsw_toplevel.npz"), X=X, Q=Q, n=np.int64(n), M=np.int64(4), efc=np.int64(20),
                        ef=np.int64(ef), k=np.int64(k), graph0=graph0, keys0=keys0, entry0=np.int64(entry0),
                        top0=np.int64(top0), D0=D0, I0=I0, graph1=graph1, keys1=keys1, entry1=np.int64(idx.entry_node_id),
                        top1=np.int64(idx.max_level), D1=D1, I1=I1)


def gen_ivf(Storage, Node):
    """Reference IvfIndex (src/vector_db/indexing/ivf.py): build (kmeans2 on the seeded global numpy RNG),
    add, search at several n_probe, delete + search.  Blobs keep every cluster well populated."""
    import vector_db.indexing.ivf as ref_ivf
    rng = np.random.default_rng(2024)
    n, dim, k, nq, top_k = 1500, 24, 12, 20, 10
    centres = rng.standard_normal((k, dim)).astype(np.float32) * 2.0
    X = (centres[rng.integers(0, k, n)] + 0.7 * rng.standard_normal((n, dim))).astype(np.float32)
    X[40] = X[7]                                # duplicate embedding: stable-sort tie (ivf.py:193)
    ids = (np.arange(n, dtype=np.int64) * 2 + 11)
    Q = (centres[rng.integers(0, k, nq)] + 0.9 * rng.standard_normal((nq, dim))).astype(np.float32)
    Q[0] = X[7]
    idx = ref_ivf.IvfIndex(k=k, storage=Storage())
    np.random.seed(5)
    state = np.random.get_state()
    idx.build_index([Node(id=int(ids[i]), embedding=X[i]) for i in range(n)])
    np.random.set_state(state)
    init_rows = np.random.mtrand._rand.choice(n, size=k, replace=False)     # what scipy's _kpoints drew
    centroids = np.asarray(idx.centroids)
    lists0 = [list(l) for l in idx.inverted_lists]
    # add (ivf.py:87-117)
    A = (centres[rng.integers(0, k, 30)] + 0.7 * rng.standard_normal((30, dim))).astype(np.float32)
    add_ids = np.arange(30, dtype=np.int64) + 100000
    add_cluster = np.empty(30, np.int64)
    for j in range(30):
        idx.add(Node(id=int(add_ids[j]), embedding=A[j]))
        add_cluster[j] = next(c for c, l in enumerate(idx.inverted_lists) if l and l[-1] == int(add_ids[j]))
    res = {}
    for n_probe in (1, 3, k):
        D = np.full((nq, top_k), np.inf, np.float32)
        I = np.full((nq, top_k), -1, np.int64)
        for qi in range(nq):
            out = idx.search(Q[qi], n_probe=n_probe, top_k=top_k)
            D[qi, :len(out)] = [d for _, d in out]
            I[qi, :len(out)] = [nd.id for nd, _ in out]
        res[f"D{n_probe}"], res[f"I{n_probe}"] = D, I
    cdist = np.stack([np.linalg.norm(centroids - Q[qi], axis=1) for qi in range(nq)])
    # delete (ivf.py:119-137) then search again
    dead = np.concatenate([res[f"I{k}"][:, 0][:8], ids[100:110]])
    for i in dead:
        idx.delete(int(i))
    Dd = np.full((nq, top_k), np.inf, np.float32)
    Id = np.full((nq, top_k), -1, np.int64)
    for qi in range(nq):
        out = idx.search(Q[qi], n_probe=3, top_k=top_k)
        Dd[qi, :len(out)] = [d for _, d in out]
        Id[qi, :len(out)] = [nd.id for nd, _ in out]
    flat = np.array([i for l in lists0 for i in l], dtype=np.int64)
    ptr = np.cumsum([0] + [len(l) for l in lists0]).astype(np.int64)
    np.savez_compressed(os.path.join(OUT, "ivf.npz"), X=X, ids=ids, Q=Q, k=np.int64(k), init_rows=init_rows,
                        centroids=centroids, list_ptr=ptr, list_ids=flat, A=A, add_ids=add_ids,
                        add_cluster=add_cluster, cdist=cdist, dead=dead, Dd=Dd, Id=Id, **res)


def gen_metadata_filter():
    """Reference StorageService.filter_by_metadata (storage_service.py:106-128) over an MMapNodeStorage in a
    temporary directory: metadata with ints / floats / bools / strings / None / lists / missing keys, and the id
    sets the reference returns for a list of filters."""
    import json
    import tempfile

    from vector_db.services.storage_service import StorageService
    from vector_db.types import Node
    rng = np.random.default_rng(77)
    colors = ["red", "green", "blue", None]
    mds = []
    for i in range(120):
        md = {"parity": i % 2, "color": colors[i % 4], "score": float(i % 5) / 2, "flag": bool(i % 3 == 0),
              "tags": ["a", "b"] if i % 6 == 0 else ["a"]}
        if i % 7 == 0:
            md.pop("color")
        if i % 11 == 0:
            md = {}
        mds.append(md)
    filters = [{"parity": 1}, {"parity": 1.0}, {"parity": True}, {"color": "red"}, {"color": None}, {"color": "purple"},
               {"score": 1}, {"score": 0.5, "flag": True}, {"flag": 1}, {"tags": ["a", "b"]}, {"tags": ["a"]},
               {"nokey": None}, {"nokey": 3}, {}, {"parity": 0, "color": "green", "flag": False}]
    with tempfile.TemporaryDirectory() as tmp:
        svc = StorageService(os.path.join(tmp, "db.npy"), dim=8, capacity=256)
        for i, md in enumerate(mds):
            svc.save(Node(id=i, embedding=rng.standard_normal(8).astype(np.float32) + 1.0, metadata=md, content=f"d{i}"))
        deleted = [5, 18, 40]
        for i in deleted:
            svc.delete(i)
        expect = [sorted(svc.filter_by_metadata(f)) for f in filters]
    with open(os.path.join(OUT, "metadata_filter.json"), "w") as fh:
        json.dump({"metadatas": mds, "deleted": deleted, "filters": filters, "expect": expect}, fh)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref_hnsw, ref_pq, Storage, Node, ref_dist, topk_sorted = _import_reference()
    gen_exact(ref_dist, topk_sorted)
    gen_pq(ref_pq)
    gen_pq8(ref_pq)
    gen_hnsw(ref_hnsw, Storage, Node, "small", n=400, dim=16, M=8, efc=50, nq=24, ef=50, k=10, seed=11)
    gen_hnsw(ref_hnsw, Storage, Node, "medium", n=2000, dim=48, M=16, efc=100, nq=32, ef=64, k=10, seed=12)
    gen_ivf(Storage, Node)
    gen_metadata_filter()
    gen_cosine(ref_dist, topk_sorted)
    gen_adc(ref_dist)
    gen_hnsw_toplevel(ref_hnsw, Storage, Node)
    print("golden fixtures written to", os.path.abspath(OUT))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "metadata":
    _import_reference()
    gen_metadata_filter()
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "pins":        # r2 additions only
    _ref = _import_reference()
    gen_cosine(_ref[4], _ref[5])
    gen_adc(_ref[4])
    gen_hnsw_toplevel(_ref[0], _ref[2], _ref[3])
elif __name__ == "__main__":
    main()
