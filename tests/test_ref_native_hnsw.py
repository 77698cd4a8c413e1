"""The reference's native HNSWIndex (native/index/HNSWIndex.cpp) — the one beam search of the reference that can run on this image —
compiled UNMODIFIED into oracle/_ref (oracle/Makefile; its three BLAS calls go through oracle/shim/cblas.h to the reference's own
fallback loops). CPU: it builds, answers, and its returned distances agree with the oracle's exact distance (north_star: 1e-5
relative). GPU: the product walks the REFERENCE-BUILT graph ("same persisted graph") and finds at least what the reference finds."""
import numpy as np
import pytest


def recall(ids, gt):
    return float(np.mean([len(set(a) & set(b)) / gt.shape[1] for a, b in zip(ids, gt)]))


@pytest.fixture(scope="module")
def native(po):
    if not po.ref_available() or not hasattr(po.ref(), "ref_hnsw_create"):
        pytest.skip("oracle/_ref without the native HNSWIndex (rebuild with `make -C oracle ref` where /root/reference exists)")
    n, dim, M = 3000, 32, 16
    X, Q = po.synth(42, n, dim), po.synth(43, 200, dim)
    r = po.RefHnsw(dim, n, M=M, efC=64, ef=64)
    r.add_batch(X, np.arange(n, dtype=np.int32))
    yield r, X, Q, M
    r.close()


def test_native_hnsw_reference_builds_and_answers(po, native):
    r, X, Q, M = native
    k = 10
    assert r.size() == X.shape[0]
    ids, d = r.search_batch(Q, k)
    gt, gd = po.flat_search(X, Q, k)
    assert (ids >= 0).all() and (np.diff(d, axis=1) >= 0).all()                    # full lists, ascending (HNSWIndex.cpp:340-357)
    assert recall(ids, gt) >= 0.7                                                   # uniform 32-d rows, ef ~ 54 (HNSWConfig::getEfSearch)
    # a6 exact distance: the reference's values (AVX2 order R1 / the BLAS form of the batch path) vs the oracle's R1, 1e-5 relative
    ex = np.array([[po.l2sq(Q[i], X[j], 0) for j in ids[i]] for i in range(40)], dtype=np.float32)
    assert np.max(np.abs(ex - d[:40]) / ex) < 1e-5
    i1, d1, c = r.search(Q[0], k)                                                   # single-query entry = the batch entry
    assert c == k and np.array_equal(i1, ids[0])
    levels, adj, entry = r.graph()
    assert levels[entry] == levels.max()                                            # entry point sits on the top level
    a0 = adj[0]
    assert ((a0 >= -1) & (a0 < X.shape[0])).all() and not (a0 == np.arange(X.shape[0])[:, None]).any()
    assert (a0 >= 0).sum(1).max() <= 2 * M                                          # Mmax0 = 2M (HNSWIndex.cpp pruneNeighbors)


@pytest.mark.gpu
def test_product_walks_the_reference_built_graph(po, native):
    import __graft_entry__ as ge
    ge.build_product()
    import vector_db_b200 as v
    r, X, Q, M = native
    n, dim, k = X.shape[0], X.shape[1], 10
    rid, _ = r.search_batch(Q, k)
    gt, _ = po.flat_search(X, Q, k)
    levels, adj, entry = r.graph()
    g = v.NativeHnswPqIndex(dim, 2 * n, M=2 * M, ef=64, pqM=8, level_seed=1234)
    g.addBatch(np.arange(n, dtype=np.int32), X)                                     # trains + encodes (300-insert trigger); own graph replaced below
    L = g.max_level
    # the product's convention (Java's, and the native class's): levels[i] = top layer of node i; adj[l][i][0..hnsw_m) padded with -1
    A = np.full((L, n, 2 * M), -1, dtype=np.int32)
    for lv, a in adj.items():
        if lv < L:            # the Java level cap (log(maxElements)/log(M)) can be below the native one: layers above it are dropped,
            A[lv] = a         # the native entry point (on every layer) stays the entry
    g.set_graph(np.minimum(levels, L - 1), list(A), entry)
    ids, d = g.searchBatch(Q, k)
    rg, rr = recall(ids, gt), recall(rid, gt)
    assert rg >= rr - 0.02, (rg, rr)                                                # same graph, beam >= the reference's
    assert (ids >= 0).all() and (np.diff(d, axis=1) >= 0).all()
    g.close()
