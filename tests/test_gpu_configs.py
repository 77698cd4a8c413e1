"""GPU parity on BASELINE.json's REAL shapes (C1, C3, C4, C5 of SURVEY.md §8), driver-run (`-m gpu`), through the C ABI.

The small-shape suite (test_gpu_parity.py) pins every kernel bit for bit; this file holds the configurations the
headline numbers are quoted on to the same bar:
  C1  10 000 x 512-d, HNSW+PQ (64, 8), 1 000 queries, k = 10      (README benchmark; RecallOptimizationTest.java:35-60)
  C3  100 000 x 512-d, HNSW+PQ recommendedConfig(512), 256-query sample of the 4 096-query batch
  C4  1 000 000 x 768-d flat, k = 100, 64 queries, sq-L2 and cosine (ground truth of RecallOptimizationTest.java:151-164)
  C5  1 000 000 x 1000-d, PQ (125, 8), ADC scan -> top-1000 -> fp32 rerank -> k = 10 (PQIndex.cpp:146-213 + rerank)
Tolerances (north_star): PQ codes and exact-search ids bit-exact (ties by id), ADC / exact distances bit-equal here
(the oracle reproduces the reduction order), HNSW Recall@10 within 0.5 pt of the reference algorithm on the same graph.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CORES = os.cpu_count() or 8
STRICT = dict(trav_store=1, trav_batch=1)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def recall(a, b):
    return float(np.mean([len(set(x) & set(y)) / len(y) for x, y in zip(a, b)]))


@pytest.fixture(scope="module")
def v():
    import __graft_entry__ as ge
    ge.build_product()
    import vector_db_b200
    return vector_db_b200


def _build_hnswpq(v, po, n, dim, M, nq, seed=1234):
    """product build (GPU k-means + encode, graph construction), then the SAME state loaded into the oracle"""
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    g = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=seed)
    g.addBatch(np.arange(n, dtype=np.int32), X)
    g.buildIndex()
    levels, adj, entry = g.graph()
    cb, codes = g.codebooks(), g.codes()
    o = po.OracleHnsw(po.KIND_HNSWPQ, dim, 2 * n, pq_m=M, level_seed=seed)
    o.import_state(X, levels, adj, entry, codebooks=cb, codes=codes)
    return X, Q, g, o, (levels, adj, entry, cb, codes)


def _strict_twin(v, X, state, dim, M, n, seed=1234):
    levels, adj, entry, cb, codes = state
    s = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=seed, **STRICT)
    s.addBatch(np.arange(n, dtype=np.int32), X)
    s.set_codebooks(cb)
    s.set_graph(levels, list(adj), entry)
    s.set_codes(codes)
    return s


def _diff_report(ids, jid, d, jd, k):
    """per-query account of the id differences: how many are swaps inside the same top-k set, how many queries found a
    strictly better / worse k-th exact distance than the Java-order walk"""
    diff_q = np.where((ids != jid).any(axis=1))[0]
    same_set = sum(1 for q in diff_q if set(ids[q]) == set(jid[q]))
    better = sum(1 for q in diff_q if d[q, k - 1] < jd[q, k - 1])
    worse = sum(1 for q in diff_q if d[q, k - 1] > jd[q, k - 1])
    return dict(queries_differing=int(diff_q.size), same_set=same_set, gpu_better_kth=better, gpu_worse_kth=worse)


def test_c1_readme_shape_default_and_strict(v, po):
    n, dim, M, nq, k = 10_000, 512, 64, 1000, 10
    X, Q, g, o, state = _build_hnswpq(v, po, n, dim, M, nq)
    # PQ codes: bit-exact against the oracle's encoder on the product-trained codebooks (Java order, HnswPqIndex.java:355-366)
    assert np.array_equal(state[4], po.pq_encode(X, state[3], po.ARITH_JAVA_SEQ))
    ids, d = g.searchBatch(Q, k)
    jid, jd, jst = o.search(Q, k, nthreads=CORES)
    gt, _ = po.flat_search(X, Q, k, nthreads=CORES)
    rg, rj = recall(ids, gt), recall(jid, gt)
    assert rg >= 0.95 and abs(rg - rj) <= 0.005, (rg, rj)              # the metric's gate + north_star's 0.5 pt
    match = float((ids == jid).mean())
    assert match >= 0.99, (match, _diff_report(ids, jid, d, jd, k))
    same = ids == jid
    assert np.array_equal(bits(d[same]), bits(jd[same]))                # re-scored in Java's sequential order
    # strict mode: fp32 rows, one candidate per round == the Java pop order; identical to the oracle twin
    s = _strict_twin(v, X, state, dim, M, n)
    sid, sd = s.searchBatch(Q, k)
    o.set_arith(po.ARITH_GPU_TREE)
    tid, td, _ = o.search(Q, k, nthreads=CORES)
    assert np.array_equal(sid, tid) and np.array_equal(bits(sd), bits(td))
    assert (sid == jid).mean() >= 0.995
    s.close(); g.close()


def test_c3_headline_shape_sample(v, po):
    n, dim, M, nq, k = 100_000, 512, 64, 256, 10
    X, Q, g, o, state = _build_hnswpq(v, po, n, dim, M, nq)
    ids, d = g.searchBatch(Q, k)
    jid, jd, jst = o.search(Q, k, nthreads=CORES)
    gt, _ = po.flat_search(X, Q, k, nthreads=CORES)
    rg, rj = recall(ids, gt), recall(jid, gt)
    assert abs(rg - rj) <= 0.005, (rg, rj)
    # id agreement with the Java pop order. Measured on the B200 (tools/c3_idmatch.py -> profiles/r2_c3_idmatch.log): with ONE
    # candidate per round the ids are 100 % the oracle's for bf16 AND fp32 rows; with 64 candidates per round 8 of 256 queries
    # differ on bf16 rows (98.6 %) and 10 on int8 rows (97.7 %; the default here: dim % 128 == 0 and uniform rows pass the outlier
    # check), and in all but one of them the GPU's k-th exact distance is strictly SMALLER (the batched round expands a few
    # candidates the sequential loop would have skipped, +0.8 % distance evaluations). The graph is built on the device and is not
    # the same from run to run (like Java's own racy build), and with it the share of queries near such a boundary: over a dozen
    # runs the id agreement of this sample ranged from 96.5 % to 99.1 %. So the SHARP bar of the default mode is the direction of
    # the differences — at most 2 queries of the sample with a worse k-th distance than Java's, recall not below Java's (0.2 pt
    # slack) — next to >= 95 % equal ids.
    assert g.stats()["trav_format"] == 3                   # VDB_TRAV_AUTO picked int8
    match = float((ids == jid).mean())
    rep = _diff_report(ids, jid, d, jd, k)
    assert match >= 0.95, (match, rep)
    assert rep["gpu_worse_kth"] <= 2 and rg >= rj - 0.002, (match, rep, rg, rj)
    # forced bf16 rows: >= 96 % and at most two worse
    b16 = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=1234, trav_store=4)
    b16.addBatch(np.arange(n, dtype=np.int32), X)
    b16.set_codebooks(state[3]); b16.set_graph(state[0], list(state[1]), state[2]); b16.set_codes(state[4])
    fid, fd = b16.searchBatch(Q, k)
    b16.close()
    rep16 = _diff_report(fid, jid, fd, jd, k)
    assert float((fid == jid).mean()) >= 0.96 and rep16["gpu_worse_kth"] <= 2, rep16
    # one candidate per round on the bf16 rows (half the bytes of the strict fp32 mode): the Java pop order, ids identical
    b1 = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=1234, trav_batch=1, trav_store=4)
    b1.addBatch(np.arange(n, dtype=np.int32), X)
    b1.set_codebooks(state[3]); b1.set_graph(state[0], list(state[1]), state[2]); b1.set_codes(state[4])
    bid, bd = b1.searchBatch(Q[:64], k)
    b1.close()
    # bf16 rounding can flip one comparison early in a walk, after which that query's whole list may differ (the graph is built on the
    # device and differs from run to run, so WHICH query does varies): the bar is per id over the sample, with room for one such query
    assert (bid == jid[:64]).mean() >= 0.95
    same = ids == jid
    assert np.array_equal(bits(d[same]), bits(jd[same]))
    st = g.stats()
    assert abs(st["dist_evals"] / nq - jst[0] / nq) / (jst[0] / nq) < 0.05      # same amount of work as the Java walk
    # strict twin on a 64-query subset: exact
    s = _strict_twin(v, X, state, dim, M, n)
    sid, sd = s.searchBatch(Q[:64], k)
    o.set_arith(po.ARITH_GPU_TREE)
    tid, td, _ = o.search(Q[:64], k, nthreads=CORES)
    assert np.array_equal(sid, tid) and np.array_equal(bits(sd), bits(td))
    s.close(); g.close()


@pytest.mark.parametrize("metric,nq", [(0, 64), (1, 32)], ids=["sql2", "cosine"])
def test_c4_flat_1m_768(v, po, metric, nq):
    n, dim, k = 1_000_000, 768, 100
    f = v.NativeFlatIndex(dim, n, metric=metric)
    f.fill_synthetic(42, n)                                             # rows generated in HBM by the oracle's generator
    Q = po.synth(43, nq, dim)
    ids, d = f.searchBatch(Q, k)
    st = f.stats()
    f.close()
    X = po.synth(42, n, dim)
    if metric == 1:
        oi, od = po.flat_search(po.java_normalize_rows(X), po.java_normalize_rows(Q), k, metric=1, nthreads=CORES)
    else:
        oi, od = po.flat_search(X, Q, k, metric=0, nthreads=CORES)
    assert np.array_equal(ids, oi)
    assert np.array_equal(bits(d), bits(od))
    assert st["flat_certified"] + st["flat_fallback"] == nq and st["flat_certified"] >= nq * 0.9, st


def test_c5_adc_scan_rerank_1m_1000(v, po):
    n, dim, M, R, k, nq = 1_000_000, 1000, 125, 1000, 10, 64
    p = v.NativePqIndex(dim, n, M=M, rerank=R)
    S = po.synth(41, 10000, dim)
    p.train(S)
    p.fill_synthetic(42, n)
    p.encode_all()
    Q = po.synth(43, nq, dim)
    ids, d = p.searchBatch(Q, k)
    cb, codes = p.codebooks(), p.codes()
    p.close()
    assert np.array_equal(bits(cb), bits(po.pq_train(S, M, 25, po.TRAIN_NATIVE_PQ)))      # PQIndex.cpp:29-105 k-means, bit-equal
    X = po.synth(42, n, dim)
    sample = np.concatenate([np.arange(0, 8192), np.arange(n - 8192, n)])
    assert np.array_equal(codes[sample], po.pq_encode(X[sample], cb, po.ARITH_NATIVE_AVX2))   # codes bit-exact (row sample)
    oi, od = po.pq_flat_search(Q, cb, codes, k, po.ARITH_NATIVE_AVX2, rerank=R, base=X, exact_order=1, nthreads=CORES)   # re-score in order R1 (euclideanDistanceAVX2)
    assert np.array_equal(ids, oi)
    assert np.array_equal(bits(d), bits(od))
    # no-rerank form == PQIndex::search itself (ADC distances, (dist,id) order) on a smaller query set
    p2 = v.NativePqIndex(dim, n, M=M)
    p2.set_codebooks(cb)
    p2.fill_synthetic(42, n)
    p2.encode_all()
    ids2, d2 = p2.searchBatch(Q[:16], k)
    p2.close()
    oi2, od2 = po.pq_flat_search(Q[:16], cb, codes, k, po.ARITH_NATIVE_AVX2, nthreads=CORES)
    assert np.array_equal(ids2, oi2) and np.array_equal(bits(d2), bits(od2))


@pytest.mark.parametrize("shape", ["rowscale", "spiky"])
def test_auto_traversal_format_on_heavy_tailed_rows(v, po, shape):
    """VDB_TRAV_AUTO (the default) on non-uniform data. `rowscale`: Gaussian rows times a log-normal per-row scale (norms spread
    over two orders of magnitude) — a per-row int8 scale absorbs it, the outlier check passes, int8 rows are used and recall must
    stay within 0.5 pt of the Java-order walk. `spiky`: every row carries a few components 40 x larger than the rest — int8 would
    spend its 8 bits on the spikes, the check must reject it and fall back to bf16."""
    n, dim, M, nq, k = 20000, 128, 16, 200, 10
    rng = np.random.default_rng(7)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    if shape == "rowscale":
        X *= np.exp(rng.normal(0.0, 1.0, (n, 1))).astype(np.float32)
        Q *= np.exp(rng.normal(0.0, 1.0, (nq, 1))).astype(np.float32)
    else:
        spikes = rng.integers(0, dim, (n, 3))
        X[np.arange(n)[:, None], spikes] *= 40.0
    g = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=1234)
    g.addBatch(np.arange(n, dtype=np.int32), X)
    g.buildIndex()
    ids, d = g.searchBatch(Q, k)
    assert g.stats()["trav_format"] == (3 if shape == "rowscale" else 4)
    levels, adj, entry = g.graph()
    o = po.OracleHnsw(po.KIND_HNSWPQ, dim, 2 * n, pq_m=M, level_seed=1234)
    o.import_state(X, levels, adj, entry, codebooks=g.codebooks(), codes=g.codes())
    jid, jd, _ = o.search(Q, k, nthreads=CORES)
    gt, _ = po.flat_search(X, Q, k, nthreads=CORES)
    rg, rj = recall(ids, gt), recall(jid, gt)
    assert abs(rg - rj) <= 0.005, (rg, rj)
    assert (ids == jid).mean() >= 0.95
    same = ids == jid
    assert np.array_equal(bits(d[same]), bits(jd[same]))                # exact re-score: distances are Java's bit for bit
    g.close()
