"""CPU suite: pins the oracle (oracle/oracle.cpp) against
  (1) golden vectors produced by the UNMODIFIED reference compiled here (tests/golden/ref_native.npz),
  (2) the compiled reference itself when oracle/_ref is present,
  (3) the reference's own known-answer tests (VectorTest.java, HnswIndexTest.java, RecallOptimizationTest.java)."""
import math

import numpy as np
import pytest


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


# ---------- (1) golden vectors from the compiled reference ----------
def test_distance_orders_match_reference_golden(po, golden_ref):
    for d in (3, 8, 13, 64):
        a, b = po.synth(7, 20, d), po.synth(8, 20, d)
        for name, order in (("seq", 0), ("r1", 1), ("r2", 2)):
            got = np.array([po.l2sq(a[i], b[i], order) for i in range(20)], dtype=np.float32)
            assert np.array_equal(bits(got), bits(golden_ref["%s_%d" % (name, d)])), (name, d)
        got = np.array([po.lib().orc_native_cosine(a[i], b[i], d) for i in range(20)], dtype=np.float32)
        assert np.array_equal(bits(got), bits(golden_ref["ncos_%d" % d]))


def test_pq_train_encode_search_match_reference_golden(po, golden_ref):
    n, dim, M, nq, k, ntrain = [int(v) for v in golden_ref["meta"]]
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    # PQIndex::train (PQIndex.cpp:29-105) and HNSWPQIndex::train (HNSWPQIndex.cpp:127-244): bit-exact codebooks
    cb_pq = po.pq_train(X[:ntrain], M, 25, po.TRAIN_NATIVE_PQ)
    assert np.array_equal(bits(cb_pq), bits(golden_ref["pq_codebooks"]))
    cb_hp = po.pq_train(X[:ntrain], M, 25, po.TRAIN_NATIVE_HNSWPQ)
    assert np.array_equal(bits(cb_hp), bits(golden_ref["hnswpq_codebooks"]))
    # encode (R1 argmin)
    assert np.array_equal(po.pq_encode(X, cb_pq, po.ARITH_NATIVE_AVX2), golden_ref["pq_codes"])
    assert np.array_equal(po.pq_encode(X[:256], cb_hp, po.ARITH_NATIVE_AVX2), golden_ref["hnswpq_codes"])
    # PQIndex::search: LUT (R2) + ADC (R3) + partial_sort on (dist,id)
    ids, d = po.pq_flat_search(Q, cb_pq, golden_ref["pq_codes"], k, po.ARITH_NATIVE_AVX2, ids=golden_ref["ids"])
    assert np.array_equal(ids, golden_ref["pq_ids"])
    assert np.array_equal(bits(d), bits(golden_ref["pq_dists"]))


# ---------- (2) live comparison with oracle/_ref (only where it was built) ----------
def test_oracle_vs_compiled_reference_live(po):
    if not po.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    n, dim, M = 800, 48, 6
    X, Q = po.synth(5, n, dim), po.synth(6, 8, dim)
    ref = po.RefPQ(dim, 2000, M)
    ref.train(X[:500])
    ids = np.arange(n, dtype=np.int32)[::-1].copy()
    ref.add_batch(X, ids)
    cb = po.pq_train(X[:500], M, 25, po.TRAIN_NATIVE_PQ)
    assert np.array_equal(bits(cb), bits(ref.codebooks()))
    codes = po.pq_encode(X, cb, po.ARITH_NATIVE_AVX2)
    assert np.array_equal(codes, ref.codes())
    rid, rd = ref.search_batch(Q, 5)
    oid, od = po.pq_flat_search(Q, cb, codes, 5, po.ARITH_NATIVE_AVX2, ids=ids)
    assert np.array_equal(rid, oid) and np.array_equal(bits(rd), bits(od))
    # untrained PQIndex::searchBatch writes -1 (PQIndex.cpp:259-270)
    r2 = po.RefPQ(dim, 100, M)
    rid, rd = r2.search_batch(Q, 5)
    assert (rid == -1).all()
    r2.close(); ref.close()
    # adcDistanceScalar == sequential R3 (ADCUtils.cpp:11-17)
    lut = po.pq_lut(Q[0], cb, po.ARITH_NATIVE_AVX2)
    got = po.adc_scan(lut, codes[:50])
    want = np.array([po.ref().ref_adc_scalar(lut, codes[i], M, 256) for i in range(50)], dtype=np.float32)
    assert np.array_equal(bits(got), bits(want))


# ---------- (3) the reference's own known-answer tests ----------
def test_java_vector_known_answers(po):
    L = po.lib()
    f = lambda *v: np.array(v, dtype=np.float32)
    # VectorTest.java:39-109 (tolerance 1e-4 there; exact values here)
    assert abs(L.orc_java_euclidean(f(1, 0, 0), f(0, 1, 0), 3) - math.sqrt(2.0)) < 1e-4
    assert abs(L.orc_java_cosine_similarity(f(1, 0, 0), f(0, 1, 0), 3) - 0.0) < 1e-4
    assert abs(L.orc_java_cosine_similarity(f(1, 0, 0), f(1, 1, 0), 3) - 1 / math.sqrt(2.0)) < 1e-4
    assert abs(L.orc_java_norm(f(3, 4), 2) - 5.0) < 1e-4
    out = np.empty(2, np.float32)
    L.orc_java_normalize(f(3, 4), 2, out)
    assert np.allclose(out, [0.6, 0.8], atol=1e-4)


def test_java_random_known_stream(po):
    # java.util.Random(42): first nextInt() values are documented widely; nextInt(bound)/nextFloat derive from next()
    ints = np.empty(4, np.int32); fl = np.empty(4, np.float32)
    po.lib().orc_java_random_probe(42, 10, 4, ints, fl)
    # new Random(42): nextInt(10) -> 0, then nextFloat() ...; check determinism + ranges + the known first draw
    assert ints[0] == 0
    assert ((0 <= ints) & (ints < 10)).all() and ((0 <= fl) & (fl < 1)).all()
    ints2 = np.empty(3, np.int32); fl2 = np.empty(3, np.float32)
    po.lib().orc_java_random_probe(0, 1 << 30, 3, ints2, fl2)
    assert ((0 <= ints2) & (ints2 < (1 << 30))).all()


def test_compression_config_recommended(po):
    # CompressionConfig.java:114-130 and SURVEY §2 row 2
    for dim, m in ((512, 64), (1000, 125), (768, 96), (128, 16)):
        assert po.lib().orc_recommended_pq_m(dim) == m
    assert po.lib().orc_java_search_ef(10000, 10, 64) == 1500
    assert po.lib().orc_java_search_ef(100000, 10, 64) == 5000


def test_hnsw_index_behaviour_like_HnswIndexTest(po):
    # HnswIndexTest.java:89-164: d=10, 100 vectors, k=10 -> exactly k ascending results; planted all-ones neighbour top-1
    rng = np.random.default_rng(42)
    X = (rng.random((100, 10), dtype=np.float32) * 2 - 1).astype(np.float32)
    h = po.OracleHnsw(po.KIND_HNSW, 10, 1000)
    h.add_batch(X)
    ids, d, _ = h.search(X[:5] * 0.9, 10)
    assert (ids >= 0).all()
    assert (np.diff(d, axis=1) >= 0).all()
    e = po.OracleHnsw(po.KIND_HNSW, 10, 1000)
    ids, d, _ = e.search(X[:1], 10)
    assert (ids == -1).all()                               # empty index -> empty list (HnswIndexTest.java:120-129)
    target = np.ones((1, 10), dtype=np.float32)
    h.add_batch(target, ids=[9999])
    ids, d, _ = h.search(np.full((1, 10), 0.9, dtype=np.float32), 5)
    assert ids[0, 0] == 9999


def test_hnswpq_recall_like_RecallOptimizationTest(po):
    # RecallOptimizationTest.java:227-290 quick variant asserts Recall@10 >= 0.70; scaled down (d=64, N=2000) for CPU time
    N, D, M, Q, k = 2000, 64, 8, 40, 10
    X, Qs = po.synth(42, N, D), po.synth(43, Q, D)
    h = po.OracleHnsw(po.KIND_HNSWPQ, D, 2 * N, pq_m=M)
    h.add_batch(X)
    assert h.trained                                        # training fired at min(10000, maxElements/10) = 400 samples
    ids, d, st = h.search(Qs, k)
    gt, gd = po.flat_search(X, Qs, k)
    rec = np.mean([len(set(a) & set(b)) / k for a, b in zip(ids, gt)])
    assert rec >= 0.85
    assert (ids >= 0).all() and (np.diff(d, axis=1) >= 0).all()       # exactly K results, ascending
    # self query returns itself first (RecallDiagnosticTest.java:207-257, test_hnsw.cpp:34-55)
    ids, d, _ = h.search(X[:3], k)
    assert list(ids[:, 0]) == [0, 1, 2] and (d[:, 0] == 0).all()
    # first 399 rows were added raw: all-zero codes (HnswPqIndex.java:104,391)
    codes = h.codes()
    assert (codes[:399] == 0).all() and codes[399:].any()


def test_java_oracle_regression_golden(po, golden_java):
    n, dim, M, nq, k, maxel = [int(v) for v in golden_java["meta"]]
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    h = po.OracleHnsw(po.KIND_HNSWPQ, dim, maxel, pq_m=M, level_seed=99)
    h.add_batch(X)
    assert np.array_equal(bits(h.codebooks()), bits(golden_java["hp_codebooks"]))
    assert np.array_equal(h.codes(), golden_java["hp_codes"])
    ids, d, _ = h.search(Q, k)
    assert np.array_equal(ids, golden_java["hp_ids"]) and np.array_equal(bits(d), bits(golden_java["hp_dists"]))
    # traversal under the product's warp-tree arithmetic gives the same ids and the same (re-scored) distances
    h.set_arith(po.ARITH_GPU_TREE)
    ids2, d2, _ = h.search(Q, k)
    assert np.array_equal(ids2, ids) and np.array_equal(bits(d2), bits(d))
    c = po.OracleHnsw(po.KIND_HNSW, dim, maxel, level_seed=99)
    c.add_batch(X)
    ids, d, _ = c.search(Q, k)
    assert np.array_equal(ids, golden_java["hc_ids"]) and np.array_equal(bits(d), bits(golden_java["hc_dists"]))
    assert np.array_equal(bits(po.pq_train(X[:300], M, 25, po.TRAIN_JAVA)), bits(golden_java["java_codebooks"]))


def test_synthetic_generator_properties(po):
    a = po.fill_uniform(42, 0, 100000)
    assert a.min() >= -1 and a.max() < 1 and abs(a.mean()) < 0.01 and abs(a.var() - 1 / 3) < 0.01
    b = po.fill_uniform(42, 500, 100)
    assert np.array_equal(a[500:600], b)                     # counter based: any slice regenerates identically
