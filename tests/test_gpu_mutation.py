"""GPU tests (-m gpu) of the CRUD side of the path, through the C ABI, against the oracle's
restatement of the Java classes:
  * removeVector (index/HnswPqIndex.java:542-596, HnswIndex.java:254-317): same graph, same entry point, same answers after
    a mixed add / remove / add sequence, including Java's code-row reuse after a removal and the duplicate-id rule;
  * incremental addVector: only the adjacency rows an insert touched travel to the device — results must equal a full upload;
  * snapshot as a SHARED format: a file written by vdb_save is read by the oracle, a file written by the oracle is read
    by vdb_load, and both sides answer identically from it; corrupt files are rejected, not executed."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
STRICT = dict(trav_store=1, trav_batch=1)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def v():
    import __graft_entry__ as ge
    ge.build_product()
    import vector_db_b200
    return vector_db_b200


def _same_graph(g, o):
    levels, adj, entry = g.graph()
    ol, oadj, oentry = o.export_graph()
    return entry == oentry and np.array_equal(levels, ol) and all(np.array_equal(adj[l], oadj[l]) for l in range(len(oadj)))


def _mutate(po, g, o, X, script):
    for op, i in script:
        if op == "add":
            a, b = g.addVector(int(i), X[i]), o.add(int(i), X[i])
        else:
            a, b = g.removeVector(int(i)), o.remove(int(i))
        assert a == b, (op, i, a, b)


@pytest.mark.parametrize("kind", ["hnswpq", "hnsw"])
def test_remove_vector_matches_java_semantics(v, po, kind):
    n, dim, M, k = 700, 32, 4, 8
    X, Q = po.synth(42, n + 60, dim), po.synth(43, 40, dim)
    maxel = 3000                                           # training trigger at 300 inserts (HnswPqIndex.java:172)
    if kind == "hnswpq":
        g = v.NativeHnswPqIndex(dim, maxel, pqM=M, level_seed=11, **STRICT)
        o = po.OracleHnsw(po.KIND_HNSWPQ, dim, maxel, pq_m=M, level_seed=11)
    else:
        g = v.NativeHnswIndex(dim, maxel, level_seed=11, **STRICT)
        o = po.OracleHnsw(po.KIND_HNSW, dim, maxel, level_seed=11)
    g.addBatch(np.arange(n, dtype=np.int32), X[:n])
    o.add_batch(X[:n], np.arange(n, dtype=np.int32))
    ids0, d0 = g.searchBatch(Q, k)
    entry = g.graph()[2]
    rng = np.random.default_rng(5)
    victims = [int(x) for x in rng.choice(n, 40, replace=False)]
    if entry not in victims:
        victims[3] = entry                                 # the entry point goes too (ConcurrentHashMap-order replacement)
    script = [("rm", i) for i in victims[:20]] + [("add", n + j) for j in range(30)] + [("rm", i) for i in victims[20:]]
    script += [("rm", victims[0]), ("add", 5 if 5 not in victims else 6), ("add", n + 31), ("rm", n + 2), ("add", n + 2)]
    _mutate(po, g, o, X, script)
    assert g.size() == o.live
    assert _same_graph(g, o)                               # inserts / removals use Java's arithmetic on both sides
    o.set_arith(po.ARITH_GPU_TREE)                         # searches: strict mode's lane order, which the oracle twin reproduces bit for bit
    ids, d = g.searchBatch(Q, k)
    oi, od, _ = o.search(Q, k)
    assert np.array_equal(ids, oi) and np.array_equal(bits(d), bits(od))
    assert not np.isin(ids, [x for x in victims if x != n + 2]).any()
    if kind == "hnswpq":
        assert np.array_equal(g.codes(), o.codes())        # incl. the rows whose Java code row was overwritten after a removal
    # rebuildIndex drops the removed vectors for good and rebuilds by ascending id
    o.set_arith(po.ARITH_JAVA_SEQ)
    g.buildIndex(); o.rebuild()
    o.set_arith(po.ARITH_GPU_TREE)
    assert g.size() == o.live == o.size
    assert _same_graph(g, o)
    ids, d = g.searchBatch(Q, k)
    oi, od, _ = o.search(Q, k)
    assert np.array_equal(ids, oi) and np.array_equal(bits(d), bits(od))
    g.close()


def test_incremental_add_uploads_only_changed_rows(v, po):
    n, dim, M, k = 1500, 32, 4, 10
    X, Q = po.synth(42, n, dim), po.synth(43, 50, dim)
    g = v.NativeHnswPqIndex(dim, 4000, pqM=M, level_seed=3)
    g.addBatch(np.arange(1000, dtype=np.int32), X[:1000])
    g.searchBatch(Q[:1], k)                                # first search: full upload
    for i in range(1000, n):                               # every add is followed by a search, like VectorDatabase callers do
        assert g.addVector(i, X[i])
        if i % 50 == 0:
            g.searchBatch(Q[:2], k)
    ids, d = g.searchBatch(Q, k)
    levels, adj, entry = g.graph()
    h = v.NativeHnswPqIndex(dim, 4000, pqM=M, level_seed=3)      # same state, uploaded in one go
    h.addBatch(np.arange(n, dtype=np.int32), X)
    h.set_codebooks(g.codebooks()); h.set_graph(levels, list(adj), entry); h.set_codes(g.codes())
    ids2, d2 = h.searchBatch(Q, k)
    assert np.array_equal(ids, ids2) and np.array_equal(bits(d), bits(d2))
    assert not g.addVector(7, X[7])                        # duplicate id: false, nothing changes (HnswPqIndex.java:157-159)
    assert g.size() == n
    g.close(); h.close()


def test_snapshot_is_a_shared_format(v, po, tmp_path):
    n, dim, M, k = 900, 32, 4, 10
    X, Q = po.synth(42, n, dim), po.synth(43, 30, dim)
    g = v.NativeHnswPqIndex(dim, 3000, pqM=M, level_seed=21, **STRICT)
    g.addBatch(np.arange(n, dtype=np.int32), X)
    for i in (3, 77, 500):
        assert g.removeVector(i)
    ids, d = g.searchBatch(Q, k)
    p1 = str(tmp_path / "gpu.vdb")
    g.save(p1)
    # library -> oracle
    o = po.oracle_from_snapshot(p1)
    o.set_arith(po.ARITH_GPU_TREE)
    oi, od, _ = o.search(Q, k)
    assert np.array_equal(ids, oi) and np.array_equal(bits(d), bits(od))
    # oracle -> library (the oracle mutates the state first, so the file is genuinely its own)
    o.set_arith(po.ARITH_JAVA_SEQ)                         # mutations in Java's arithmetic, searches in the strict twin's
    assert o.remove(10) and o.add(5000, X[10])
    o.set_arith(po.ARITH_GPU_TREE)
    p2 = str(tmp_path / "oracle.vdb")
    po.oracle_to_snapshot(o, p2, 3000, 64, 64, level_seed=21)
    h = v.NativeHnswPqIndex(dim, 3000, pqM=M, level_seed=21, **STRICT)
    h.load(p2)
    assert h.size() == o.live and h.trained
    hi, hd = h.searchBatch(Q, k)
    oi, od, _ = o.search(Q, k)
    assert np.array_equal(hi, oi) and np.array_equal(bits(hd), bits(od))
    # the loaded index keeps working like the saved one: same level stream, same training state
    o.set_arith(po.ARITH_JAVA_SEQ)
    assert h.addVector(6000, X[20]) and o.add(6000, X[20])
    o.set_arith(po.ARITH_GPU_TREE)
    hi, hd = h.searchBatch(Q, k)
    oi, od, _ = o.search(Q, k)
    assert np.array_equal(hi, oi) and np.array_equal(bits(hd), bits(od))
    # corrupt files are rejected before anything reaches a kernel
    raw = bytearray(open(p1, "rb").read())
    s = po.snapshot_read(p1)
    bad = dict(s); bad["adj"] = s["adj"].copy(); bad["adj"][0, 5, 0] = 10 ** 6
    p3 = str(tmp_path / "bad_adj.vdb"); po.snapshot_write(p3, bad)
    bad2 = dict(s); bad2["hdr"] = s["hdr"].copy(); bad2["hdr"]["n"] = -5
    p4 = str(tmp_path / "bad_n.vdb"); po.snapshot_write(p4, bad2)
    p5 = str(tmp_path / "trunc.vdb"); open(p5, "wb").write(raw[: len(raw) // 2])
    for path in (p3, p4, p5):
        z = v.NativeHnswPqIndex(dim, 3000, pqM=M, level_seed=21)
        with pytest.raises(Exception):
            z.load(path)
        assert z.size() == 0
        z.close()
    c = v.NativeHnswIndex(dim, 3000)                       # other index type / metric: refused
    with pytest.raises(Exception):
        c.load(p1)
    c.close()
    g.close(); h.close()
