"""vdb_config.n_devices (csrc/multi_gpu.cu): one vdb_search_batch from ONE process fans out over the id-range
shards and merges — what a JVM caller gets through nativeSearchBatch (native/bridge/VectorDBJNI.cpp:184-200). The bar: the
answer is the single-GPU answer BIT FOR BIT for any number of shards (PQ + rerank included: the shards agree on one global
shortlist before re-scoring). On a one-GPU box the shards share the device (VDB_MULTI_SAME_DEVICE=1: same code path); with >= 2
GPUs visible the same test also runs on real peers."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def v():
    import __graft_entry__ as ge
    ge.build_product()
    import vector_db_b200
    return vector_db_b200


def _modes():
    import torch
    return ["same"] + (["peers"] if torch.cuda.device_count() >= 2 else [])


@pytest.mark.parametrize("metric", [0, 1], ids=["sql2", "cosine"])
def test_flat_multi_device_equals_single(v, po, monkeypatch, metric):
    import torch
    n, dim, nq, k = 30001, 96, 1300, 20                     # ragged shards, two query sub-batches
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    ref = v.NativeFlatIndex(dim, n, metric=metric)
    ref.addBatch(np.arange(n, dtype=np.int32), X)
    ri, rd = ref.searchBatch(Q, k)
    ref.close()
    for mode in _modes():
        g = 3 if mode == "same" else min(torch.cuda.device_count(), 4)
        if mode == "same":
            monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
        else:
            monkeypatch.delenv("VDB_MULTI_SAME_DEVICE", raising=False)
        f = v.NativeFlatIndex(dim, n, metric=metric, n_devices=g)
        f.addBatch(np.arange(n, dtype=np.int32), X)
        ids, d = f.searchBatch(Q, k)
        assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd)), mode
        tq = torch.from_numpy(Q).cuda()
        di, dd = f.search_device(tq, k)                       # device buffers on the front device
        torch.cuda.synchronize()
        assert np.array_equal(di.cpu().numpy(), ri) and np.array_equal(bits(dd.cpu().numpy()), bits(rd)), mode
        assert f.size() == n
        f.close()


@pytest.mark.parametrize("rerank", [0, 300])
def test_pq_multi_device_equals_single_for_any_shard_count(v, po, monkeypatch, rerank):
    import torch
    n, dim, M, nq, k = 40003, 64, 8, 1100, 10
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    ref = v.NativePqIndex(dim, n, M=M, rerank=rerank)
    ref.train(X[:2000]); ref.addBatch(np.arange(n, dtype=np.int32), X)
    ri, rd = ref.searchBatch(Q, k)
    cb = ref.codebooks()
    ref.close()
    for mode in _modes():
        counts = (2, 5) if mode == "same" else (min(torch.cuda.device_count(), 4),)
        for g in counts:
            if mode == "same":
                monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
            else:
                monkeypatch.delenv("VDB_MULTI_SAME_DEVICE", raising=False)
            f = v.NativePqIndex(dim, n, M=M, rerank=rerank, n_devices=g)
            f.train(X[:2000])                                 # trained on shard 0, copied to the others
            assert np.array_equal(bits(f.codebooks()), bits(cb))
            f.addBatch(np.arange(n, dtype=np.int32), X)
            ids, d = f.searchBatch(Q, k)
            assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd)), (mode, g)
            f.close()
    # rows generated on the devices (the bench path): shards generate their own id range
    monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
    a = v.NativePqIndex(dim, n, M=M, rerank=rerank); a.set_codebooks(cb); a.fill_synthetic(42, n); a.encode_all()
    b = v.NativePqIndex(dim, n, M=M, rerank=rerank, n_devices=4); b.set_codebooks(cb); b.fill_synthetic(42, n); b.encode_all()
    ai, ad = a.searchBatch(Q[:200], k); bi, bd = b.searchBatch(Q[:200], k)
    assert np.array_equal(ai, bi) and np.array_equal(bits(ad), bits(bd)) and np.array_equal(ai, ri[:200])
    a.close(); b.close()


def test_pq_multi_device_local_shortlists_retry_when_rows_are_not_spread_evenly(v, po, monkeypatch):
    """shards keep LOCAL shortlists of ~2R/g + 64 keys (csrc/multi_gpu.cu); when the whole global shortlist sits in one shard the
    boundary check must notice and the sub-batch is repeated with full-length lists — the answer stays the single-GPU answer"""
    monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
    n, dim, M, nq, k, R, g = 40000, 64, 8, 300, 10, 300, 5
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    c = X[7].copy()
    Q = (c[None, :] + 0.01 * Q).astype(np.float32)
    X = np.ascontiguousarray(X[np.argsort(((X - c[None, :]) ** 2).sum(axis=1), kind="stable")])
    X[n // g:] += np.float32(6.0)                                  # everything near the queries lives in shard 0
    T = np.ascontiguousarray(X[::5][:2000])
    ref = v.NativePqIndex(dim, n, M=M, rerank=R)
    ref.train(T); ref.addBatch(np.arange(n, dtype=np.int32), X)
    ri, rd = ref.searchBatch(Q, k)
    cb = ref.codebooks()
    ref.close()
    f = v.NativePqIndex(dim, n, M=M, rerank=R, n_devices=g)
    f.set_codebooks(cb); f.addBatch(np.arange(n, dtype=np.int32), X)
    ids, d = f.searchBatch(Q, k)
    assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd))
    assert f.stats()["shortlist_retries"] > 0
    f.close()
    # evenly spread rows: no retry
    X2 = po.synth(44, n, dim)
    f = v.NativePqIndex(dim, n, M=M, rerank=R, n_devices=g)
    f.set_codebooks(cb); f.addBatch(np.arange(n, dtype=np.int32), X2)
    f.searchBatch(po.synth(45, nq, dim), k)
    assert f.stats()["shortlist_retries"] == 0
    f.close()


def test_multi_device_edge_shapes(v, po, monkeypatch):
    """tiny and ragged cases: fewer rows than shards ask for, k beyond a shard's rows, a shortlist longer than a shard, rows arriving in
    several calls that straddle shard boundaries — always the single-GPU answer"""
    monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
    dim, M = 32, 4
    Q = po.synth(43, 7, dim)
    for n, g, k in ((10, 3, 5), (50, 4, 20), (301, 5, 64)):
        X = po.synth(42, n, dim)
        ref = v.NativeFlatIndex(dim, 400)
        ref.addBatch(np.arange(n, dtype=np.int32), X)
        ri, rd = ref.searchBatch(Q, k)
        ref.close()
        f = v.NativeFlatIndex(dim, 400, n_devices=g)                       # capacity 400 -> shards of 100 / 80 rows: later shards stay empty
        for lo in range(0, n, 37):                                         # several calls, pieces straddle shard boundaries
            hi = min(n, lo + 37)
            f.addBatch(np.arange(lo, hi, dtype=np.int32), X[lo:hi])
        ids, d = f.searchBatch(Q, k)
        assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd)), (n, g, k)
        assert f.size() == n
        f.close()
    # add -> search -> add -> search: rows that arrive after the first search go to the last shard (until the next rebuild) — far
    # beyond an even share of the capacity; the answer stays the single index's
    n, g, k = 390, 4, 15
    X = po.synth(42, n, dim)
    ref = v.NativeFlatIndex(dim, 400)
    f = v.NativeFlatIndex(dim, 400, n_devices=g)
    for lo, hi in ((0, 40), (40, 200), (200, 390)):
        for idx in (ref, f):
            idx.addBatch(np.arange(lo, hi, dtype=np.int32), X[lo:hi])
        ri, rd = ref.searchBatch(Q, k)
        ids, d = f.searchBatch(Q, k)
        assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd)), hi
    with pytest.raises(Exception):
        f.addBatch(np.arange(390, 420, dtype=np.int32), po.synth(44, 30, dim))      # the FRONT index enforces max_elements
    f.buildIndex()                                                                  # rebuildIndex splits the rows evenly again
    ids, d = f.searchBatch(Q, k)
    assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd)) and f.size() == n
    ref.close(); f.close()
    # PQ + rerank: shortlist (R = 200) longer than every shard (<= 80 rows), k beyond a shard
    n, g, k, R = 301, 5, 90, 200
    X = po.synth(42, n, dim)
    ref = v.NativePqIndex(dim, 400, M=M, rerank=R)
    ref.train(X[:256]); ref.addBatch(np.arange(n, dtype=np.int32), X)
    ri, rd = ref.searchBatch(Q, k)
    cb = ref.codebooks()
    ref.close()
    f = v.NativePqIndex(dim, 400, M=M, rerank=R, n_devices=g)
    f.set_codebooks(cb)
    for lo in range(0, n, 37):
        hi = min(n, lo + 37)
        f.addBatch(np.arange(lo, hi, dtype=np.int32), X[lo:hi])
    ids, d = f.searchBatch(Q, k)
    assert np.array_equal(ids, ri) and np.array_equal(bits(d), bits(rd))
    f.close()


def test_multi_device_rejects_graph_indexes_and_unsupported_calls(v, monkeypatch):
    monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
    with pytest.raises(Exception, match="replica"):
        v.NativeHnswPqIndex(64, 1000, pqM=8, n_devices=2)
    f = v.NativeFlatIndex(32, 1000, n_devices=2)
    with pytest.raises(Exception, match="multi-device"):
        f.save("/tmp/x.vdb")
    ids, d = f.searchBatch(np.zeros((3, 32), np.float32), 5)          # empty index: empty lists
    assert (ids == -1).all() and np.isinf(d).all()
    f.close()
