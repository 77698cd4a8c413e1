"""World-size-2 gloo test of the host-side sharding logic: id-range shard plan, per-rank local top-k
with global ids, all_gather of the small lists, (dist,id) merge. The local searcher here is the ORACLE's flat search
(checker) so the test runs without a GPU; on a GPU box the same ShardedSearcher drives vdb_search_batch_device and
the CUDA merge kernel (tests/test_gpu_parity.py::test_sharded_merge_kernel)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, dim, nq, k, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle as po
    from vector_db_b200.sharded import ShardedSearcher, merge_topk_host, plan_shards
    lo, hi = plan_shards(n, world)[rank]
    X = po.fill_uniform(42, lo * dim, (hi - lo) * dim).reshape(hi - lo, dim)      # this rank generates only its shard
    Q = po.synth(43, nq, dim)

    def local_search(q, kk):
        ids, d = po.flat_search(X, q.numpy(), kk, ids=np.arange(lo, hi, dtype=np.int32))
        return torch.from_numpy(ids), torch.from_numpy(d)

    def merge(all_d, all_i, kk):
        d, i = merge_topk_host(all_d.numpy(), all_i.numpy(), kk)
        return torch.from_numpy(d), torch.from_numpy(i)

    s = ShardedSearcher(None, rank, world, merge_fn=merge)
    ids, d = s.search(torch.from_numpy(Q), k, local_search=local_search)
    if rank == 0:
        ret["ids"], ret["d"] = ids.numpy().copy(), d.numpy().copy()
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_search_world2_matches_single_index():
    from oracle import pyoracle as po
    po.build()
    n, dim, nq, k = 2001, 32, 9, 7          # odd n: ragged shards
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(2, 29517, n, dim, nq, k, ret), nprocs=2, join=True)
        ids, d = ret["ids"], ret["d"]
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    gi, gd = po.flat_search(X, Q, k)
    assert np.array_equal(ids, gi)
    assert np.array_equal(d.view(np.uint32), gd.view(np.uint32))


def _pq_data(po, n, dim, nq, skew):
    """rows, queries and training sample of the PQ shard tests. skew: the queries sit in one small cluster, the rows are stored in
    order of their distance to it and the last two thirds are moved far away, so the whole global shortlist lives in shard 0 — the
    case the local shortlists cannot cover"""
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    if skew:
        c = X[7].copy()
        Q = (c[None, :] + 0.01 * Q).astype(np.float32)
        X = np.ascontiguousarray(X[np.argsort(((X - c[None, :]) ** 2).sum(axis=1), kind="stable")])
        X[n // 3:] += np.float32(6.0)
    return X, Q, np.ascontiguousarray(X[::5][:600])


class _StubIndex:                                       # what search_pq reads from the local index on the CPU path
    def __init__(self, rerank):
        self.rerank = rerank


def _worker_pq(rank, world, port, n, dim, M, nq, k, R, skew, ret):
    """two-phase PQ search over id-range shards on CPU: the local phases are the ORACLE's (checker), the exchange, the local
    shortlist length, the boundary certainty rule and the retry are the product's host code (vector_db_b200/sharded.py)"""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle as po
    from vector_db_b200.sharded import ShardedSearcher, kth_key_host, merge_topk_host, plan_shards
    lo, hi = plan_shards(n, world)[rank]
    Xall, Q, T = _pq_data(po, n, dim, nq, skew)
    cb = po.pq_train(T, M, 25, po.TRAIN_NATIVE_PQ)
    X = Xall[lo:hi]
    codes = po.pq_encode(X, cb, po.ARITH_NATIVE_AVX2)
    gids = np.arange(lo, hi, dtype=np.int32)
    state = {"lens": []}

    def key_of(d, i):                                   # the library's key: ordered fp32 bits << 32 | (id ^ 0x80000000)
        b = d.view(np.uint32).astype(np.uint64)
        b = np.where(b & 0x80000000, ~b & 0xffffffff, b | 0x80000000)
        return (b << np.uint64(32)) | (i.astype(np.int64).astype(np.uint64) ^ np.uint64(0x80000000)) & np.uint64(0xffffffff)

    def phase1(q, kk, L):
        ids, d = po.pq_flat_search(q.numpy(), cb, codes, L, po.ARITH_NATIVE_AVX2, ids=gids)     # local top-L by (ADC dist, id)
        keys = key_of(d, ids)
        keys[ids < 0] = np.uint64(0xffffffffffffffff)
        state["ids"], state["keys"] = ids, keys
        state["lens"].append(L)
        return torch.from_numpy(keys.view(np.int64))

    def phase2(q, kk, kth):
        keep = state["keys"] <= kth.numpy().view(np.uint64)[:, None]
        out_i = np.full((nq, kk), -1, np.int32); out_d = np.full((nq, kk), np.inf, np.float32)
        for qi in range(nq):
            cand = state["ids"][qi][keep[qi]]
            if cand.size:
                i2, d2 = po.flat_search(Xall[cand], q.numpy()[qi:qi + 1], min(kk, cand.size), order=1, ids=cand.astype(np.int32))
                out_i[qi, :i2.shape[1]] = i2[0]; out_d[qi, :i2.shape[1]] = d2[0]
        return torch.from_numpy(out_i), torch.from_numpy(out_d)

    def merge(all_d, all_i, kk):
        d, i = merge_topk_host(all_d.numpy(), all_i.numpy(), kk)
        return torch.from_numpy(d), torch.from_numpy(i)

    def kth_fn(ak, r):
        kth, unc = kth_key_host(ak.numpy().view(np.uint64), r)
        return torch.from_numpy(kth.view(np.int64)), unc

    s = ShardedSearcher(_StubIndex(R), rank, world, merge_fn=merge)
    ids, d = s.search_pq(torch.from_numpy(Q), k, phase1=phase1, phase2=phase2, kth_fn=kth_fn)
    if rank == 0:
        ret["ids"], ret["d"], ret["lens"], ret["retries"] = ids.numpy().copy(), d.numpy().copy(), list(state["lens"]), s.shortlist_retries
    dist.barrier()
    dist.destroy_process_group()


def _check_pq(world, port, n, dim, M, nq, k, R, skew):
    from oracle import pyoracle as po
    po.build()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker_pq, args=(world, port, n, dim, M, nq, k, R, skew, ret), nprocs=world, join=True)
        ids, d, lens, retries = ret["ids"], ret["d"], list(ret["lens"]), ret["retries"]
    X, Q, T = _pq_data(po, n, dim, nq, skew)
    cb = po.pq_train(T, M, 25, po.TRAIN_NATIVE_PQ)
    codes = po.pq_encode(X, cb, po.ARITH_NATIVE_AVX2)
    gi, gd = po.pq_flat_search(Q, cb, codes, k, po.ARITH_NATIVE_AVX2, rerank=R, base=X, exact_order=1)
    assert np.array_equal(ids, gi)                       # the sharded answer IS the single-index answer (global top-R, then rerank)
    assert np.array_equal(d.view(np.uint32), gd.view(np.uint32))
    return lens, retries


def test_sharded_pq_rerank_world2_equals_single_index():
    lens, retries = _check_pq(2, 29519, 3001, 32, 4, 6, 5, 40, False)
    assert lens == [40] and retries == 0                 # two shards: full-length lists


def test_sharded_pq_rerank_world3_local_shortlists_and_retry():
    from vector_db_b200.sharded import local_shortlist_len
    n, dim, M, nq, k, R = 3001, 32, 4, 6, 5, 400
    L = local_shortlist_len(R, 3, k)
    assert k <= L < R
    lens, retries = _check_pq(3, 29521, n, dim, M, nq, k, R, False)
    assert lens == [L] and retries == 0                  # evenly spread rows: the short local lists are certified, no retry
    lens, retries = _check_pq(3, 29523, n, dim, M, nq, k, R, True)
    assert lens == [L, R] and retries > 0                # the whole shortlist in shard 0: detected, repeated with full lists, same answer
