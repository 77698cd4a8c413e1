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
