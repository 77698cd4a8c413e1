"""Multi-GPU sharding of the flat / PQ query path: one process per GPU (torch.distributed), base
vectors and PQ codes split by contiguous id range, query batch replicated, every rank returns its local top-k with
GLOBAL ids, and only those small lists cross NVLink: one all_gather of [nq][k] (dist,id) per rank followed by the
(dist,id) merge kernel (vdb_merge_topk_device). The reference has no multi-device path at all (SURVEY.md §2
"Parallelism / communication components: none exist"); this is new design, the merge order is the reference's
(dist,id) rule (PQIndex.cpp:200-205).

HNSW graphs do not shard by id range (a walk touches arbitrary nodes): for those, `shard_queries` splits the QUERY
batch over replicas and gathers the results — no data-path collective."""
import numpy as np


def plan_shards(n_total, world_size):
    """contiguous id ranges [start, stop) per rank; sizes differ by at most one row"""
    base, rem = divmod(int(n_total), int(world_size))
    out, start = [], 0
    for r in range(world_size):
        size = base + (1 if r < rem else 0)
        out.append((start, start + size))
        start += size
    return out


def shard_queries(nq, world_size):
    return plan_shards(nq, world_size)


def merge_topk_host(dists, ids, k):
    """numpy statement of the merge rule, used by the CPU (gloo) tests and as documentation of the kernel's contract:
    dists/ids [g][nq][k] -> [nq][k] ordered by (dist, id); id < 0 = empty slot."""
    g, nq, kk = dists.shape
    d = np.transpose(dists, (1, 0, 2)).reshape(nq, g * kk)
    i = np.transpose(ids, (1, 0, 2)).reshape(nq, g * kk)
    out_d = np.full((nq, k), np.inf, dtype=np.float32)
    out_i = np.full((nq, k), -1, dtype=np.int32)
    for q in range(nq):
        valid = i[q] >= 0
        order = np.lexsort((i[q][valid], d[q][valid]))[:k]
        out_d[q, :order.size] = d[q][valid][order]
        out_i[q, :order.size] = i[q][valid][order]
    return out_d, out_i


def kth_key_host(keys, R):
    """numpy statement of vdb_select_kth_key_ex_device: keys [g][nq][L] (uint64, each list ascending, 2^64-1 = empty) -> (the R-th
    smallest key per query = the boundary of the GLOBAL shortlist, the number of queries whose boundary is uncertain because a full
    list of a shard ends below it — only possible when L < R)"""
    g, nq, L = keys.shape
    kth = np.sort(np.transpose(keys, (1, 0, 2)).reshape(nq, g * L), axis=1)[:, min(R, g * L) - 1].copy()
    last = keys[:, :, L - 1]                                    # [g][nq]
    unc = int(((last != np.uint64(0xFFFFFFFFFFFFFFFF)) & (last < kth[None, :])).any(axis=0).sum()) if L < R else 0
    return kth, unc


def local_shortlist_len(R, g, k):
    """length of a shard's LOCAL shortlist: about twice its expected share of the global top-R (rows spread evenly: Binomial(R, 1/g),
    2 R / g + 64 is more than ten standard deviations up); the same rule as csrc/multi_gpu.cu"""
    return R if g <= 2 else min(R, max(k, 2 * ((R + g - 1) // g) + 64))


class ShardedSearcher:
    """Wraps one local GpuIndex shard per rank. `search(d_queries, k)` returns the merged global top-k on every rank."""

    def __init__(self, local_index, rank, world_size, group=None, merge_fn=None):
        self.index, self.rank, self.world, self.group = local_index, rank, world_size, group
        self._merge_fn = merge_fn            # injectable for CPU tests; default = CUDA merge kernel
        self.shortlist_retries = 0           # queries that needed full-length shard shortlists (search_pq)

    def _merge(self, all_d, all_i, k):
        if self._merge_fn is not None:
            return self._merge_fn(all_d, all_i, k)
        import torch
        from . import _lib
        g, nq, _ = all_d.shape
        out_d = torch.empty((nq, k), dtype=torch.float32, device=all_d.device)
        out_i = torch.empty((nq, k), dtype=torch.int32, device=all_d.device)
        st = torch.cuda.current_stream(all_d.device).cuda_stream
        _lib.check(_lib.lib().vdb_merge_topk_device(all_d.data_ptr(), all_i.data_ptr(), g, nq, k, out_d.data_ptr(),
                                                    out_i.data_ptr(), st))
        return out_d, out_i

    def search_pq(self, d_queries, k, phase1=None, phase2=None, kth_fn=None):
        """PQ shards (with or without rerank) in two phases so that the answer does not depend on the number of shards: every rank
        scans its codes (phase 1: local top-L keys, L = local_shortlist_len), ONE all_gather of the keys, the global R-th key per
        query bounds the global shortlist, every rank re-scores only its own rows inside it (phase 2) and the top-k lists are merged
        as in `search`. If a shard's local list ends below the boundary of some query (rows not spread evenly) the batch is repeated
        with full-length lists — every rank sees the same gathered keys, so all take the same decision.
        phase1(q, k, L) -> uint64 keys [nq][L]; phase2(q, k, kth) -> (ids, dists); kth_fn(all_keys [g][nq][L], R) -> (kth [nq],
        uncertain count); injectable for the CPU (gloo) tests."""
        import torch
        import torch.distributed as dist
        if phase1 is None:
            from . import _lib
            import ctypes as C
            L_, h = _lib.lib(), self.index._h
            st = torch.cuda.current_stream(d_queries.device).cuda_stream
            nq = d_queries.shape[0]

            def phase1(q, kk, L):
                R = C.c_int32(L)
                keys = torch.empty((nq, L), dtype=torch.int64, device=q.device)
                _lib.check(L_.vdb_pq_shortlist_device(h, q.data_ptr(), nq, kk, keys.data_ptr(), C.byref(R), st))
                assert R.value == L, (R.value, L)
                return keys

            def phase2(q, kk, kth):
                ids = torch.empty((nq, kk), dtype=torch.int32, device=q.device)
                d = torch.empty((nq, kk), dtype=torch.float32, device=q.device)
                _lib.check(L_.vdb_pq_rerank_device(h, q.data_ptr(), nq, kk, 0 if kth is None else kth.data_ptr(), ids.data_ptr(), d.data_ptr(), st))
                return ids, d

            def kth_fn(all_keys, R):
                out = torch.empty((nq,), dtype=torch.int64, device=all_keys.device)
                unc = torch.zeros((1,), dtype=torch.int32, device=all_keys.device)
                _lib.check(L_.vdb_select_kth_key_ex_device(all_keys.data_ptr(), all_keys.shape[0], nq, all_keys.shape[2], R, out.data_ptr(),
                                                           unc.data_ptr(), st))
                return out, int(unc.item())
        R = max(self.index_rerank(), k)
        if self.world == 1:
            phase1(d_queries, k, R)
            return phase2(d_queries, k, None)
        for L in dict.fromkeys((local_shortlist_len(R, self.world, k), R)):       # local lists first, full-length lists if needed
            keys = phase1(d_queries, k, L)
            nq = keys.shape[0]
            all_keys = torch.empty((self.world * nq, L), dtype=keys.dtype, device=keys.device)
            dist.all_gather_into_tensor(all_keys, keys.contiguous(), group=self.group)
            kth, unc = kth_fn(all_keys.view(self.world, nq, L), R)
            if unc == 0 or L == R:
                break
            self.shortlist_retries += unc
        return self.search(d_queries, k, local_search=lambda q, kk: phase2(q, kk, kth))

    def index_rerank(self):
        return int(getattr(self.index, "rerank", 0))

    def search(self, d_queries, k, local_search=None):
        import torch
        import torch.distributed as dist
        ids, d = (local_search or self.index.search_device)(d_queries, k)
        if self.world == 1:
            return ids, d
        nq = d.shape[0]
        all_d = torch.empty((self.world * nq, k), dtype=d.dtype, device=d.device)
        all_i = torch.empty((self.world * nq, k), dtype=ids.dtype, device=ids.device)
        dist.all_gather_into_tensor(all_d, d.contiguous(), group=self.group)      # rank-major concatenation
        dist.all_gather_into_tensor(all_i, ids.contiguous(), group=self.group)
        out_d, out_i = self._merge(all_d.view(self.world, nq, k), all_i.view(self.world, nq, k), k)
        return out_i, out_d
