#!/usr/bin/env python
"""tests/manual/shard_check.py — multi-GPU check of the id-range-sharded flat / PQ path, one process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/manual/shard_check.py
1. parity: sharded flat search (tensor-core path per shard + NCCL all_gather of the [nq][k] lists + (dist,id) merge kernel)
   against the CPU oracle's brute force over the whole logical dataset; same for the PQ ADC scan + rerank;
2. timing (CUDA events, max over ranks): C4 (1M x 768 flat, k=100, 4096 queries) and a C5-shaped PQ scan, rows split N/g."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def timed(torch, dist, fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c4-rows", type=int, default=1_000_000)
    ap.add_argument("--c5-rows", type=int, default=4_000_000)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist.init_process_group("nccl")
    from oracle import pyoracle as po      # checker: CPU brute force / reference-order scan
    from vector_db_b200 import synth as sy
    import vector_db_b200 as v
    from vector_db_b200.sharded import ShardedSearcher, plan_shards
    dev = torch.cuda.current_device()
    out = {"world": world}

    # ---- 1. parity, flat
    n, dim, nq, k = 200_000, 96, 64, 50
    lo, hi = plan_shards(n, world)[rank]
    f = v.NativeFlatIndex(dim, hi - lo, id_offset=lo, device=dev)
    f.fill_synthetic(42, hi - lo)
    Q = sy.synth(43, nq, dim)
    tq = torch.from_numpy(Q).cuda()
    ids, d = ShardedSearcher(f, rank, world).search(tq, k)
    if rank == 0:
        oi, od = po.flat_search(sy.synth(42, n, dim), Q, k, nthreads=16)
        out["flat_parity"] = {"ids_equal": float((ids.cpu().numpy() == oi).mean()),
                              "dist_bits_equal": float((d.cpu().numpy().view(np.uint32) == od.view(np.uint32)).mean())}
    st = f.stats(); out.setdefault("flat_stats", {})["rank0"] = [st["flat_certified"], st["flat_fallback"]] if rank == 0 else None
    f.close()

    # ---- 1b. parity, PQ scan + rerank
    n, dim, M, k, R = 100_000, 128, 16, 10, 200
    lo, hi = plan_shards(n, world)[rank]
    p = v.NativePqIndex(dim, hi - lo, M=M, id_offset=lo, rerank=R, device=dev)
    p.train(sy.synth(41, 3000, dim))
    p.fill_synthetic(42, hi - lo)
    p.encode_all()
    Q = sy.synth(43, 32, dim)
    tq = torch.from_numpy(Q).cuda()
    ids, d = ShardedSearcher(p, rank, world).search(tq, k)
    if rank == 0:
        # the reference order: per-shard ADC top-R -> fp32 rerank -> top-k, then the (dist,id) merge == global rerank of the union
        X = sy.synth(42, n, dim)
        cb = p.codebooks()
        parts = []
        for (a0, a1) in plan_shards(n, world):
            codes = po.pq_encode(X[a0:a1], cb, po.ARITH_NATIVE_AVX2)
            oi, od = po.pq_flat_search(Q, cb, codes, k, po.ARITH_NATIVE_AVX2, ids=np.arange(a0, a1, dtype=np.int32), rerank=R, base=X[a0:a1],
                                       exact_order=1, nthreads=16)
            parts.append((od, oi))
        from vector_db_b200.sharded import merge_topk_host
        md, mi = merge_topk_host(np.stack([x[0] for x in parts]), np.stack([x[1] for x in parts]), k)
        out["pq_parity"] = {"ids_equal": float((ids.cpu().numpy() == mi).mean()),
                            "dist_bits_equal": float((d.cpu().numpy().view(np.uint32) == md.view(np.uint32)).mean())}
    p.close()

    # ---- 2. timing, C4
    if a.c4_rows > 0:
        n, dim, nq, k = a.c4_rows, 768, 4096, 100
        lo, hi = plan_shards(n, world)[rank]
        f = v.NativeFlatIndex(dim, hi - lo, id_offset=lo, device=dev)
        f.fill_synthetic(42, hi - lo)
        tq = torch.from_numpy(sy.synth(43, nq, dim)).cuda()
        s = ShardedSearcher(f, rank, world)
        ms = timed(torch, dist, lambda: s.search(tq, k))
        st = f.stats()
        out["c4"] = {"rows": n, "ms": ms, "qps": nq / ms * 1e3, "tflops": 2.0 * nq * n * dim / (ms * 1e-3) / 1e12,
                     "fallback_rank0": st["flat_fallback"]}
        f.close()

    # ---- 2b. timing, C5-shaped
    if a.c5_rows > 0:
        n, dim, M, k, R, nq = a.c5_rows, 1000, 125, 10, 1000, 1024
        lo, hi = plan_shards(n, world)[rank]
        p = v.NativePqIndex(dim, hi - lo, M=M, id_offset=lo, rerank=R, device=dev)
        p.train(sy.synth(41, 10000, dim))
        p.fill_synthetic(42, hi - lo)
        p.encode_all()
        tq = torch.from_numpy(sy.synth(43, nq, dim)).cuda()
        s = ShardedSearcher(p, rank, world)
        ms = timed(torch, dist, lambda: s.search(tq, k), iters=3, warm=1)
        out["c5"] = {"rows": n, "ms": ms, "qps": nq / ms * 1e3, "scan_gbs_aggregate": nq * n * M / (ms * 1e-3) / 1e9}
        p.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
