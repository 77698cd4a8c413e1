#!/usr/bin/env python
"""tools/c3_idmatch.py — where do the C3 id differences against the Java-order oracle come from? (GPU box only)
Builds the C3 index once, then searches the same graph / codebooks / codes with different traversal-row formats and round
batches; prints one JSON line per variant: id match, recall, work, kernel ms.
    python tools/c3_idmatch.py [--n 100000] [--nq 256]"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--M", type=int, default=64)
    ap.add_argument("--nq", type=int, default=256)
    ap.add_argument("--timing-nq", type=int, default=4096)
    a = ap.parse_args()
    import torch
    from oracle import pyoracle as po
    import vector_db_b200 as v
    n, dim, M, nq, k = a.n, a.dim, a.M, a.nq, 10
    cores = os.cpu_count() or 8
    X, Q = po.synth(42, n, dim), po.synth(43, max(nq, a.timing_nq), dim)
    t0 = time.time()
    g = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=1234)
    g.addBatch(np.arange(n, dtype=np.int32), X)
    g.buildIndex()
    print(json.dumps({"build_s": time.time() - t0}), flush=True)
    levels, adj, entry = g.graph()
    cb, codes = g.codebooks(), g.codes()
    g.close()
    o = po.OracleHnsw(po.KIND_HNSWPQ, dim, 2 * n, pq_m=M, level_seed=1234)
    o.import_state(X, levels, adj, entry, codebooks=cb, codes=codes)
    jid, jd, jst = o.search(Q[:nq], k, nthreads=cores)
    gt, _ = po.flat_search(X, Q[:nq], k, nthreads=cores)
    rec = lambda ids: float(np.mean([len(set(x) & set(y)) / k for x, y in zip(ids, gt)]))
    print(json.dumps({"oracle_recall": rec(jid), "oracle_evals_per_q": float(jst[0]) / nq}), flush=True)
    tq = torch.from_numpy(Q[:a.timing_nq]).cuda()
    variants = [("default", {}), ("bf16_b1", dict(trav_batch=1)), ("bf16_b8", dict(trav_batch=8)), ("bf16_b16", dict(trav_batch=16)),
                ("bf16_b32", dict(trav_batch=32)), ("f32_b64", dict(trav_store=1, trav_batch=64)), ("f32_b1", dict(trav_store=1, trav_batch=1)),
                ("int8", dict(trav_store=3)), ("int8_b32", dict(trav_store=3, trav_batch=32))]
    for name, kw in variants:
        s = v.NativeHnswPqIndex(dim, 2 * n, pqM=M, level_seed=1234, **kw)
        s.addBatch(np.arange(n, dtype=np.int32), X)
        s.set_codebooks(cb); s.set_graph(levels, list(adj), entry); s.set_codes(codes)
        ids, d = s.searchBatch(Q[:nq], k)
        st0 = s.stats()
        ms = []
        if not (kw.get("trav_batch") == 1 and a.timing_nq > 1024):
            for _ in range(3):
                s.search_device(tq, k); torch.cuda.synchronize(); ms.append(s.stats()["last_kernel_ms"])
        diffq = np.where((ids != jid).any(axis=1))[0]
        out = {"variant": name, "id_match": float((ids == jid).mean()), "recall": rec(ids), "evals_per_q": st0["dist_evals"] / nq,
               "queries_differing": int(diffq.size), "same_set": int(sum(1 for q in diffq if set(ids[q]) == set(jid[q]))),
               "gpu_better_kth": int(sum(1 for q in diffq if d[q, k - 1] < jd[q, k - 1])),
               "gpu_worse_kth": int(sum(1 for q in diffq if d[q, k - 1] > jd[q, k - 1])),
               "kernel_ms_%dq" % a.timing_nq: min(ms) if ms else None}
        print(json.dumps(out), flush=True)
        s.close()


if __name__ == "__main__":
    main()
