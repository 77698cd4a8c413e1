#!/usr/bin/env python
"""tools/build_time.py — wall time of rebuildIndex (vdb_build) on the GPU vs the host builder, plus Recall@10 of both graphs.
    python tools/build_time.py [--n 100000] [--dim 512] [--M 64] [--host 1]"""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000); ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--M", type=int, default=64); ap.add_argument("--host", type=int, default=1); ap.add_argument("--nq", type=int, default=1024)
    a = ap.parse_args()
    from oracle import pyoracle as po
    import vector_db_b200 as v
    X, Q = po.synth(42, a.n, a.dim), po.synth(43, a.nq, a.dim)
    gt, _ = po.flat_search(X, Q, 10, nthreads=os.cpu_count())
    for mode in (["gpu", "gpu_div16", "gpu_div256", "host"] if a.host else ["gpu"]):
        for e in ("VDB_HOST_GRAPH_BUILD", "VDB_GB_NO_INTRA", "VDB_GB_DIV"): os.environ.pop(e, None)
        if mode == "host": os.environ["VDB_HOST_GRAPH_BUILD"] = "1"
        if "div16" in mode: os.environ["VDB_GB_DIV"] = "16"
        if "div256" in mode: os.environ["VDB_GB_DIV"] = "256"
        g = v.NativeHnswPqIndex(a.dim, 2 * a.n, pqM=a.M, level_seed=1234)
        g.addBatch(np.arange(a.n, dtype=np.int32), X)
        t0 = time.time(); g.buildIndex(); t1 = time.time()
        ids, _ = g.searchBatch(Q, 10)
        lv, adj, e = g.graph()
        rec = float(np.mean([len(set(x) & set(y)) / 10 for x, y in zip(ids, gt)]))
        # forward-list quality: the last 300 rows received almost no back-links, so adj[0][u] is what their insert found;
        # overlap with the exact 32 nearest rows
        last = np.arange(a.n - 300, a.n)
        tn, _ = po.flat_search(X, X[last], 33, nthreads=os.cpu_count())
        ov = float(np.mean([len(set(adj[0][u][adj[0][u] >= 0]) & set(t[1:])) / 32 for u, t in zip(last, tn)]))
        print(json.dumps({"mode": mode, "n": a.n, "dim": a.dim, "build_s": t1 - t0, "recall_at_10": rec,
                          "mean_degree_l0": float((adj[0] >= 0).sum(1).mean()), "indegree_max_l0": int(np.bincount(adj[0][adj[0] >= 0], minlength=a.n).max()),
                          "indegree_zero_l0": int((np.bincount(adj[0][adj[0] >= 0], minlength=a.n) == 0).sum()), "evals_per_q": g.stats()["dist_evals"] / a.nq, "fwd_list_overlap_with_exact_32nn": ov, "kernel_ms": g.stats()["last_kernel_ms"]}), flush=True)
        g.close()


if __name__ == "__main__":
    main()
