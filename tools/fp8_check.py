#!/usr/bin/env python
"""tools/fp8_check.py — recall / id agreement / kernel time of the fp8 and int8 traversal rows (trav_store = 2 / 3) against the bf16
default and fp32 rows on the SAME graph, codebooks and codes. Ground truth: the exact flat search of the library.
    python tools/fp8_check.py [--n 100000] [--dim 512] [--nq 1184] [--k 10] [--pqm 64]
Prints one JSON line."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def recall(ids, gt):
    return float(np.mean([len(set(a.tolist()) & set(b.tolist())) / len(b) for a, b in zip(ids, gt)]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--nq", type=int, default=1184)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--pqm", type=int, default=64)
    ap.add_argument("--skip-fp32", action="store_true")
    a = ap.parse_args()
    import torch
    from vector_db_b200 import synth as sy
    import vector_db_b200 as v
    X, Q = sy.synth(42, a.n, a.dim), sy.synth(43, a.nq, a.dim)
    ids_all = np.arange(a.n, dtype=np.int32)
    base = v.NativeHnswPqIndex(a.dim, 2 * a.n, pqM=a.pqm, level_seed=1234)
    base.addBatch(ids_all, X)
    t0 = time.time(); base.buildIndex(); build_s = time.time() - t0
    levels, adj, entry = base.graph()
    cb, codes = base.codebooks(), base.codes()
    f = v.NativeFlatIndex(a.dim, a.n)
    f.addBatch(ids_all, X)
    gt, _ = f.searchBatch(Q, a.k)
    f.close()
    tq = torch.from_numpy(Q).cuda()
    out = {"n": a.n, "dim": a.dim, "nq": a.nq, "k": a.k, "build_s": build_s}
    res = {}
    modes = [("bf16", 0), ("fp8", 2), ("int8", 3)] + ([] if a.skip_fp32 else [("fp32", 1)])
    for name, store in modes:
        if store == 0:
            g = base
        else:
            g = v.NativeHnswPqIndex(a.dim, 2 * a.n, pqM=a.pqm, level_seed=1234, trav_store=store)
            g.addBatch(ids_all, X)
            g.set_codebooks(cb)
            g.set_graph(levels, adj, entry)
            g.set_codes(codes)
        ids, d = g.searchBatch(Q, a.k)
        ms = []
        for _ in range(3):
            g.search_device(tq, a.k); torch.cuda.synchronize()
            ms.append(g.stats()["last_kernel_ms"])
        st = g.stats()
        res[name] = ids
        out[name] = {"recall": recall(ids, gt), "kernel_ms": min(ms), "dist_evals_per_query": st["dist_evals"] / st["queries"]}
        if store != 0:
            g.close()
    for m in ("fp8", "int8"):
        out[m]["id_match_vs_bf16"] = float((res[m] == res["bf16"]).mean())
    if "fp32" in res:
        for m in ("fp8", "int8", "bf16"):
            out[m]["id_match_vs_fp32"] = float((res[m] == res["fp32"]).mean())
    base.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
