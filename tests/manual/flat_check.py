#!/usr/bin/env python
"""tests/manual/flat_check.py — bring-up check of the tensor-core flat search against the CPU oracle (GPU box only).
    python tests/manual/flat_check.py [--n 50000] [--dim 768] [--nq 200] [--k 100] [--metric 0]"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=50000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--nq", type=int, default=200)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--metric", type=int, default=0)
    ap.add_argument("--arith", type=int, default=0)
    ap.add_argument("--big", type=int, default=0, help="also time N=big rows generated on the device (no oracle)")
    a = ap.parse_args()
    import torch
    from oracle import pyoracle as po      # checker: CPU brute force / reference-order scan
    from vector_db_b200 import synth as sy
    import vector_db_b200 as v
    X, Q = sy.synth(42, a.n, a.dim), sy.synth(43, a.nq, a.dim)
    f = v.NativeFlatIndex(a.dim, a.n, metric=a.metric, arith=a.arith)
    f.addBatch(np.arange(a.n, dtype=np.int32), X)
    t0 = time.time()
    ids, d = f.searchBatch(Q, a.k)
    t1 = time.time()
    st = f.stats()
    if a.metric == 1:
        oi, od = po.flat_search(po.java_normalize(X), po.java_normalize(Q), a.k, metric=1, nthreads=16)
    else:
        oi, od = po.flat_search(X, Q, a.k, metric=0, order=a.arith, nthreads=16)
    same_ids = float((ids == oi).mean())
    same_d = float((d.view(np.uint32) == od.view(np.uint32)).mean())
    out = dict(arith=a.arith, n=a.n, dim=a.dim, nq=a.nq, k=a.k, metric=a.metric, ids_equal=same_ids, dist_bits_equal=same_d, certified=st["flat_certified"],
               fallback=st["flat_fallback"], kernel_ms=st["last_kernel_ms"], total_ms=st["last_total_ms"], wall_s=t1 - t0)
    if same_ids < 1.0:
        bad = np.argwhere(ids != oi)[:5]
        out["first_bad"] = [[int(q), int(j), int(ids[q, j]), int(oi[q, j]), float(d[q, j]), float(od[q, j])] for q, j in bad]
    f.close()
    print(json.dumps(out), flush=True)
    if a.big:
        f = v.NativeFlatIndex(a.dim, a.big, metric=a.metric)
        f.fill_synthetic(42, a.big)
        tq = torch.from_numpy(sy.synth(43, 4096, a.dim)).cuda()
        res = []
        for it in range(3):
            f.search_device(tq, a.k); torch.cuda.synchronize()
            s = f.stats()
            res.append((s["last_kernel_ms"], s["last_total_ms"]))
        s = f.stats()
        print(json.dumps(dict(big=a.big, nq=4096, ms=res, tflops=[2.0 * 4096 * a.big * a.dim / (m[0] * 1e-3) / 1e12 for m in res],
                              certified=s["flat_certified"], fallback=s["flat_fallback"])), flush=True)
        f.close()


if __name__ == "__main__":
    main()
