#!/usr/bin/env python
"""tools/prof.py — small, fixed workloads for ncu captures (one kernel family per invocation).
    python tools/prof.py hnsw [--n 100000] [--nq 592] [--snapshot gpurun_out/c3.snap]
    python tools/prof.py adc  [--rows 2000000] [--nq 148]
    python tools/prof.py flat [--n 200000] [--nq 1024]
Prints one JSON line with CUDA-event timings; under ncu only the kernels matter."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["hnsw", "adc", "flat"])
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--rows", type=int, default=2000000)
    ap.add_argument("--nq", type=int, default=592)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--snapshot", default="")
    ap.add_argument("--trav-store", type=int, default=0)
    ap.add_argument("--trav-batch", type=int, default=0)
    a = ap.parse_args()
    import torch
    from vector_db_b200 import synth as sy
    import vector_db_b200 as v
    out = {"what": a.what}
    if a.what == "hnsw":
        idx = v.NativeHnswPqIndex(a.dim, 2 * a.n, pqM=64 if a.dim == 512 else 0, level_seed=1234, trav_store=a.trav_store, trav_batch=a.trav_batch)
        if a.snapshot and os.path.exists(a.snapshot):
            idx.load(a.snapshot)
        else:
            X = sy.synth(42, a.n, a.dim)
            idx.addBatch(np.arange(a.n, dtype=np.int32), X)
            t0 = time.time(); idx.buildIndex(); out["build_s"] = time.time() - t0
            if a.snapshot:
                idx.save(a.snapshot)
        tq = torch.from_numpy(sy.synth(43, a.nq, a.dim)).cuda()
        ms = []
        for _ in range(a.iters):
            idx.search_device(tq, a.k); torch.cuda.synchronize()
            ms.append(idx.stats()["last_kernel_ms"])
        st = idx.stats()
        out.update(kernel_ms=ms, dist_evals_per_query=st["dist_evals"] / st["queries"], expansions_per_query=st["expansions"] / st["queries"])
    elif a.what == "adc":
        dim, M = 1000, 125
        p = v.NativePqIndex(dim, a.rows, M=M)
        p.train(sy.synth(41, 2000, dim))
        p.fill_synthetic(42, a.rows)
        p.encode_all()
        tq = torch.from_numpy(sy.synth(43, a.nq, dim)).cuda()
        ms = []
        for _ in range(a.iters):
            p.search_device(tq, a.k); torch.cuda.synchronize()
            ms.append(p.stats()["last_kernel_ms"])
        out.update(kernel_ms=ms, gbs=[a.nq * a.rows * M / (m * 1e-3) / 1e9 for m in ms])
    else:
        f = v.NativeFlatIndex(a.dim, a.n)
        f.fill_synthetic(42, a.n)
        tq = torch.from_numpy(sy.synth(43, a.nq, a.dim)).cuda()
        ms = []
        for _ in range(a.iters):
            f.search_device(tq, a.k); torch.cuda.synchronize()
            ms.append(f.stats()["last_total_ms"])
        out.update(total_ms=ms, tflops=[2.0 * a.nq * a.n * a.dim / (m * 1e-3) / 1e12 for m in ms])
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
