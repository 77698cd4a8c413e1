"""Generates tests/golden/*.npz. Run HERE (container with /root/reference): python tests/golden/make_golden.py

ref_native.npz   outputs of the UNMODIFIED reference compiled into oracle/_ref (native/index/PQIndex.cpp,
                 native/index/HNSWPQIndex.cpp, native/compute/DistanceAVX2.cpp) on counter-based synthetic inputs.
                 These pin the oracle restatement and the CUDA path on the GPU box, where /root/reference is absent.
java_oracle.npz  outputs of the oracle's restatement of the Java classes (no JVM exists here, so this file is a
                 regression pin of the restatement, not a pin against Java — see DESIGN.md "parity pinning").
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def ref_native():
    assert po.ref_available(), "run `make -C oracle ref` first"
    R = po.ref()
    n, dim, M, nq, k = 1500, 64, 8, 16, 10
    X = po.synth(42, n, dim)
    Q = po.synth(43, nq, dim)
    ids = (np.arange(n, dtype=np.int32) * 7 + 3) % 20011          # non-monotonic external ids
    g = {}
    # distances: R1 (euclideanDistanceAVX2), R2 (batchEuclideanDistanceAVX2), scalar, native cosine
    dims = [3, 8, 13, 64]
    for d in dims:
        a, b = po.synth(7, 20, d), po.synth(8, 20, d)
        g["r1_%d" % d] = np.array([R.ref_l2sq_avx2(a[i], b[i], d) for i in range(20)], dtype=np.float32)
        g["seq_%d" % d] = np.array([R.ref_l2sq_scalar(a[i], b[i], d) for i in range(20)], dtype=np.float32)
        out = np.empty(20, np.float32)
        for i in range(20):
            o1 = np.empty(1, np.float32); R.ref_batch_l2(a[i], b[i], 1, d, o1); out[i] = o1[0]
        g["r2_%d" % d] = out
        g["ncos_%d" % d] = np.array([R.ref_cosine_avx2(a[i], b[i], d) for i in range(20)], dtype=np.float32)
    # PQIndex: train (flavour 2), addBatch (encode, R1), searchBatch (LUT R2, ADC R3, (dist,id) top-k)
    pq = po.RefPQ(dim, 4096, M)
    pq.train(X[:1000])
    pq.add_batch(X, ids)
    g["pq_codebooks"] = pq.codebooks()
    g["pq_codes"] = pq.codes()
    g["pq_ids"], g["pq_dists"] = pq.search_batch(Q, k)
    pq.close()
    # HNSWPQIndex: train (flavour 1) + encode
    hp = po.RefHnswPq(dim, 4096, M)
    hp.train(X[:1000])
    g["hnswpq_codebooks"] = hp.codebooks()
    g["hnswpq_codes"] = hp.encode(X[:256])
    hp.close()
    g["meta"] = np.array([n, dim, M, nq, k, 1000], dtype=np.int64)
    g["ids"] = ids
    np.savez_compressed(os.path.join(OUT, "ref_native.npz"), **g)
    print("ref_native.npz", {k_: v.shape for k_, v in g.items()})


def java_oracle():
    g = {}
    # HnswPqIndex: incremental add (train fires at min(10000, maxElements/10) = 150 samples), then search
    n, dim, M, nq, k = 600, 32, 4, 12, 10
    X = po.synth(42, n, dim)
    Q = po.synth(43, nq, dim)
    h = po.OracleHnsw(po.KIND_HNSWPQ, dim, 1500, pq_m=M, level_seed=99)
    h.add_batch(X)
    g["hp_codebooks"] = h.codebooks()
    g["hp_codes"] = h.codes()
    g["hp_levels"], adj, entry = h.export_graph()
    g["hp_adj"] = np.stack(adj)
    g["hp_entry"] = np.array([entry], dtype=np.int32)
    g["hp_ids"], g["hp_dists"], st = h.search(Q, k)
    g["hp_stats"] = st
    h.rebuild()
    g["hp_rebuilt_ids"], g["hp_rebuilt_dists"], _ = h.search(Q, k)
    g["hp_rebuilt_codes"] = h.codes()
    h.close()
    # HnswIndex (cosine)
    c = po.OracleHnsw(po.KIND_HNSW, dim, 1500, level_seed=99)
    c.add_batch(X)
    g["hc_levels"], adj, entry = c.export_graph()
    g["hc_adj"] = np.stack(adj)
    g["hc_entry"] = np.array([entry], dtype=np.int32)
    g["hc_norms"] = c.norms()
    g["hc_ids"], g["hc_dists"], _ = c.search(Q, k)
    c.close()
    # Java-flavour PQ training on 300 samples
    g["java_codebooks"] = po.pq_train(X[:300], M, 25, po.TRAIN_JAVA)
    g["meta"] = np.array([n, dim, M, nq, k, 1500], dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "java_oracle.npz"), **g)
    print("java_oracle.npz", {k_: v.shape for k_, v in g.items()})


if __name__ == "__main__":
    ref_native()
    java_oracle()
