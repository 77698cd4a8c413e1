#!/usr/bin/env python
"""tools/native_hnswpq_baseline.py — BASELINE.md §3 B1, the part that is NOT in the default bench run: the reference's native
HNSWPQIndex (native/index/HNSWPQIndex.cpp, compiled unmodified into oracle/_ref) at the C1 shape: train, 10 000 sequential adds
(its BFS-based insert takes ~30 s on 8 cores), search / searchBatch (hard-wired to 4 threads, HNSWPQIndex.cpp:782-808), Recall@10
against the exact sq-L2 ranking. CPU only. Its recall is near zero on this data (SURVEY.md finding 6), which is why the CPU arm of
the metric is the restatement of the JAVA HnswPqIndex and not this class."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po

n, dim, nq, k = 10000, 512, 1000, 10
X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
r = po.RefHnswPq(dim, 2 * n, 64)
t = time.time(); r.train(X[:2000]); train_s = time.time() - t
t = time.time()
for i in range(n):
    r.add(i, X[i])
add_s = time.time() - t
r.search_batch(Q[:50], k)
t = time.time(); ids, d = r.search_batch(Q, k); batch_s = time.time() - t
gt, _ = po.flat_search(X, Q, k, nthreads=os.cpu_count())
rec = float(np.mean([len(set(a) & set(b)) / k for a, b in zip(ids, gt)]))
print(json.dumps({"what": "native HNSWPQIndex (unmodified reference) at C1: 10k x 512-d, PQ 64x8, 1000 queries, k=10", "cores": os.cpu_count(),
                  "train_s": train_s, "add_s": add_s, "searchBatch_qps_4_threads": nq / batch_s, "recall_at_10": rec}))
