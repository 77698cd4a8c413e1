#!/usr/bin/env python
"""tools/c5_shard_prof.py [rows] [g] — ONE rank's share of the sharded pipelines at g GPUs, on one GPU, for the per-kernel split
(run under `ncu --metrics gpu__time_duration.sum --clock-control none --csv`):
  C5: rows = 10 M / g (default 1.25 M) x 1000-d, PQ 125 x 8, 1024 queries: phase 1 (LUT -> ADC scan -> local top-1000 keys),
      a local shortlist of 2R/g + 64 keys, the global R-th key emulated by the rank's own (R/g)-th key (so about R/g candidates
      survive, as on g shards of i.i.d. rows), phase 2 (mask -> fp32 rerank of the survivors -> k = 10);
  C4: 1 M / g rows x 768-d flat, 4096 queries, k = 100.
Prints the batch times (CUDA events)."""
import ctypes as C, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vector_db_b200 import synth as sy, _lib
import vector_db_b200 as v
g = int(sys.argv[2]) if len(sys.argv) > 2 else 8
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000 // g
L = _lib.lib()
out = {"g": g}

def timed(fn, it=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

dim, M, k, R, nq = 1000, 125, 10, 1000, 1024
p = v.NativePqIndex(dim, rows, M=M, rerank=R)
p.train(sy.synth(41, 10000, dim)); p.fill_synthetic(42, rows); p.encode_all()
tq = torch.from_numpy(sy.synth(43, nq, dim)).cuda()
from vector_db_b200.sharded import local_shortlist_len
Lloc = local_shortlist_len(R, g, k)                    # what a shard keeps when the index is spread over g GPUs
keys = torch.empty((nq, Lloc), dtype=torch.int64, device="cuda")
kth = torch.empty((nq,), dtype=torch.int64, device="cuda")
ids = torch.empty((nq, k), dtype=torch.int32, device="cuda"); d = torch.empty((nq, k), dtype=torch.float32, device="cuda")
def phase1():
    Rv = C.c_int32(Lloc)
    _lib.check(L.vdb_pq_shortlist_device(p._h, tq.data_ptr(), nq, k, keys.data_ptr(), C.byref(Rv), 0))

def kth_fn():
    _lib.check(L.vdb_select_kth_key_device(keys.data_ptr(), 1, nq, R // g, kth.data_ptr(), 0))

def phase2():
    _lib.check(L.vdb_pq_rerank_device(p._h, tq.data_ptr(), nq, k, kth.data_ptr(), ids.data_ptr(), d.data_ptr(), 0))

out["c5_rows_per_rank"] = rows
out["c5_local_shortlist"] = Lloc
out["c5_phase1_ms"] = timed(phase1)
out["c5_kth_ms"] = timed(kth_fn)
out["c5_phase2_ms"] = timed(phase2)
out["c5_unsharded_search_ms"] = timed(lambda: p.search_device(tq, k))
p.close()

rows4, dim4, nq4, k4 = 1_000_000 // g, 768, 4096, 100
f = v.NativeFlatIndex(dim4, rows4)
f.fill_synthetic(42, rows4)
tq4 = torch.from_numpy(sy.synth(43, nq4, dim4)).cuda()
out["c4_rows_per_rank"] = rows4
out["c4_search_ms"] = timed(lambda: f.search_device(tq4, k4), 5)
f.close()
print(json.dumps(out))
