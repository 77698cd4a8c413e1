#!/usr/bin/env python
"""tools/c5_shard_prof.py [rows] — one rank's share of the sharded C5 pipeline (default 1.25 M rows = 10 M / 8 GPUs):
LUT build -> ADC scan -> top-1000 -> fp32 rerank -> k = 10, 1024 queries. Run under
`ncu --metrics gpu__time_duration.sum` for the per-kernel split; prints the batch time."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vector_db_b200 import synth as sy
import vector_db_b200 as v
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1250000
dim, M, k, R, nq = 1000, 125, 10, 1000, 1024
p = v.NativePqIndex(dim, rows, M=M, rerank=R)
p.train(sy.synth(41, 10000, dim))
p.fill_synthetic(42, rows)
p.encode_all()
tq = torch.from_numpy(sy.synth(43, nq, dim)).cuda()
p.search_device(tq, k); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); p.search_device(tq, k); e1.record(); torch.cuda.synchronize()
print(json.dumps({"rows": rows, "ms_per_batch": e0.elapsed_time(e1), "scan_kernel_ms": p.stats()["last_kernel_ms"]}))
