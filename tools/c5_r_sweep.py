import sys, json
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from vector_db_b200 import synth as sy
import vector_db_b200 as v
rows, dim, M, k, nq = 1_250_000, 1000, 125, 10, 1024
tq = torch.from_numpy(sy.synth(43, nq, dim)).cuda()
cb = None
for R in (1000, 500, 300, 125, 0):
    p = v.NativePqIndex(dim, rows, M=M, rerank=R)
    if cb is None:
        p.train(sy.synth(41, 10000, dim)); cb = p.codebooks()
    else:
        p.set_codebooks(cb)
    p.fill_synthetic(42, rows); p.encode_all()
    for _ in range(2):
        p.search_device(tq, k); torch.cuda.synchronize()
    ms = []
    for _ in range(3):
        p.search_device(tq, k); torch.cuda.synchronize(); ms.append(p.stats()["last_kernel_ms"])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); p.search_device(tq, k); e1.record(); torch.cuda.synchronize()
    print(json.dumps({"R": R, "scan_kernel_ms": min(ms), "search_ms": e0.elapsed_time(e1)}))
    p.close()
