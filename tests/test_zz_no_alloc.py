"""vdb_flat_search_noalloc_device (include/vdb.h, csrc/noalloc.cu): the order-exact flat search with its scratch in a fixed
static __device__ array — no counterpart in the reference, added for callers that
cannot call cudaMalloc. Checked through the C ABI against the existing entry point (vdb_search_batch_device on a
VDB_INDEX_FLAT index over the same rows) to the bar the flat tests use: ids equal, distance bits equal. Two calls in a row
with different shapes, so the second one finds the first one's contents in the shared scratch.

This file is named test_zz_* so that pytest collects it LAST: a failure here must
not hide the rest of the suite. It passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def v():
    import __graft_entry__ as ge
    ge.build_product()
    import vector_db_b200
    return vector_db_b200


def _noalloc(v, torch, tx, tq, k, order):
    from vector_db_b200 import _lib
    nq = tq.shape[0]
    ids = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    d = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = _lib.lib().vdb_flat_search_noalloc_device(tx.data_ptr(), tx.shape[0], tx.shape[1], order, tq.data_ptr(), nq, k,
                                                   ids.data_ptr(), d.data_ptr(), st)
    _lib.check(rc)
    torch.cuda.synchronize()
    return ids.cpu().numpy(), d.cpu().numpy()


def _existing(v, X, Q, k, order):
    f = v.NativeFlatIndex(X.shape[1], X.shape[0], arith=order)
    f.addBatch(np.arange(X.shape[0], dtype=np.int32), X)
    ids, d = f.searchBatch(Q, k)
    f.close()
    return ids, d


def test_noalloc_flat_search_matches_existing_entry_point_twice(v, po):
    import torch
    from vector_db_b200 import _lib
    from vector_db_b200 import synth as sy
    torch.cuda.set_device(0)
    # call 1: several chunks per query (tensor-core path on the index side: the answers must still be identical)
    X1, Q1 = sy.synth(42, 20000, 96), sy.synth(43, 50, 96)
    tx1, tq1 = torch.from_numpy(X1).cuda(), torch.from_numpy(Q1).cuda()
    i1, d1 = _noalloc(v, torch, tx1, tq1, 10, 0)
    e1, ed1 = _existing(v, X1, Q1, 10, 0)
    assert np.array_equal(i1, e1) and np.array_equal(bits(d1), bits(ed1))
    # call 2, straight after, other shape / k / reduction order: the scratch still holds call 1's lists
    X2, Q2 = sy.synth(44, 3000, 40), sy.synth(45, 7, 40)
    X2[100] = X2[17]                                    # an exact tie: broken by id
    tx2, tq2 = torch.from_numpy(X2).cuda(), torch.from_numpy(Q2).cuda()
    i2, d2 = _noalloc(v, torch, tx2, tq2, 33, 1)
    e2, ed2 = _existing(v, X2, Q2, 33, 1)
    assert np.array_equal(i2, e2) and np.array_equal(bits(d2), bits(ed2))
    # and the first request again: same answer as before
    i3, d3 = _noalloc(v, torch, tx1, tq1, 10, 0)
    assert np.array_equal(i3, i1) and np.array_equal(bits(d3), bits(d1))
    # the CPU oracle agrees too (the checker, not the thing under test)
    oi, od = po.flat_search(X1, Q1, 10, nthreads=4)
    assert np.array_equal(i1, oi) and np.array_equal(bits(d1), bits(od))
    # k > n: the tail is padded with id -1 / +inf like every other entry point
    X4 = sy.synth(46, 5, 16)
    tx4, tq4 = torch.from_numpy(X4).cuda(), torch.from_numpy(sy.synth(47, 2, 16)).cuda()
    i4, d4 = _noalloc(v, torch, tx4, tq4, 8, 0)
    assert (i4[:, 5:] == -1).all() and np.isinf(d4[:, 5:]).all() and (np.sort(i4[:, :5], axis=1) == np.arange(5)).all()
    # beyond the fixed scratch: a clear error, nothing launched
    big = torch.empty((40000, 16), dtype=torch.float32, device="cuda")
    out_i = torch.empty((40000, 16), dtype=torch.int32, device="cuda")
    out_d = torch.empty((40000, 16), dtype=torch.float32, device="cuda")
    rc = _lib.lib().vdb_flat_search_noalloc_device(tx4.data_ptr(), 5, 16, 0, big.data_ptr(), 40000, 16, out_i.data_ptr(), out_d.data_ptr(), None)
    assert rc == 1 and b"scratch" in _lib.lib().vdb_last_error()
    rc = _lib.lib().vdb_flat_search_noalloc_device(tx4.data_ptr(), 5, 16, 0, tq4.data_ptr(), 2, 2000, out_i.data_ptr(), out_d.data_ptr(), None)
    assert rc == 1 and b"limits" in _lib.lib().vdb_last_error()
