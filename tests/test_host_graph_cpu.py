"""CPU test of the product's host graph builder (vector_db_b200/csrc/host_graph.cpp) against the oracle's
restatement of the Java insert / remove (index/HnswPqIndex.java:391-596, 1043-1064): in exact-order mode (strict mode of the
library) the graph is the oracle's graph LINK FOR LINK — levels, adjacency lists in order, entry point — after a build and
after a mixed remove / add script that also removes the entry point. Runs without a GPU through tests/probes/hostgraph_probe.cpp."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hg(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("probe") / "hgprobe.so")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-mavx2", "-mfma", os.path.join(ROOT, "tests", "probes", "hostgraph_probe.cpp"),
                           os.path.join(ROOT, "vector_db_b200", "csrc", "host_graph.cpp"), "-o", so])
    L = C.CDLL(so)
    L.hg_create.restype = C.c_void_p; L.hg_create.argtypes = [C.c_int] * 4 + [C.c_uint64, C.c_int]
    L.hg_insert.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int]
    L.hg_remove.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.hg_export.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    for f in ("hg_size", "hg_entry", "hg_max_level", "hg_dirty_count", "hg_clear_dirty", "hg_destroy"):
        getattr(L, f).argtypes = [C.c_void_p]
    return L


def _same(L, h, o):
    n, lm = L.hg_size(h), L.hg_max_level(h)
    lv = np.zeros(n, np.int32); adj = np.zeros((lm, n, 32), np.int32)
    L.hg_export(h, lv.ctypes.data, adj.ctypes.data)
    ol, oadj, oe = o.export_graph()
    return L.hg_entry(h) == oe and np.array_equal(lv, ol) and all(np.array_equal(adj[l], oadj[l]) for l in range(lm))


def _first_live(id2row, skip, maxel):          # ConcurrentHashMap<>(maxElements) iteration order (HnswPqIndex.java:577-584)
    table, want = 1, maxel + (maxel >> 1) + 1
    while table < want:
        table <<= 1
    return min((((i ^ (i >> 16)) & 0x7fffffff) & (table - 1), r) for i, r in id2row.items() if i != skip)[1]


@pytest.mark.parametrize("kind", ["hnswpq", "hnsw"])
def test_host_builder_graph_equals_oracle_graph_link_for_link(po, hg, kind):
    n, dim, maxel = 500, 24, 3000
    X = po.synth(42, n + 40, dim)
    if kind == "hnswpq":
        o = po.OracleHnsw(po.KIND_HNSWPQ, dim, maxel, pq_m=4, level_seed=11)
        h = hg.hg_create(dim, 32, maxel, 0, 11, 1)
        V, norms = X, np.zeros(len(X), np.float32)
    else:
        o = po.OracleHnsw(po.KIND_HNSW, dim, maxel, level_seed=11)
        h = hg.hg_create(dim, 32, maxel, 1, 11, 1)
        V = po.java_normalize_rows(X)                                   # normalised on insert (HnswIndex.java:154)
        norms = np.array([po.lib().orc_java_norm(np.ascontiguousarray(v), dim) for v in V], dtype=np.float32)
    id2row = {}

    def add(i):
        if i in id2row:
            assert not o.add(i, X[i]); return
        assert o.add(i, X[i]); id2row[i] = hg.hg_size(h)
        ef, variant = ((200 if o.trained else 64), 1) if kind == "hnswpq" else (400, 2)       # :417 / :498 / HnswIndex.java:22
        hg.hg_insert(h, V[i].ctypes.data, float(norms[i]), ef, variant)

    def rm(i):
        if i not in id2row:
            assert not o.remove(i); return
        row = id2row[i]
        ne = _first_live(id2row, i, maxel) if (row == hg.hg_entry(h) and len(id2row) > 1) else -1
        assert o.remove(i); hg.hg_remove(h, row, ne); del id2row[i]

    for i in range(n):
        add(i)
    assert _same(hg, h, o)
    hg.hg_clear_dirty(h)
    add(n)                                                               # one insert dirties its own lists + <= 32 back-links per level
    assert 0 < hg.hg_dirty_count(h) <= 3 * 33 and _same(hg, h, o)
    rng = np.random.default_rng(5)
    victims = [int(x) for x in rng.choice(n, 30, replace=False)]
    victims[3] = o.ext_ids()[o.entry]                                    # the entry point goes too
    for i in victims[:15]:
        rm(i)
    for j in range(1, 25):
        add(n + j)
    for i in victims[15:]:
        rm(i)
    rm(victims[0]); add(7 if 7 in id2row else 8); rm(n + 2); add(n + 2)
    assert _same(hg, h, o)
    assert len(id2row) == o.live
    hg.hg_destroy(h)
