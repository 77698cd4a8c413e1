"""Pins the Java half of the oracle (oracle/oracle.cpp restates index/HnswPqIndex.java and index/HnswIndex.java; no JVM
exists on this image) on things that do NOT come from the restatement itself:
  1. hand-derived traces on 5..150-node graphs (tests/golden/micro_fixtures.json), one per Java rule: the three termination
     rules, the stale / live `furthest`, the first-improving-neighbour `break`, the rerank / sort / cut tail;
  2. an independent pure-Python restatement of the same Java methods written from the Java source (heapq + dicts, below),
     compared with the C++ one on random small graphs — two readings of the Java must agree on ids, order and work;
  3. java.util.Random: a Python restatement of the javadoc algorithm anchored on widely published values, then compared
     with the oracle's LCG over long streams (nextInt(), nextInt(bound) for power-of-two / odd / rejection-prone bounds,
     nextFloat());
  4. the direct PQ distance (a3, order R5) against the COMPILED reference's HNSWPQIndex::computeDistancePQ
     (native/index/HNSWPQIndex.cpp:269-289, oracle/_ref) on the reference's own trained codebooks and codes."""
import heapq
import json
import math
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
FX = json.load(open(os.path.join(HERE, "golden", "micro_fixtures.json")))


def _vectors(fx):
    if "positions" in fx:
        return np.array([[p, 0, 0, 0] for p in fx["positions"]], dtype=np.float32), np.zeros(4, dtype=np.float32)
    a = np.radians(np.array(fx["angles_deg"], dtype=np.float64))
    X = np.stack([np.cos(a), np.sin(a), np.zeros_like(a), np.zeros_like(a)], axis=1).astype(np.float32)
    return X, np.array([1, 0, 0, 0], dtype=np.float32)


def _oracle_with_graph(po, kind, X, adj, entry, level=0, max_elements=1100):
    """max_elements 1100 -> maxLevel = int(ln(1100)/ln(32)) + 1 = 3: the fixture graph is placed on `level`"""
    n = X.shape[0]
    o = po.OracleHnsw(kind, 4, max_elements, pq_m=2 if kind == po.KIND_HNSWPQ else 0)
    L = o.max_level
    A = np.full((L, n, 32), -1, dtype=np.int32)
    for i, nb in enumerate(adj):
        A[level, i, :len(nb)] = nb
    o.import_state(X, np.full(n, L - 1, dtype=np.int32), A, entry)
    return o


@pytest.mark.parametrize("fx", FX["fixtures"], ids=[f["name"] for f in FX["fixtures"]])
def test_termination_rules_hand_traces(po, fx):
    X, q = _vectors(fx)
    o = _oracle_with_graph(po, fx["kind"], X, fx["adj"], fx["entry"], fx["level"])
    idx, d, st = o.search_layer(q, fx["entry"], fx["ef"], fx["level"], fx["variant"])
    assert list(idx) == fx["expect_idx"], fx["trace"]
    assert int(st[0]) == fx["expect_evals"], fx["trace"]
    assert (np.diff(d) >= 0).all()


@pytest.mark.parametrize("fx", FX["greedy"], ids=[f["name"] for f in FX["greedy"]])
def test_first_improving_neighbour_hand_traces(po, fx):
    X, q = _vectors(fx)
    o = _oracle_with_graph(po, fx["kind"], X, fx["adj"], fx["entry"], fx["level"])
    node, st = o.search_closest(q, fx["entry"], fx["level"])
    assert node == fx["expect_node"] and int(st[0]) == fx["expect_evals"], fx["trace"]


def test_search_nearest_rerank_tail_and_padding(po):
    fx = FX["search_nearest"][0]
    X, q = _vectors(fx)
    o = _oracle_with_graph(po, po.KIND_HNSWPQ, X, fx["adj"], fx["entry"], 0)
    ids, d, _ = o.search(q.reshape(1, -1), 3)
    assert list(ids[0]) == fx["k3"]["ids"] and list(d[0]) == fx["k3"]["dists"]
    ids, d, _ = o.search(q.reshape(1, -1), 10)
    assert list(ids[0]) == fx["k10"]["ids"] and list(d[0][:6]) == fx["k10"]["dists"] and np.isinf(d[0][6:]).all()


def test_search_nearest_line_graph_break_at_pool_100(po):
    fx = FX["search_nearest"][1]
    n = fx["n"]
    X = np.array([[i + 1, 0, 0, 0] for i in range(n)], dtype=np.float32)
    adj = [[j for j in (i + 1, i - 1) if 0 <= j < n] for i in range(n)]
    adj[149] = [148, 10]
    o = _oracle_with_graph(po, po.KIND_HNSWPQ, X, adj, fx["entry"], 0)
    # untrained index -> searchLayer (variant 1, `continue`); the fixture is about searchLayerWithSize (trained): give it codebooks
    cb = np.zeros((2, 256, 2), dtype=np.float32)
    o.import_state(X, np.zeros(n, dtype=np.int32), _adj_table(o, adj), fx["entry"], codebooks=cb, codes=np.zeros((n, 2), dtype=np.uint8))
    ids, d, st = o.search(np.zeros((1, 4), dtype=np.float32), 1)
    assert list(ids[0]) == fx["k1"]["ids"] and list(d[0]) == fx["k1"]["dists"]
    assert int(st[0]) == fx["expect_evals"], fx["trace"]


def _adj_table(o, adj, level=0):
    A = np.full((o.max_level, len(adj), 32), -1, dtype=np.int32)
    for i, nb in enumerate(adj):
        A[level, i, :len(nb)] = nb
    return A


# ---------------------------------------------------------------------------------------------------------------------
# 2. independent Python restatement of the Java methods (written from the Java source; float32 arithmetic through numpy)
# ---------------------------------------------------------------------------------------------------------------------
F = np.float32


def _l2sq_java(a, b):                      # HnswPqIndex.java:741-750: sequential sum += diff * diff
    s = F(0)
    for x, y in zip(a, b):
        df = F(x) - F(y)
        s = F(s + F(df * df))
    return s


class PyJavaHnswPq:
    """searchLayer / searchLayerWithSize / searchLayerClosest / searchNearest of index/HnswPqIndex.java on a given graph
    (untrained index = exact distances everywhere). Heap ties by (distance, id) like the C++ oracle states it does."""

    def __init__(self, X, adj_levels, entry, ef=64):
        self.X, self.adj, self.entry, self.ef = X, adj_levels, entry, ef
        self.max_level = len(adj_levels)
        self.evals = 0

    def dist(self, i, q):
        self.evals += 1
        return _l2sq_java(self.X[i], q)

    def closest(self, q, ep, level):       # :761-792
        cur, curd = ep, self.dist(ep, q)
        changed = True
        while changed:
            changed = False
            nb = self.adj[level].get(cur)
            if not nb:
                return cur
            for n in nb:
                d = self.dist(n, q)
                if d < curd:
                    curd, cur, changed = d, n, True
                    break
        return cur

    def layer(self, q, ep, ef, level, with_size):      # :829-895 (with_size False) / :905-974 (True)
        d0 = self.dist(ep, q)
        results = [(-d0, -ep)]             # max-heap through negation; ties: larger id is "further"
        cands = [(d0, ep)]
        visited = {ep}
        furthest = d0
        while cands:
            cd, c = heapq.heappop(cands)
            if cd > furthest and len(results) >= ef:
                if with_size:
                    break
                continue
            nb = self.adj[level].get(c)
            if not nb:
                continue
            for n in nb:
                if n in visited:
                    continue
                visited.add(n)
                d = self.dist(n, q)
                if len(results) < ef or d < furthest:
                    heapq.heappush(cands, (d, n))
                    heapq.heappush(results, (-d, -n))
                    if with_size:
                        if len(results) > ef:
                            heapq.heappop(results)
                        furthest = -results[0][0]
                    elif len(results) > ef:
                        heapq.heappop(results)
                        furthest = -results[0][0]
        return sorted((-a, -b) for a, b in results)

    def search(self, q, k):                # :599-669, untrained branch
        n = len(self.X)
        ef = min(min(max(max(int(n * 0.15), k * 100), self.ef), n), 5000)
        cur = self.entry
        for level in range(self.max_level - 1, 0, -1):
            cur = self.closest(q, cur, level)
        c = self.layer(q, cur, ef, 0, False)
        c = sorted((_l2sq_java(self.X[i], q), i) for _, i in c)
        return c[:k]


def _random_graph(rng, n, L, deg):
    levels = np.minimum(rng.geometric(0.6, n) - 1, L - 1)
    levels[0] = L - 1
    adj = []
    for l in range(L):
        members = [i for i in range(n) if levels[i] >= l]
        d = {}
        for i in members:
            others = [j for j in members if j != i]
            d[i] = list(rng.choice(others, size=min(deg, len(others)), replace=False)) if others else []
        adj.append(d)
    return levels.astype(np.int32), adj


@pytest.mark.parametrize("seed", range(6))
def test_cpp_oracle_equals_independent_python_restatement(po, seed):
    rng = np.random.default_rng(seed)
    n, dim, deg = 90 + 10 * seed, 6, 4 + seed % 3
    X = rng.uniform(-1, 1, (n, dim)).astype(np.float32)
    o = po.OracleHnsw(po.KIND_HNSWPQ, dim, 1100, pq_m=2)
    L = o.max_level
    levels, adj = _random_graph(rng, n, L, deg)
    A = np.full((L, n, 32), -1, dtype=np.int32)
    for l in range(L):
        for i, nb in adj[l].items():
            A[l, i, :len(nb)] = nb
    o.import_state(X, levels, A, 0)
    py = PyJavaHnswPq(X, adj, 0)
    Q = rng.uniform(-1, 1, (8, dim)).astype(np.float32)
    for q in Q:
        for ef in (3, 10, 200):
            for variant, with_size in ((0, True), (1, False)):
                py.evals = 0
                want = py.layer(q, 5, ef, 0, with_size)
                idx, d, st = o.search_layer(q, 5, ef, 0, variant)
                assert [i for _, i in want] == list(idx) and [float(x) for x, _ in want] == [float(x) for x in d]
                assert py.evals == int(st[0])
        py.evals = 0
        want = py.search(q, 5)
        ids, d, st = o.search(q.reshape(1, -1), 5)
        assert [i for _, i in want] == [int(i) for i in ids[0] if i >= 0]
        assert [float(x) for x, _ in want] == [float(x) for x, i in zip(d[0], ids[0]) if i >= 0]


# ---------------------------------------------------------------------------------------------------------------------
# 3. java.util.Random — javadoc algorithm restated in Python ints, anchored on published values
# ---------------------------------------------------------------------------------------------------------------------
class PyJavaRandom:
    MASK = (1 << 48) - 1

    def __init__(self, seed):
        self.seed = (seed ^ 0x5DEECE66D) & self.MASK

    def next(self, bits):
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & self.MASK
        v = self.seed >> (48 - bits)
        return v - (1 << bits) if v >= (1 << (bits - 1)) and bits == 32 else v     # (int) cast: only next(32) can be negative

    def nextInt(self, bound=None):
        if bound is None:
            return self.next(32)
        r = self.next(31)
        m = bound - 1
        if (bound & m) == 0:
            return (bound * r) >> 31
        u = r
        while True:
            r = u % bound
            if u - r + m < (1 << 31):      # Java: `u - r + m < 0` detects int overflow
                return r
            u = self.next(31)

    def nextFloat(self):
        return np.float32(self.next(24)) / np.float32(1 << 24)


def test_java_random_anchored_and_long_streams(po):
    a = FX["java_random_anchors"]
    r0 = PyJavaRandom(0)
    assert [r0.nextInt(), r0.nextInt()] == a["seed0_nextInt"]            # the classic `new Random(0)` values
    assert PyJavaRandom(42).nextInt() == a["seed42_nextInt_first"]
    assert PyJavaRandom(42).nextInt(10) == a["seed42_nextInt10_first"]
    L = po.lib()
    n = 4000
    ints, fl = np.empty(n, np.int32), np.empty(n, np.float32)
    for seed in (0, 42, 42 + 63, -7, 2 ** 40 + 12345):
        L.orc_java_random_stream(seed, 0, 0, n, ints, fl)
        r = PyJavaRandom(seed)
        assert list(ints) == [r.nextInt() for _ in range(n)]
        # bounds: power of two, small odd, the k-means++ sample counts, and one just above 2^30 (rejection loop taken often)
        for bound in (1, 16, 1 << 30, 10, 2000, 10000, 5000, (1 << 30) + 1, 2 ** 31 - 1):
            L.orc_java_random_stream(seed, 1, bound, n, ints, fl)
            r = PyJavaRandom(seed)
            assert list(ints) == [r.nextInt(bound) for _ in range(n)], (seed, bound)
        L.orc_java_random_stream(seed, 2, 0, n, ints, fl)
        r = PyJavaRandom(seed)
        assert [float(x) for x in fl] == [float(r.nextFloat()) for _ in range(n)]
    # literal values (from the anchored restatement above) kept in the test so a regression in BOTH is visible
    L.orc_java_random_stream(42, 1, 256, 12, ints, fl)
    assert list(ints[:12]) == [PyJavaRandom42_256[i] for i in range(12)]


_r = PyJavaRandom(42)
PyJavaRandom42_256 = [_r.nextInt(256) for _ in range(12)]


# ---------------------------------------------------------------------------------------------------------------------
# 4. direct PQ distance (a3, R5) against the compiled reference
# ---------------------------------------------------------------------------------------------------------------------
def test_pq_direct_distance_equals_compiled_reference(po):
    if not po.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine and no prebuilt library)")
    n, dim, M = 600, 64, 8
    X, Q = po.synth(42, n, dim), po.synth(43, 20, dim)
    r = po.RefHnswPq(dim, 2 * n, M)
    r.train(X[:400])
    cb = r.codebooks()
    codes = r.encode(X)
    for i in range(n):
        r.add(i, X[i])
    R = po.ref()
    for q in Q:
        for node in range(0, n, 7):
            want = R.ref_hnswpq_pq_distance(r.h, np.ascontiguousarray(q), node)           # HNSWPQIndex::computeDistancePQ
            got = po.pq_direct(q, codes[node], cb)
            assert np.float32(want).view(np.uint32) == np.float32(got).view(np.uint32), (node, want, got)
    r.close()
