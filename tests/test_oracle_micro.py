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
