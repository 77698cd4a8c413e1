"""CPU suite: the C-ABI library loads and exports every symbol include/vdb.h declares (no compute calls without a
GPU), the product fails loudly without CUDA, and the host-side mirror of the reference interface behaves like the
reference's value types."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build_lib():
    import __graft_entry__ as ge
    ge.build_product()


def test_library_exports_every_declared_symbol():
    _build_lib()
    from vector_db_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "vdb.h")).read()
    declared = set(re.findall(r"\b(vdb_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    L = _lib.lib()
    for name in sorted(declared):
        assert hasattr(L, name), "libvectordb.so does not export %s" % name
    assert declared == set(_lib.SYMBOLS)
    assert b"sm_100a" in L.vdb_version()


def test_struct_layout_matches_header():
    import ctypes as C
    from vector_db_b200._lib import VdbConfig, VdbStats
    # ask the C compiler what include/vdb.h lays out, and compare with the ctypes mirror
    import subprocess, tempfile
    src = ('#include <stdio.h>\n#include <stddef.h>\n#include "vdb.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n", sizeof(vdb_config), '
           'offsetof(vdb_config, level_seed), offsetof(vdb_config, trav_store), sizeof(vdb_stats), offsetof(vdb_stats, last_kernel_ms));return 0;}')
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")])
        want = [int(x) for x in subprocess.check_output([os.path.join(d, "t")]).split()]
    got = [C.sizeof(VdbConfig), VdbConfig.level_seed.offset, VdbConfig.trav_store.offset, C.sizeof(VdbStats), VdbStats.last_kernel_ms.offset]
    assert got == want, (got, want)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _build_lib()
    import vector_db_b200 as v
    with pytest.raises(v.VdbError) as e:
        v.NativeHnswPqIndex(64, 1000)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "vector_db_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in src and "liboracle" not in src and "import oracle" not in src, f


def test_compression_config_mirror():
    from vector_db_b200 import CompressionConfig
    # RecallOptimizationTest.java:53-73: sub-dim >= 4, ratio in [4, 64]
    c = CompressionConfig.recommendedConfig(512)
    assert c.pqSubspaces == 64 and c.pqBits == 8 and c.enabled
    assert 512 // c.pqSubspaces >= 4 and 4 <= c.getCompressionRatio(512) <= 64
    assert c.getCompressionRatio(512) == 32.0                       # README.md:39
    assert CompressionConfig.recommendedConfig(1000).pqSubspaces == 125
    assert CompressionConfig.recommendedConfig(768).pqSubspaces == 96
    assert CompressionConfig.recommendedConfig(128).pqSubspaces == 16
    assert CompressionConfig.highRecallConfig(128).pqSubspaces == 32
    assert CompressionConfig.defaultConfig().getCompressionRatio(512) == 1.0


def test_search_result_mirror():
    from vector_db_b200 import SearchResult
    # SearchResultTest.java: similarity = round(1/(1+0.5 d) * 1e4) / 1e4, ordering by distance
    assert SearchResult(1, 0.0).similarity == 1.0
    assert SearchResult(1, 2.0).similarity == 0.5
    r = sorted([SearchResult(1, 0.5), SearchResult(2, 0.1)], key=lambda s: s.distance)
    assert [x.id for x in r] == [2, 1]


def test_shard_plan_and_merge_rule():
    from vector_db_b200.sharded import merge_topk_host, plan_shards
    assert plan_shards(10, 3) == [(0, 4), (4, 7), (7, 10)]
    assert plan_shards(8, 8)[-1] == (7, 8)
    d = np.array([[[0.5, 0.7, np.inf]], [[0.5, 0.6, 0.9]]], dtype=np.float32)       # [g=2][nq=1][k=3]
    i = np.array([[[7, 1, -1]], [[3, 9, 4]]], dtype=np.int32)
    od, oi = merge_topk_host(d, i, 3)
    assert list(oi[0]) == [3, 7, 9] and list(od[0]) == [0.5, 0.5, np.float32(0.6)]     # ties broken by id


def test_product_synthetic_generator_matches_oracle():
    """vector_db_b200.synth (host twin of fill_synth_kernel, used by bench.py / tools for inputs) == the oracle's generator."""
    from oracle import pyoracle as po
    from vector_db_b200 import synth as sy
    a, b = po.synth(42, 777, 96), sy.synth(42, 777, 96)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    a, b = po.fill_uniform(43, 123456789, 4097), sy.fill_uniform(43, 123456789, 4097, chunk=1000)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert a.min() >= -1.0 and a.max() < 1.0
