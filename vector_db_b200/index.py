"""Host-side mirror of the reference's index plug-in interface for the GPU path.

Reference interfaces mirrored (names, argument meaning, error behaviour):
  * `com.vectordb.index.VectorIndex`  (src/main/java/com/vectordb/index/VectorIndex.java:11-51):
        addVector, removeVector, searchNearest, size, buildIndex
  * `com.vectordb.jni.NativeIndex`    (src/main/java/com/vectordb/jni/NativeIndex.java:18-231):
        addVector(id, float[]), search(float[], k), addBatch(int[], float[][]), searchBatch(float[][], k), close
  * `NativeHnswIndex` / `NativePqIndex` constructors (jni/NativeHnswIndex.java, jni/NativePqIndex.java) and the
    `NativeHnswPqIndex` class the reference lacks (SURVEY.md §0 finding 4).
All compute happens in libvectordb.so (CUDA); this module only marshals buffers.
"""
import ctypes as C
from collections import namedtuple

import numpy as np

from . import _lib
from ._lib import (ARITH_JAVA_SEQ, ARITH_NATIVE_AVX2, INDEX_FLAT, INDEX_HNSW, INDEX_HNSW_PQ, INDEX_PQ, METRIC_COSINE,
                   METRIC_SQL2, TRAIN_JAVA, TRAIN_NATIVE_HNSWPQ, TRAIN_NATIVE_PQ, VdbConfig, VdbError, VdbStats, check)


class SearchResult(namedtuple("SearchResult", ["id", "distance"])):
    """core/SearchResult.java:6-40 — similarity = round(1/(1+0.5*d)*1e4)/1e4, ordering by distance only."""
    __slots__ = ()

    @property
    def similarity(self):
        return round(1.0 / (1.0 + 0.5 * float(self.distance)) * 10000.0) / 10000.0


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class GpuIndex:
    """One libvectordb index handle (the `nativeHandle` of NativeIndex.java:20)."""

    def __init__(self, index_type, dimension, max_elements, **overrides):
        L = _lib.lib()
        cfg = VdbConfig()
        check(L.vdb_default_config(index_type, dimension, max_elements, C.byref(cfg)))
        for k, v in overrides.items():
            if v is None:
                continue
            if not hasattr(cfg, k):
                raise TypeError("unknown vdb_config field %r" % k)
            setattr(cfg, k, v)
        h = C.c_void_p()
        check(L.vdb_create(C.byref(cfg), C.byref(h)))   # raises when no B200 is present: no CPU fallback
        self._h = h
        self.dimension = dimension
        self.max_elements = max_elements
        self.index_type = index_type
        self.pq_m = cfg.pq_m if index_type in (INDEX_PQ, INDEX_HNSW_PQ) else 0
        if self.pq_m and dimension % self.pq_m:
            from .database import CompressionConfig
            self.pq_m = CompressionConfig.recommendedConfig(dimension).pqSubspaces
        self.hnsw_m = cfg.hnsw_m
        self.rerank = cfg.rerank
        self.n_devices = max(1, cfg.n_devices)

    # ---- lifecycle ----
    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().vdb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- VectorIndex / NativeIndex surface ----
    def addVector(self, id_or_vector, values=None):
        """VectorIndex.addVector(Vector) or NativeIndex.addVector(int, float[]). Returns False where Java would
        (capacity reached: HnswPqIndex.java:162-164); dimension mismatch raises ValueError (NativeIndex.java:41-44)."""
        if values is None:
            vid, values = id_or_vector.id, id_or_vector.values
        else:
            vid = id_or_vector
        v = _f32(values)
        if v.ndim != 1 or v.size != self.dimension:
            raise ValueError("Vector dimension mismatch: expected %d, got %d" % (self.dimension, v.size))
        if self.index_type in (INDEX_HNSW, INDEX_HNSW_PQ) and self.contains(vid):
            return False                                                     # HnswPqIndex.java:157-159
        try:
            self.addBatch(np.array([vid], dtype=np.int32), v.reshape(1, -1))
        except VdbError as e:
            if e.code == 3:
                return False
            raise
        return True

    def addBatch(self, ids, vectors):
        v = _f32(vectors)
        if v.ndim != 2 or v.shape[1] != self.dimension:
            raise ValueError("Vector dimension mismatch: expected %d" % self.dimension)
        idp = None
        if ids is not None:
            ids = np.ascontiguousarray(ids, dtype=np.int32)
            if ids.shape[0] != v.shape[0]:
                raise ValueError("ids and vectors must have same length")     # NativeIndex.java:154-156
            idp = ids.ctypes.data
        check(_lib.lib().vdb_add_batch(self._h, idp, v.ctypes.data, v.shape[0]))

    def removeVector(self, id_):
        """VectorIndex.removeVector: graph indexes follow HnswPqIndex.java:542-596; FLAT / PQ answer False like
        NativeIndex.removeVector (NativeIndex.java:47-51)."""
        removed = C.c_int32(0)
        check(_lib.lib().vdb_remove(self._h, int(id_), C.byref(removed)))
        return bool(removed.value)

    def contains(self, id_):
        return bool(_lib.lib().vdb_contains(self._h, int(id_)))

    def train(self, samples):
        s = _f32(samples)
        if s.ndim != 2 or s.shape[1] != self.dimension:
            raise ValueError("Invalid training samples")
        check(_lib.lib().vdb_train(self._h, s.ctypes.data, s.shape[0]))

    def buildIndex(self):
        check(_lib.lib().vdb_build(self._h))
        return True

    def size(self):
        return int(_lib.lib().vdb_size(self._h))

    def rows(self):
        """stored rows including removed ones (row count of codes() / graph() / set_graph buffers)"""
        return int(_lib.lib().vdb_rows(self._h))

    @property
    def trained(self):
        return bool(_lib.lib().vdb_is_trained(self._h))

    def searchBatch(self, queries, k):
        """NativeIndex.searchBatch (NativeIndex.java:187-230): returns (ids[nq][k], distances[nq][k]); empty slots are
        id=-1, distance=+inf."""
        q = _f32(queries)
        if q.ndim != 2 or q.shape[1] != self.dimension:
            raise ValueError("Query dimension mismatch: expected %d" % self.dimension)
        if k <= 0:
            raise ValueError("k must be positive")                            # NativeIndex.java:110-112
        ids = np.empty((q.shape[0], k), dtype=np.int32)
        d = np.empty((q.shape[0], k), dtype=np.float32)
        check(_lib.lib().vdb_search_batch(self._h, q.ctypes.data, q.shape[0], k, ids.ctypes.data, d.ctypes.data))
        return ids, d

    def searchNearest(self, query, k):
        """VectorIndex.searchNearest: list of SearchResult, ascending distance, unfilled slots dropped."""
        q = _f32(getattr(query, "values", query)).reshape(1, -1)
        ids, d = self.searchBatch(q, k)
        return [SearchResult(int(i), float(x)) for i, x in zip(ids[0], d[0]) if i >= 0]

    search = searchNearest

    def search_device(self, d_queries, k, stream=None):
        """Device-resident search: torch CUDA tensors in and out (no host copies)."""
        import torch
        assert d_queries.is_cuda and d_queries.dtype == torch.float32 and d_queries.is_contiguous()
        nq = d_queries.shape[0]
        ids = torch.empty((nq, k), dtype=torch.int32, device=d_queries.device)
        d = torch.empty((nq, k), dtype=torch.float32, device=d_queries.device)
        st = stream if stream is not None else torch.cuda.current_stream(d_queries.device).cuda_stream
        check(_lib.lib().vdb_search_batch_device(self._h, d_queries.data_ptr(), nq, k, ids.data_ptr(), d.data_ptr(), st))
        return ids, d

    def search_device_into(self, d_queries, k, ids, d, stream):
        check(_lib.lib().vdb_search_batch_device(self._h, d_queries.data_ptr(), d_queries.shape[0], k, ids.data_ptr(),
                                                 d.data_ptr(), stream))

    # ---- state import / export ----
    def stats(self):
        s = VdbStats()
        check(_lib.lib().vdb_get_stats(self._h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in VdbStats._fields_}

    def set_codebooks(self, cb):
        cb = _f32(cb)
        check(_lib.lib().vdb_set_codebooks(self._h, cb.ctypes.data))

    def codebooks(self):
        cb = np.empty((self.pq_m, 256, self.dimension // self.pq_m), dtype=np.float32)
        check(_lib.lib().vdb_get_codebooks(self._h, cb.ctypes.data))
        return cb

    def codes(self):
        c = np.empty((self.rows(), self.pq_m), dtype=np.uint8)
        check(_lib.lib().vdb_get_codes(self._h, c.ctypes.data))
        return c

    def set_codes(self, codes):
        c = np.ascontiguousarray(codes, dtype=np.uint8)
        assert c.shape == (self.rows(), self.pq_m)
        check(_lib.lib().vdb_set_codes(self._h, c.ctypes.data))

    def encode_all(self):
        check(_lib.lib().vdb_encode_all(self._h))

    @property
    def max_level(self):
        return int(_lib.lib().vdb_max_level(self._h))

    def set_graph(self, levels, adj_levels, entry):
        levels = np.ascontiguousarray(levels, dtype=np.int32)
        adj = np.ascontiguousarray(np.stack(adj_levels, axis=0), dtype=np.int32)
        assert adj.shape == (self.max_level, self.rows(), self.hnsw_m), adj.shape
        check(_lib.lib().vdb_set_graph(self._h, levels.ctypes.data, adj.ctypes.data, self.max_level, int(entry)))

    def graph(self):
        n = self.rows()
        levels = np.empty(n, dtype=np.int32)
        adj = np.empty((self.max_level, n, self.hnsw_m), dtype=np.int32)
        entry = C.c_int32(-1)
        check(_lib.lib().vdb_get_graph(self._h, levels.ctypes.data, adj.ctypes.data, C.byref(entry)))
        return levels, adj, entry.value

    def save(self, path):
        check(_lib.lib().vdb_save(self._h, str(path).encode()))

    def load(self, path):
        check(_lib.lib().vdb_load(self._h, str(path).encode()))

    def fill_synthetic(self, seed, n):
        check(_lib.lib().vdb_fill_synthetic(self._h, seed, n))


class NativeHnswIndex(GpuIndex):
    """jni/NativeHnswIndex.java: (dimension, maxElements, M, efConstruction, ef) — reproduces Java HnswIndex (cosine)."""

    def __init__(self, dimension, maxElements, M=32, efConstruction=400, ef=400, **kw):
        super().__init__(INDEX_HNSW, dimension, maxElements, hnsw_m=M, ef_construction=efConstruction, ef_search=ef, **kw)


class NativeHnswPqIndex(GpuIndex):
    """The class the reference lacks (no JNI entry for HNSW+PQ): reproduces Java HnswPqIndex
    (index/HnswPqIndex.java:67-108): (dimension, maxElements, M, efConstruction, ef, pqM, pqBits, pqIters)."""

    def __init__(self, dimension, maxElements, M=32, efConstruction=64, ef=64, pqM=0, pqBits=8, pqIters=25, **kw):
        super().__init__(INDEX_HNSW_PQ, dimension, maxElements, hnsw_m=M, ef_construction=efConstruction, ef_search=ef,
                         pq_m=pqM, pq_bits=pqBits, pq_iters=pqIters, **kw)


class NativePqIndex(GpuIndex):
    """jni/NativePqIndex.java:6-25: (dimension, maxElements, M=8, nBits=8); train(samples) before add."""

    def __init__(self, dimension, maxElements, M=8, nBits=8, **kw):
        super().__init__(INDEX_PQ, dimension, maxElements, pq_m=M, pq_bits=nBits, **kw)


class NativeFlatIndex(GpuIndex):
    """Brute force (the reference's ground truth loop, RecallOptimizationTest.java:151-164)."""

    def __init__(self, dimension, maxElements, metric=METRIC_SQL2, **kw):
        super().__init__(INDEX_FLAT, dimension, maxElements, metric=metric, **kw)
